// fma.rn.f32.f16 (SASS FHFMA: fp16 x fp16 + fp32 -> fp32, half selector in the operand; new on sm_100, full FFMA
// rate -- profiles/r02_simt_rate.txt) on packed fp16 pairs, shared by the two sparse-KL kernel families.
#pragma once
#include <stdint.h>

namespace cnmf {

// ---- fp16 x fp16 + fp32 (FHFMA) on packed words ---------------------------------------------------------------
__device__ __forceinline__ float fhfma_lo(uint32_t a, uint32_t b, float c) {
  asm("{\n\t.reg .b16 al, ah, bl, bh;\n\tmov.b32 {al, ah}, %1;\n\tmov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.f16 %0, al, bl, %0;\n\t}" : "+f"(c) : "r"(a), "r"(b));
  return c;
}
__device__ __forceinline__ float fhfma_hi(uint32_t a, uint32_t b, float c) {
  asm("{\n\t.reg .b16 al, ah, bl, bh;\n\tmov.b32 {al, ah}, %1;\n\tmov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.f16 %0, ah, bh, %0;\n\t}" : "+f"(c) : "r"(a), "r"(b));
  return c;
}
// q (a .b16 register holding one fp16) x low / high half of w
__device__ __forceinline__ float fhfma_q_lo(uint16_t q, uint32_t w, float c) {
  asm("{\n\t.reg .b16 wl, wh;\n\tmov.b32 {wl, wh}, %2;\n\tfma.rn.f32.f16 %0, %1, wl, %0;\n\t}" : "+f"(c) : "h"(q), "r"(w));
  return c;
}
__device__ __forceinline__ float fhfma_q_hi(uint16_t q, uint32_t w, float c) {
  asm("{\n\t.reg .b16 wl, wh;\n\tmov.b32 {wl, wh}, %2;\n\tfma.rn.f32.f16 %0, %1, wh, %0;\n\t}" : "+f"(c) : "h"(q), "r"(w));
  return c;
}
__device__ __forceinline__ uint16_t f32_to_f16_bits(float v) {
  uint16_t h;
  asm("cvt.rn.f16.f32 %0, %1;" : "=h"(h) : "f"(v));   // overflow -> inf (caught when the factor is written)
  return h;
}

}  // namespace cnmf
