// Dev probe (not part of the library): issue rate of tcgen05.mma kind::tf32 / kind::f16, M=128, N swept, operands
// resident in shared memory (no TMA in the loop). One CTA per SM; one thread issues ITERS x 4 MMAs (one 32-float
// k-block: 4 K-slices of 8 tf32 / 2 K-slices of 16 bf16 ... we keep 4 instructions per block for both), commits,
// waits.  Prints SM clocks per k-block and the wall time.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint64_t desc_sw128(uint32_t a) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((a & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1024u >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
template <int KIND>  // 0 tf32, 1 bf16
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

template <int KIND>
__global__ void __launch_bounds__(128, 1) probe(int N, int iters, int nstage, int fill, long long* clocks, int mode) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int stage_bytes = 16384 + 256 * 128;  // A 128 rows + B up to 256 rows, 128 B each
  // fill operands
  uint32_t* w = reinterpret_cast<uint32_t*>(smem);
  for (int i = threadIdx.x; i < nstage * stage_bytes / 4; i += blockDim.x) {
    uint32_t v = 0;
    if (fill == 1) { v = (i * 2654435761u) ^ (blockIdx.x * 40503u); v = (v & 0x007FE000u) | 0x3F800000u; if (KIND == 1) v = (v & 0x007F007Fu) | 0x3F803F80u; }
    w[i] = v;
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"((mode & 1) ? iters + 1 : 1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  uint32_t elected = 0;
  if (threadIdx.x < 32) {
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(elected));
  }
  if ((mode & 8) && threadIdx.x < 32) {
    // CUTLASS pattern: the whole warp runs the loop; one elected lane issues
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
    long long t0 = clock64();
    int stage = 0;
    for (int it = 0; it < iters; ++it) {
      const uint32_t sa = smem_u32(smem + stage * stage_bytes);
      const uint64_t ad = desc_sw128(sa), bd = desc_sw128(sa + 16384);
      const uint32_t d = tmem + ((it & 1) ? 256u : 0u);
      uint32_t el;
      asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(el));
      if (el) {
#pragma unroll
        for (int k = 0; k < 4; ++k) mma<KIND>(d, ad + 2 * k, bd + 2 * k, idesc, (it > 1 || k) ? 1u : 0u);
        if (mode & 1) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      }
      __syncwarp();
      if (++stage == nstage) stage = 0;
    }
    if (threadIdx.x == 0) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
    long long t1 = clock64();
    clocks[blockIdx.x] = t1 - t0;
    }
  } else
  if ((mode & 4) ? (threadIdx.x < 32 && elected) : (threadIdx.x == 0)) {
    // D=f32; tf32: A=B=2 ; bf16: A=B=1 (kind::f16 with bf16 inputs)
    const uint32_t fmt = KIND == 0 ? 2u : 1u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
    long long t0 = clock64();
    int stage = 0;
    for (int it = 0; it < iters; ++it) {
      const uint32_t sa = smem_u32(smem + stage * stage_bytes);
      const uint64_t ad = desc_sw128(sa), bd = desc_sw128(sa + 16384);
      const uint32_t d = tmem + ((it & 1) ? 256u : 0u);
#pragma unroll
      for (int k = 0; k < 4; ++k) mma<KIND>(d, ad + 2 * k, bd + 2 * k, idesc, (it > 1 || k) ? 1u : 0u);
      if (mode & 1) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      if (mode & 2) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (++stage == nstage) stage = 0;
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
    long long t1 = clock64();
    clocks[blockIdx.x] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main(int argc, char** argv) {
  int iters = argc > 1 ? atoi(argv[1]) : 4000;
  long long* d; cudaMalloc(&d, 148 * 8);
  long long h[148];
  for (int kind = 0; kind < 1; ++kind)
    for (int mode : {0, 8, 9})
      for (int nstage = 4; nstage <= 4; nstage += 3)
        for (int N : {48, 240}) {
          const int fill = 1;
          size_t smem = (size_t)nstage * (16384 + 256 * 128) + 1024;
          auto k = kind == 0 ? probe<0> : probe<1>;
          cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
          cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
          k<<<148, 128, smem>>>(N, 200, nstage, fill, d, mode);
          cudaEventRecord(e0);
          k<<<148, 128, smem>>>(N, iters, nstage, fill, d, mode);
          cudaEventRecord(e1);
          cudaError_t rc = cudaDeviceSynchronize();
          if (rc != cudaSuccess) { printf("error %s\n", cudaGetErrorString(rc)); return 1; }
          float ms; cudaEventElapsedTime(&ms, e0, e1);
          cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
          double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
          printf("mode=%d (1: commit per k-block, 2: fence per k-block, 4: elect.sync instead of tid==0) %s fill=%s stages=%d N=%3d: %.1f clk per k-block (4 MMAs), %.1f ns per k-block => %.0f MHz; ideal %d clk\n",
                 mode, kind == 0 ? "tf32 K=8 " : "bf16 K=16", fill ? "rand" : "zero", nstage, N, avg / iters, ms * 1e6 / iters,
                 avg / (ms * 1e3), 2 * N);
        }
  return 0;
}
