// Shared host/device helpers for libcnmf_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <nvtx3/nvToolsExt.h>   // header-only; a no-op unless a profiler injects itself

#include "../../include/cnmf_b200.h"

namespace cnmf {

// thread-local error text + launch counter (defined in abi.cu)
void set_error(const char* fmt, ...);
void count_launch(int64_t n = 1);
void reset_launch_count();

#define CNMF_CHECK_ARG(cond, ...)                \
  do {                                           \
    if (!(cond)) {                               \
      cnmf::set_error(__VA_ARGS__);              \
      return CNMF_B200_ERR_INVALID;              \
    }                                            \
  } while (0)

#define CNMF_CUDA(call)                                                                    \
  do {                                                                                     \
    cudaError_t _e = (call);                                                               \
    if (_e != cudaSuccess) {                                                               \
      cnmf::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__,    \
                      __LINE__);                                                           \
      return CNMF_B200_ERR_CUDA;                                                           \
    }                                                                                      \
  } while (0)

#define CNMF_LAUNCH_CHECK()                                                                \
  do {                                                                                     \
    cudaError_t _e = cudaGetLastError();                                                   \
    if (_e != cudaSuccess) {                                                               \
      cnmf::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, \
                      __LINE__);                                                           \
      return CNMF_B200_ERR_CUDA;                                                           \
    }                                                                                      \
    cnmf::count_launch();                                                                  \
  } while (0)

// Named range on the calling host thread for the profilers' timelines (Nsight Systems / Compute --nvtx): every
// entry point of the library opens one, the engines one per stage (half-step, Gram, objective, chunk phase ...).
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
#define CNMF_NVTX_CAT2(a, b) a##b
#define CNMF_NVTX_CAT(a, b) CNMF_NVTX_CAT2(a, b)
#define CNMF_NVTX(name) cnmf::NvtxRange CNMF_NVTX_CAT(_nvtx_, __LINE__)(name)

#define CNMF_TRY(call)          \
  do {                          \
    int _rc = (call);           \
    if (_rc != 0) return _rc;   \
  } while (0)

// cudaFuncSetAttribute is per device: remember, per kernel, on which devices the opt-in shared-memory size
// has been set (bit d of the mask = device ordinal d; ordinals >= 64 are simply set every time).
static inline bool needs_func_attr(unsigned long long* mask) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
  if (*mask >> dev & 1ull) return false;
  *mask |= 1ull << dev;
  return true;
}

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t align_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Dispatch a runtime Kp in {4,8,...,32} to a compile-time constant.
#define CNMF_DISPATCH_KP(kp, ...)                          \
  switch (kp) {                                            \
    case 4:  { constexpr int KP = 4;  __VA_ARGS__; } break; \
    case 8:  { constexpr int KP = 8;  __VA_ARGS__; } break; \
    case 12: { constexpr int KP = 12; __VA_ARGS__; } break; \
    case 16: { constexpr int KP = 16; __VA_ARGS__; } break; \
    case 20: { constexpr int KP = 20; __VA_ARGS__; } break; \
    case 24: { constexpr int KP = 24; __VA_ARGS__; } break; \
    case 28: { constexpr int KP = 28; __VA_ARGS__; } break; \
    case 32: { constexpr int KP = 32; __VA_ARGS__; } break; \
    default:                                               \
      cnmf::set_error("Kp=%d unsupported (multiple of 4, <= 32)", (int)(kp)); \
      return CNMF_B200_ERR_INVALID;                        \
  }

#ifdef __CUDACC__
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// The per-row update shared by the SIMT row-update kernel and the tcgen05 epilogue.
// p[k]  : (Y F^T)[row, r*KP + k]     s[k] : state[row, r*KP + k] (in/out)
// gram  : KP x KP Gram of the fixed factor (shared memory; read as broadcast)
// MU   : cnmf.py:358-372   HALS : sklearn _cdnmf_fast.pyx column sweep restated
// ginv : optional (HALS) KP reciprocals of the regularised diagonal, 1/(G_kk + l2), or a negative value where
//        G_kk + l2 < eps (component frozen at 0): the tcgen05 epilogue precomputes them once per tile, which
//        takes the division (~14 instructions per component) out of the per-row instruction stream.
// KE <= KP: only the first KE components exist (a replicate of K = KE stored in a Kp-padded stack); the padded ones are
// zero and stay untouched -- bit-identical to running all KP, whose extra terms are exact zeros.
template <int KP, int ALGO, int KE = KP>
__device__ __forceinline__ void update_row(const float (&p)[KP], float (&s)[KP],
                                           const float* __restrict__ gram, float l1, float l2,
                                           float eps, const float* __restrict__ ginv = nullptr) {
  if (ALGO == CNMF_B200_ALGO_MU) {
    float out[KP];
#pragma unroll
    for (int k = 0; k < KE; ++k) {
      float denom = 0.f;
#pragma unroll
      for (int j = 0; j < KE; ++j) denom = fmaf(s[j], gram[j * KP + k], denom);
      denom = fmaf(l2, s[k], denom);
      // (p >= 0: a sum of products of non-negative numbers, so the clamp is the identity when l1 == 0; no branch)
      float numer = fmaxf(p[k] - l1, 0.f);
      // numer * rcp(denom) (2 ulp) instead of the IEEE division sequence (~14 instructions with its slow-path
      // check, 12 of them per row in an instruction-issue-bound epilogue); the rate multiplies fp32 state whose
      // inputs carry tf32 rounding (5e-4) on the tensor-core path
      const float q = __fdividef(numer, fmaxf(denom, eps));
      float rate = denom < eps ? 0.f : q;   // a select, not a branch: the KP components stay interleaved
      out[k] = s[k] * rate;
    }
#pragma unroll
    for (int k = 0; k < KE; ++k) s[k] = out[k];
  } else {
    // HALS sweep on a running residual r_j = p_j - l1 - (s.G)_j kept up to date after every component:
    // the same arithmetic as recomputing p_k - s.G_:k before step k, but the per-step dependency is one
    // FMA instead of a KP-long dot product (the KP updates of r are independent of each other), and p is
    // dead once r is formed.
    float r[KP];
#pragma unroll
    for (int j = 0; j < KE; ++j) {
      float t = 0.f;
#pragma unroll
      for (int i = 0; i < KE; ++i) t = fmaf(s[i], gram[i * KP + j], t);
      r[j] = p[j] - t - l1;
    }
#pragma unroll
    for (int k = 0; k < KE; ++k) {
      // s + r * (1/g) in every path (the reciprocal correctly rounded): at most ~1 ulp of the quotient away from
      // the s + r / g of the restated coordinate descent, and the same arithmetic on the SIMT and tcgen05 paths
      float inv;
      if (ginv != nullptr) {
        inv = ginv[k];
      } else {
        const float gkk = gram[k * KP + k] + l2;
        inv = (gkk < eps) ? -1.f : __frcp_rn(gkk);
      }
      const float v = (inv < 0.f) ? 0.f : fmaxf(fmaf(r[k], inv, s[k]), 0.f);
      const float delta = v - s[k];
      s[k] = v;
#pragma unroll
      for (int j = 0; j < KE; ++j)
        if (j > k) r[j] = fmaf(-delta, gram[k * KP + j], r[j]);
    }
  }
}
#endif

}  // namespace cnmf
