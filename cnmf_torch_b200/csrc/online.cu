// Online-mode building blocks (replicate-batched): the inner multiplicative-update loops that run on
// cached products, and the small glue kernels around them.
//
// nmf-torch's online mode, as the reference drives it (mode='online', cnmf.py:757-771), visits the
// cells in chunks of online_chunk_size rows.  Per chunk: P = x W^T is formed once (tensor-core GEMM,
// dense_tc.cu, "store" epilogue), h is iterated against it until its relative change falls below
// h_tol -- the loop and its stopping rule are the reference's own fit_H_online (cnmf.py:355-378) --
// then the sufficient statistics A += h^T h, B += h^T x are accumulated ("accumulate" epilogue) and W
// is iterated on (A, B) until w_tol.  The two inner loops are the same computation:
//     s <- s * max(p - l1, 0) / (s G + l2 s)           (update_row, common.cuh)
// with (p, G) = (x W^T, W W^T) for h and (B, A) for W, so one kernel serves both.
#include <stdlib.h>

#include "common.cuh"

namespace cnmf {

constexpr int ON_THREADS = 128;
constexpr int ON_REPS = 4;

// One inner iteration for every column m < M of a stacked factor [R*KP x M] and every replicate that
// is still running.  Per-CTA fp64 partials of sum (new - old)^2 and sum old^2.
template <int KP>
__global__ void __launch_bounds__(ON_THREADS) online_update_kernel(
    const float* __restrict__ Pt, int64_t ldp, float* __restrict__ state, int64_t lds, const float* __restrict__ gram,
    const uint8_t* __restrict__ done, const uint8_t* __restrict__ active, int64_t M, int R, float l1, float l2,
    float eps, double* __restrict__ partials) {
  __shared__ float sgram[ON_REPS][KP * KP];
  __shared__ double red[ON_THREADS / 32][2];
  const int r0 = blockIdx.y * ON_REPS;
  for (int i = threadIdx.x; i < ON_REPS * KP * KP; i += ON_THREADS) {
    const int rr = i / (KP * KP);
    sgram[rr][i - rr * KP * KP] = (r0 + rr < R) ? gram[(int64_t)(r0 + rr) * KP * KP + (i - rr * KP * KP)] : 0.f;
  }
  __syncthreads();
  const int64_t m = (int64_t)blockIdx.x * ON_THREADS + threadIdx.x;
  for (int rr = 0; rr < ON_REPS; ++rr) {
    const int r = r0 + rr;
    if (r >= R) break;
    if (done[r] != 0 || (active != nullptr && active[r] == 0)) continue;  // uniform over the CTA
    float d2 = 0.f, n2 = 0.f;
    if (m < M) {
      float p[KP], s[KP], old[KP];
      const int64_t base = (int64_t)r * KP;
#pragma unroll
      for (int k = 0; k < KP; ++k) {
        p[k] = Pt[(base + k) * ldp + m];
        s[k] = state[(base + k) * lds + m];
        old[k] = s[k];
      }
      update_row<KP, CNMF_B200_ALGO_MU>(p, s, sgram[rr], l1, l2, eps);
#pragma unroll
      for (int k = 0; k < KP; ++k) {
        state[(base + k) * lds + m] = s[k];
        const float d = s[k] - old[k];
        d2 = fmaf(d, d, d2);
        n2 = fmaf(old[k], old[k], n2);
      }
    }
    const double dd = warp_sum((double)d2), nn = warp_sum((double)n2);
    if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5][0] = dd; red[threadIdx.x >> 5][1] = nn; }
    __syncthreads();
    if (threadIdx.x == 0) {
      double a = 0.0, b = 0.0;
      for (int w = 0; w < ON_THREADS / 32; ++w) { a += red[w][0]; b += red[w][1]; }
      partials[((int64_t)blockIdx.x * R + r) * 2 + 0] = a;
      partials[((int64_t)blockIdx.x * R + r) * 2 + 1] = b;
    }
    __syncthreads();
  }
}

// rel = ||new - old|| / (||old|| + eps) < tol  =>  the replicate's loop ends (after this update was
// applied, as at cnmf.py:374-378).  One CTA; n_pending[0] = replicates still running afterwards.
__global__ void __launch_bounds__(1024) online_flag_kernel(const double* __restrict__ partials, int nblk, int R,
                                                           float tol, float eps, uint8_t* __restrict__ done,
                                                           const uint8_t* __restrict__ active,
                                                           int* __restrict__ iters, int* __restrict__ n_pending) {
  int pending = 0;
  for (int r = threadIdx.x; r < R; r += 1024) {
    if (done[r] != 0 || (active != nullptr && active[r] == 0)) continue;
    double d2 = 0.0, n2 = 0.0;
    for (int b = 0; b < nblk; ++b) {
      d2 += partials[((int64_t)b * R + r) * 2 + 0];
      n2 += partials[((int64_t)b * R + r) * 2 + 1];
    }
    const float rel = (float)sqrt(d2) / ((float)sqrt(n2) + eps);
    iters[r] += 1;
    if (rel < tol) done[r] = 1;
    else pending = 1;
  }
  pending = __syncthreads_count(pending);
  if (threadIdx.x == 0) n_pending[0] = pending;
}

// ---------------------------------------------------------------------------------------------------------------
// The whole inner loop of a chunk in ONE launch.  The two-kernels-per-iteration form above returns to the host
// every few iterations to learn whether any replicate is still running (a 4-byte copy and a stream synchronise:
// half of an online pass was launch and sync latency).  Here a replicate is owned by `nslab` CTAs (column slabs of
// its factor) that stay resident for the entire loop: every iteration each CTA updates its columns, leaves its
// partial sums of |new - old|^2 and |old|^2 in a slot, and the CTAs of the replicate meet at a counter in global
// memory (release / acquire); each then adds the slots in slab order -- the same number in every CTA -- and the
// replicate stops, on the device, exactly where `online_flag_kernel` would have stopped it.  Launched
// cooperatively (all CTAs co-resident: the spin-wait needs it); replicates beyond the number of resident slots are
// taken in turn by the same CTAs.
__device__ __forceinline__ float onl_tf32(float v) {   // round to nearest tf32 (cvt.rna), as dense_tc.cu's operand copies
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

constexpr int ONP_THREADS = 256;
constexpr int ONP_MAX_SLABS = 8;

// REGS > 0: the CTA's columns fit REGS per thread and their state lives in registers for the whole loop (read once,
// written once): an iteration then streams only P from L2 -- a third of the bytes of the streaming form (REGS = 0).
template <int KP, int REGS>
__global__ void __launch_bounds__(ONP_THREADS, REGS > 0 ? 2 : 1) online_inner_persistent_kernel(
    const float* __restrict__ Pt, int64_t ldp, float* __restrict__ state, int64_t lds, const float* __restrict__ gram,
    const uint8_t* __restrict__ active, int64_t M, int R, float l1, float l2, float eps, float tol, int max_iter,
    int nslab, int* __restrict__ iters, unsigned int* __restrict__ arrive, double* __restrict__ slots,
    unsigned long long* __restrict__ total_iters, double* __restrict__ cross, float* __restrict__ state_op,
    float* __restrict__ state_lo, float* __restrict__ state_copy, int64_t ldo) {
  __shared__ float sgram[KP * KP];
  __shared__ double red[ONP_THREADS / 32][2];
  __shared__ int s_done;
  const int slab = blockIdx.x % nslab, rslot = blockIdx.x / nslab, rstride = gridDim.x / nslab;
  const int64_t cols_per = ((M + nslab - 1) / nslab + 3) / 4 * 4;
  const int64_t c0 = (int64_t)slab * cols_per, c1 = (c0 + cols_per < M) ? c0 + cols_per : M;
  for (int r = rslot; r < R; r += rstride) {
    if (active != nullptr && active[r] == 0) continue;   // uniform over the replicate's CTAs
    __syncthreads();
    for (int i = threadIdx.x; i < KP * KP; i += ONP_THREADS) sgram[i] = gram[(int64_t)r * KP * KP + i];
    __syncthreads();
    const int64_t base = (int64_t)r * KP;
    float sreg[REGS > 0 ? REGS : 1][KP];
    if (REGS > 0) {
#pragma unroll
      for (int j = 0; j < REGS; ++j) {
        const int64_t m = c0 + threadIdx.x + (int64_t)j * ONP_THREADS;
#pragma unroll
        for (int k = 0; k < KP; ++k) sreg[j][k] = m < c1 ? state[(base + k) * lds + m] : 0.f;
      }
    }
    int it = 0;
    while (it < max_iter) {
      double dd = 0.0, nn = 0.0;
      if (REGS > 0) {
#pragma unroll
        for (int j = 0; j < REGS; ++j) {
          const int64_t m = c0 + threadIdx.x + (int64_t)j * ONP_THREADS;
          if (m < c1) {
            float p[KP], old[KP];
#pragma unroll
            for (int k = 0; k < KP; ++k) {
              p[k] = Pt[(base + k) * ldp + m];
              old[k] = sreg[j][k];
            }
            update_row<KP, CNMF_B200_ALGO_MU>(p, sreg[j], sgram, l1, l2, eps);
            float d2 = 0.f, n2 = 0.f;
#pragma unroll
            for (int k = 0; k < KP; ++k) {
              const float d = sreg[j][k] - old[k];
              d2 = fmaf(d, d, d2);
              n2 = fmaf(old[k], old[k], n2);
            }
            dd += (double)d2;
            nn += (double)n2;
          }
        }
      } else {
      for (int64_t m = c0 + threadIdx.x; m < c1; m += ONP_THREADS) {
        float p[KP], s[KP], old[KP];
#pragma unroll
        for (int k = 0; k < KP; ++k) {
          p[k] = Pt[(base + k) * ldp + m];
          s[k] = state[(base + k) * lds + m];
          old[k] = s[k];
        }
        update_row<KP, CNMF_B200_ALGO_MU>(p, s, sgram, l1, l2, eps);
        float d2 = 0.f, n2 = 0.f;
#pragma unroll
        for (int k = 0; k < KP; ++k) {
          state[(base + k) * lds + m] = s[k];
          const float d = s[k] - old[k];
          d2 = fmaf(d, d, d2);
          n2 = fmaf(old[k], old[k], n2);
        }
        dd += (double)d2;
        nn += (double)n2;
      }
      }
      dd = warp_sum(dd);
      nn = warp_sum(nn);
      if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5][0] = dd; red[threadIdx.x >> 5][1] = nn; }
      __syncthreads();
      if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < ONP_THREADS / 32; ++w) { a += red[w][0]; b += red[w][1]; }
        if (nslab > 1) {
          double* mine = slots + (((int64_t)r * ONP_MAX_SLABS + slab) * 2 + (it & 1)) * 2;
          __stcg(mine, a);
          __stcg(mine + 1, b);
          __threadfence();
          atomicAdd(arrive + r, 1u);
          const unsigned int target = (unsigned int)nslab * (unsigned int)(it + 1);
          unsigned int seen = 0, spins = 0;
          do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(arrive + r) : "memory");
            if (seen < target && ++spins > (1u << 27)) __trap();   // a scheduling bug aborts instead of hanging
          } while (seen < target);
          a = 0.0; b = 0.0;
          for (int sl = 0; sl < nslab; ++sl) {   // slab order: every CTA of the replicate forms the same sums
            const double* q = slots + (((int64_t)r * ONP_MAX_SLABS + sl) * 2 + (it & 1)) * 2;
            a += __ldcg(q);
            b += __ldcg(q + 1);
          }
        }
        const float rel = (float)sqrt(a) / ((float)sqrt(b) + eps);   // as online_flag_kernel / cnmf.py:374-378
        s_done = rel < tol ? 1 : 0;
      }
      __syncthreads();
      ++it;
      if (s_done) break;
    }
    // write-back; on the way the tf32 operand copy of the final state (and, for W, the tf32 residual state - copy that
    // the refresh's second GEMM pass reads): two elementwise launches per chunk less
    if (REGS > 0) {
#pragma unroll
      for (int j = 0; j < REGS; ++j) {
        const int64_t m = c0 + threadIdx.x + (int64_t)j * ONP_THREADS;
        if (m < c1) {
#pragma unroll
          for (int k = 0; k < KP; ++k) {
            const float v = sreg[j][k];
            state[(base + k) * lds + m] = v;
            if (state_copy != nullptr) state_copy[(base + k) * ldo + m] = v;
            if (state_op != nullptr) {
              const float o = onl_tf32(v);
              state_op[(base + k) * ldo + m] = o;
              if (state_lo != nullptr) state_lo[(base + k) * ldo + m] = onl_tf32(v - o);
            }
          }
        }
      }
    } else if (state_op != nullptr || state_copy != nullptr) {
      for (int64_t m = c0 + threadIdx.x; m < c1; m += ONP_THREADS) {
#pragma unroll
        for (int k = 0; k < KP; ++k) {
          const float v = state[(base + k) * lds + m];
          if (state_copy != nullptr) state_copy[(base + k) * ldo + m] = v;
          if (state_op != nullptr) {
            const float o = onl_tf32(v);
            state_op[(base + k) * ldo + m] = o;
            if (state_lo != nullptr) state_lo[(base + k) * ldo + m] = onl_tf32(v - o);
          }
        }
      }
    }
    if (cross != nullptr) {
      // cross[r] += <s, p> over all columns (the cross term of the Frobenius objective, <H, X W^T>): per-CTA fp64
      // partials, met at the same counter, added in slab order by the replicate's first CTA
      double cc = 0.0;
      for (int64_t m = c0 + threadIdx.x; m < c1; m += ONP_THREADS) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < KP; ++k) acc = fmaf(state[(base + k) * lds + m], Pt[(base + k) * ldp + m], acc);
        cc += (double)acc;
      }
      cc = warp_sum(cc);
      __syncthreads();
      if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][0] = cc;
      __syncthreads();
      if (threadIdx.x == 0) {
        double a = 0.0;
        for (int w = 0; w < ONP_THREADS / 32; ++w) a += red[w][0];
        if (nslab > 1) {
          double* mine = slots + (((int64_t)r * ONP_MAX_SLABS + slab) * 2 + (it & 1)) * 2;
          __stcg(mine, a);
          __threadfence();
          atomicAdd(arrive + r, 1u);
          if (slab == 0) {
            const unsigned int target = (unsigned int)nslab * (unsigned int)(it + 1);
            unsigned int seen = 0, spins = 0;
            do {
              asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(arrive + r) : "memory");
              if (seen < target && ++spins > (1u << 27)) __trap();
            } while (seen < target);
            a = 0.0;
            for (int sl = 0; sl < nslab; ++sl)
              a += __ldcg(slots + (((int64_t)r * ONP_MAX_SLABS + sl) * 2 + (it & 1)) * 2);
          }
        }
        if (slab == 0) cross[r] += a;
      }
    }
    if (slab == 0 && threadIdx.x == 0) {
      iters[r] += it;
      if (total_iters != nullptr) atomicMax(total_iters, (unsigned long long)it);
    }
  }
}

__global__ void axpy_kernel(float* __restrict__ y, const float* __restrict__ x, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] += x[i];
}

int launch_tc_half_step(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf, float* state,
                        float* state_op, int64_t lds, const float* gram, const uint8_t* active, int R, int Kp, int algo,
                        float l1, float l2, float eps, double* partials, cudaStream_t st);
int launch_round_tf32_2d(float* dst, const float* src, int64_t rows, int64_t cols, int64_t ld, cudaStream_t st);
int launch_gram(const float* F, int64_t ld, int R, int Kp, int64_t L, float* gram, void* ws, cudaStream_t st);
int launch_residual_tf32(float* dst, const float* master, const float* rounded, int64_t rows, int64_t cols, int64_t ld,
                         cudaStream_t st);

}  // namespace cnmf

using namespace cnmf;

extern "C" {

int cnmf_b200_dense_gemm_tf32(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                              float* out, int64_t ldo, const uint8_t* active, int32_t R, int32_t Kp, int32_t accumulate,
                              void* stream) {
  CNMF_NVTX("cnmf_b200_dense_gemm_tf32");
  CNMF_CHECK_ARG(A && F && out, "dense_gemm_tf32: null pointer");
  CNMF_CHECK_ARG(M > 0 && Kd > 0 && R > 0, "dense_gemm_tf32: empty shape");
  CNMF_CHECK_ARG(lda % 4 == 0 && ldf % 4 == 0 && aligned16(A) && aligned16(F),
                 "dense_gemm_tf32: operands need 16-byte aligned rows (ld multiple of 4, base 16-byte aligned)");
  CNMF_CHECK_ARG(lda >= Kd && ldf >= Kd && ldo >= M, "dense_gemm_tf32: leading dimension too small");
  // the gram pointer is not read by these epilogues; pass a valid address
  return launch_tc_half_step(A, lda, M, Kd, F, ldf, out, nullptr, ldo, F, active, R, Kp, accumulate ? 3 : 2, 0.f, 0.f,
                             0.f, nullptr, (cudaStream_t)stream);
}

int cnmf_b200_round_tf32_2d(float* dst, const float* src, int64_t rows, int64_t cols, int64_t ld, void* stream) {
  CNMF_NVTX("cnmf_b200_round_tf32_2d");
  CNMF_CHECK_ARG(dst && src && rows >= 0 && cols >= 0 && ld >= cols, "round_tf32_2d: bad argument");
  return launch_round_tf32_2d(dst, src, rows, cols, ld, (cudaStream_t)stream);
}

int cnmf_b200_axpy(float* y, const float* x, int64_t n, void* stream) {
  CNMF_NVTX("cnmf_b200_axpy");
  CNMF_CHECK_ARG(y && x && n >= 0, "axpy: bad argument");
  if (n == 0) return 0;
  int64_t blocks = ceil_div(n, 256 * 4);
  if (blocks > 148 * 8) blocks = 148 * 8;
  axpy_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(y, x, n);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int64_t cnmf_b200_online_inner_workspace_bytes(int64_t M, int32_t R) {
  return ceil_div(M, ON_THREADS) * (int64_t)R * 2 * 8;
}

int cnmf_b200_online_inner(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram, uint8_t* done,
                           const uint8_t* active, int64_t M, int32_t R, int32_t Kp, float l1, float l2, float eps,
                           float tol, int32_t n_iters, int32_t* iters, int32_t* n_pending, void* workspace,
                           void* stream) {
  CNMF_NVTX("cnmf_b200_online_inner");
  reset_launch_count();
  CNMF_CHECK_ARG(Pt && state && gram && done && iters && n_pending && workspace, "online_inner: null pointer");
  CNMF_CHECK_ARG(M > 0 && R > 0 && ldp >= M && lds >= M && n_iters >= 0, "online_inner: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  double* partials = reinterpret_cast<double*>(workspace);
  const int nblk = (int)ceil_div(M, ON_THREADS);
  dim3 grid((unsigned)nblk, (unsigned)ceil_div(R, ON_REPS));
  for (int it = 0; it < n_iters; ++it) {
    CNMF_DISPATCH_KP(Kp, (online_update_kernel<KP><<<grid, ON_THREADS, 0, st>>>(Pt, ldp, state, lds, gram, done, active,
                                                                                 M, R, l1, l2, eps, partials)));
    CNMF_LAUNCH_CHECK();
    online_flag_kernel<<<1, 1024, 0, st>>>(partials, nblk, R, tol, eps, done, active, iters, n_pending);
    CNMF_LAUNCH_CHECK();
  }
  return 0;
}

int64_t cnmf_b200_online_loop_workspace_bytes(int32_t R) {
  return align_up((int64_t)R * 4, 256) + (int64_t)R * ONP_MAX_SLABS * 2 * 2 * 8;
}

/* The inner loop to convergence in one cooperative launch (see online_inner_persistent_kernel): no host round
 * trips.  iters[r] += iterations run by replicate r; max_iters_seen (optional, device, 64-bit) is raised to the
 * largest count (what the host-driven form reports as "inner iterations").  Replaces the loop of
 * fit_H_online (cnmf.py:355-378) for R replicates at once. */
static int online_loop_impl(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram,
                            const uint8_t* active, int64_t M, int32_t R, int32_t Kp, float l1, float l2, float eps,
                            float tol, int32_t max_iter, int32_t* iters, unsigned long long* max_iters_seen,
                            void* workspace, double* cross, float* state_op, float* state_lo, float* state_copy, int64_t ldo,
                            void* stream);

int cnmf_b200_online_loop(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram,
                          const uint8_t* active, int64_t M, int32_t R, int32_t Kp, float l1, float l2, float eps,
                          float tol, int32_t max_iter, int32_t* iters, unsigned long long* max_iters_seen,
                          void* workspace, void* stream) {
  CNMF_NVTX("cnmf_b200_online_loop");
  return online_loop_impl(Pt, ldp, state, lds, gram, active, M, R, Kp, l1, l2, eps, tol, max_iter, iters, max_iters_seen,
                          workspace, nullptr, nullptr, nullptr, nullptr, 0, stream);
}

static int online_loop_impl(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram,
                            const uint8_t* active, int64_t M, int32_t R, int32_t Kp, float l1, float l2, float eps,
                            float tol, int32_t max_iter, int32_t* iters, unsigned long long* max_iters_seen,
                            void* workspace, double* cross, float* state_op, float* state_lo, float* state_copy, int64_t ldo,
                            void* stream) {
  reset_launch_count();
  CNMF_CHECK_ARG(Pt && state && gram && iters && workspace, "online_loop: null pointer");
  CNMF_CHECK_ARG(M > 0 && R > 0 && ldp >= M && lds >= M && max_iter >= 0, "online_loop: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* arrive = reinterpret_cast<unsigned int*>(workspace);
  double* slots = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) + align_up((int64_t)R * 4, 256));
  CNMF_CUDA(cudaMemsetAsync(arrive, 0, (size_t)R * 4, st));
  int dev = 0, sms = 148;
  CNMF_CUDA(cudaGetDevice(&dev));
  CNMF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // The register form (a CTA's columns <= 4 per thread, state in registers, 2 CTAs per SM) when the replicates' slabs
  // fit the resident CTAs that way; else the streaming form with its own residency.
  auto plan = [&](int resident, int* nslab_out, int* grid_out) {
    int ns = resident / R;
    const int max_slabs_cols = (int)ceil_div(M, ONP_THREADS);
    if (ns > ONP_MAX_SLABS) ns = ONP_MAX_SLABS;
    if (ns > max_slabs_cols) ns = max_slabs_cols;
    if (ns < 1) ns = 1;
    const int rslots = R * ns <= resident ? R : resident / ns;
    *nslab_out = ns;
    *grid_out = rslots * ns;
  };
  const void* kern = nullptr;
  int nslab = 1, grid = 1;
  static const bool no_regs = getenv("CNMF_B200_ONLINE_NOREG") != nullptr;   // dev knob
  if (Kp <= 16 && !no_regs) {
    int per_sm_r = 0;
    CNMF_DISPATCH_KP(Kp, CNMF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
                             &per_sm_r, online_inner_persistent_kernel<KP, 4>, ONP_THREADS, 0)));
    if (per_sm_r > 2) per_sm_r = 2;   // leave room for the kernels of other streams
    if (per_sm_r >= 1) {
      plan(per_sm_r * sms, &nslab, &grid);
      const int64_t cols_per = (ceil_div(M, nslab) + 3) / 4 * 4;
      const int cpt = (int)ceil_div(cols_per, ONP_THREADS);
      if (cpt <= 2) {
        CNMF_DISPATCH_KP(Kp, kern = reinterpret_cast<const void*>(online_inner_persistent_kernel<KP, 2>));
      } else if (cpt <= 4) {
        CNMF_DISPATCH_KP(Kp, kern = reinterpret_cast<const void*>(online_inner_persistent_kernel<KP, 4>));
      }
    }
  }
  if (kern == nullptr) {
    int per_sm = 0;
    CNMF_DISPATCH_KP(Kp, { kern = reinterpret_cast<const void*>(online_inner_persistent_kernel<KP, 0>);
                           CNMF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
                               &per_sm, online_inner_persistent_kernel<KP, 0>, ONP_THREADS, 0)); });
    if (per_sm > 2) per_sm = 2;
    CNMF_CHECK_ARG(per_sm >= 1, "online_loop: kernel does not fit an SM");
    plan(per_sm * sms, &nslab, &grid);
  }
  int64_t M_ = M;
  int R_ = R, max_iter_ = max_iter;
  void* args[] = {(void*)&Pt, (void*)&ldp, (void*)&state, (void*)&lds, (void*)&gram, (void*)&active, (void*)&M_, (void*)&R_,
                  (void*)&l1, (void*)&l2, (void*)&eps, (void*)&tol, (void*)&max_iter_, (void*)&nslab, (void*)&iters,
                  (void*)&arrive, (void*)&slots, (void*)&max_iters_seen, (void*)&cross, (void*)&state_op, (void*)&state_lo,
                  (void*)&state_copy, (void*)&ldo};
  CNMF_CUDA(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(ONP_THREADS), args, 0, st));
  count_launch();
  return 0;
}

/* One chunk of an online pass as ONE call (the kernels are the ones above and the tensor-core GEMM; what this saves
 * is the host: a dozen library calls per chunk from Python cost more than the kernels they enqueue).  The Gram and
 * the statistics GEMM read COMPACT copies of the chunk's usages (Hs, Hopc: [R*Kp x ldp]) that the inner loop writes
 * beside the resident rows: in the resident array a row of the chunk is ldh floats from the next -- 4 MB at a
 * million cells, one page per row for every TMA box.
 *   phase CNMF_B200_ONLINE_USAGE   : P = x W^T, inner loop on the chunk's usages, then the chunk's statistics
 *                                    Ac = h^T h and Bc = h^T x;
 *   phase CNMF_B200_ONLINE_REFRESH : P = x W^T to fp32 accuracy (a second pass on W - tf32(W)), inner loop (max_iter
 *                                    may be 0: evaluation only), cross += <h, P>, gramH_acc += h^T h: the refresh of H
 *                                    at the end of a pass and, on the way, the data-dependent terms of the objective;
 *   phase CNMF_B200_ONLINE_SPECTRA : A += Ac, B += Bc (after the caller's all-reduce when cells are sharded), inner
 *                                    loop on W against (A, B), then W's tf32 copy, its residual and Gram. */
int cnmf_b200_online_chunk(const cnmf_b200_online_chunk_args* a, int32_t phase, void* stream) {
  CNMF_NVTX("cnmf_b200_online_chunk");
  reset_launch_count();
  CNMF_CHECK_ARG(a != nullptr, "online_chunk: null arguments");
  CNMF_CHECK_ARG(a->R > 0 && a->Kp > 0 && a->n_genes > 0, "online_chunk: empty shape");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rk = (int64_t)a->R * a->Kp;
  int64_t n = 0;
  if (phase == CNMF_B200_ONLINE_USAGE || phase == CNMF_B200_ONLINE_REFRESH) {
    CNMF_CHECK_ARG(a->c > 0 && a->x && a->Wop && a->Ht && a->Hs && a->Pt && a->gramW && a->iters_h && a->ws_loop &&
                       a->ws_gram && a->Ac, "online_chunk: null pointer (usage phase)");
    CNMF_CHECK_ARG(a->ldp >= a->c && a->ldp % 4 == 0, "online_chunk: ldp too small or not a multiple of 4");
    CNMF_TRY(cnmf_b200_dense_gemm_tf32(a->x, a->ldx, a->c, a->n_genes, a->Wop, a->ldw, a->Pt, a->ldp, a->active, a->R,
                                       a->Kp, 0, stream));
    n += 1;
    const bool refresh = phase == CNMF_B200_ONLINE_REFRESH;
    if (refresh && a->Wlo != nullptr) {
      CNMF_TRY(cnmf_b200_dense_gemm_tf32(a->x, a->ldx, a->c, a->n_genes, a->Wlo, a->ldw, a->Pt, a->ldp, a->active, a->R,
                                         a->Kp, 1, stream));
      n += 1;
    }
    // the loop reads the chunk's usages from the resident array once, keeps them in registers, and writes back the
    // resident rows plus compact copies (fp32 for the Gram, tf32 for the statistics GEMM)
    CNMF_TRY(online_loop_impl(a->Pt, a->ldp, a->Ht, a->ldh, a->gramW, a->active, a->c, a->R, a->Kp, a->l1_H, a->l2_H,
                              a->eps, a->h_tol, a->max_iter, a->iters_h, nullptr, a->ws_loop,
                              refresh ? a->cross : nullptr, refresh ? nullptr : a->Hopc, nullptr, a->Hs, a->ldp, stream));
    CNMF_TRY(launch_gram(a->Hs, a->ldp, a->R, a->Kp, a->c, a->Ac, a->ws_gram, st));
    n += 3;
    if (refresh) {
      if (a->gramH_acc != nullptr) { CNMF_TRY(cnmf_b200_axpy(a->gramH_acc, a->Ac, (int64_t)a->R * a->Kp * a->Kp, stream)); n += 1; }
    } else {
      CNMF_CHECK_ARG(a->xt && a->Hopc && a->Bc, "online_chunk: null pointer (statistics)");
      CNMF_TRY(cnmf_b200_dense_gemm_tf32(a->xt, a->ldxt, a->n_genes, a->c, a->Hopc, a->ldp, a->Bc, a->ldw, a->active,
                                         a->R, a->Kp, 0, stream));
      n += 1;
    }
  } else if (phase == CNMF_B200_ONLINE_SPECTRA) {
    CNMF_CHECK_ARG(a->A && a->Ac && a->Bst && a->Bc && a->Wall && a->Wop && a->gramW && a->iters_w && a->ws_loop &&
                       a->ws_gram, "online_chunk: null pointer (spectra phase)");
    const int64_t na = (int64_t)a->R * a->Kp * a->Kp, nb = rk * a->ldw;
    if (a->Bst == a->A + na && a->Bc == a->Ac + na) {   // [A | B] and [Ac | Bc] each in one buffer: one launch
      CNMF_TRY(cnmf_b200_axpy(a->A, a->Ac, na + nb, stream));
      n -= 1;
    } else {
      CNMF_TRY(cnmf_b200_axpy(a->A, a->Ac, na, stream));
      CNMF_TRY(cnmf_b200_axpy(a->Bst, a->Bc, nb, stream));
    }
    CNMF_TRY(online_loop_impl(a->Bst, a->ldw, a->Wall, a->ldw, a->A, a->active, a->n_genes, a->R, a->Kp, a->l1_W,
                              a->l2_W, a->eps, a->w_tol, a->max_iter, a->iters_w, nullptr, a->ws_loop, nullptr, a->Wop,
                              a->Wlo, nullptr, a->ldw, stream));
    CNMF_TRY(launch_gram(a->Wall, a->ldw, a->R, a->Kp, a->n_genes, a->gramW, a->ws_gram, st));
    n += 5;
  } else {
    set_error("online_chunk: unknown phase %d", phase);
    return CNMF_B200_ERR_INVALID;
  }
  reset_launch_count();
  count_launch(n);
  return 0;
}

}  // extern "C"
