// Kullback-Leibler multiplicative updates on CSR X, replicate-batched (SDDMM + SpMM fused):
//
//   H half-step, cell i, replicate r:   s_ij = sum_k H[i,r,k] W[r,k,j]   only at the non-zeros of row i
//                                       numer_k = sum_j (x_ij / max(s_ij, eps)) W[r,k,j]
//                                       H[i,r,k] <- H[i,r,k] * numer_k / (sum_j W[r,k,j] + l2 H[i,r,k])
//   W half-step: the same kernel on CSR(X^T) with the roles of the two factors swapped.
//
// WH is never formed.  Layout: both factors "replicate-interleaved", Hr [N x R*Kp] and Wt [G x R*Kp], so
// that the R*Kp values one non-zero needs from the fixed factor are one contiguous, 16-byte aligned run
// of row j (thread r reads its Kp floats with 128-bit loads; a warp reads 32 consecutive replicates).
// One thread = (row, replicate): no shuffles, no atomics, deterministic.
//
// Replaces run_nmf(beta_loss='kullback-leibler') (external nmf-torch, called at
// /root/reference/src/cnmf/cnmf.py:819); the update is sklearn/decomposition/_nmf.py:562-607 /
// 640-678 restated with the reference's epsilon rule (cnmf.py:370-371).
#include <cuda_fp16.h>

#include "common.cuh"

namespace cnmf {

constexpr int KL_THREADS = 128;  // replicates per CTA (one per thread)
constexpr int KL_ROWS = 4;       // max CSR rows per CTA (one after the other); 1 when there are few, long rows

template <int KP>
__device__ __forceinline__ void load_vec(const float* __restrict__ p, float (&v)[KP]) {
#pragma unroll
  for (int c = 0; c < KP; c += 4) {
    const float4 t = *reinterpret_cast<const float4*>(p + c);
    v[c] = t.x; v[c + 1] = t.y; v[c + 2] = t.z; v[c + 3] = t.w;
  }
}

// The fixed factor is gathered once per non-zero and replicate: that gather (nnz * R * Kp values per
// half-step) is what bounds the kernel, so the gathered copy is kept in fp16 (11-bit significand, the same
// operand precision as the tf32 copies of the dense path; values below 6e-5 become subnormal, an absolute
// error < 3e-8 that is irrelevant inside s = h.w).  Masters stay fp32; every half-step writes the fp16
// copy of the factor it updates next to the fp32 state.
template <int KP>
__device__ __forceinline__ void load_vec_h(const __half* __restrict__ p, float (&v)[KP]) {
#pragma unroll
  for (int c = 0; c < KP; c += 4) {
    const uint2 t = *reinterpret_cast<const uint2*>(p + c);
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&t.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&t.y));
    v[c] = a.x; v[c + 1] = a.y; v[c + 2] = b.x; v[c + 3] = b.y;
  }
}

__global__ void to_half_kernel(const float* __restrict__ src, __half* __restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = __float2half_rn(fminf(src[i], 65504.f));
}

template <int KP>
__global__ void __launch_bounds__(KL_THREADS) kl_half_step_kernel(
    const int* __restrict__ row_ptr, const int* __restrict__ col_idx, const float* __restrict__ val, int64_t M,
    const __half* __restrict__ F, float* __restrict__ state, __half* __restrict__ state_h,
    const float* __restrict__ fsum, const uint8_t* __restrict__ active, int R, float l1, float l2, float eps,
    int rows_per_cta) {
  const int r = blockIdx.y * KL_THREADS + threadIdx.x;
  if (r >= R) return;
  if (active != nullptr && active[r] == 0) return;
  const int64_t rk = (int64_t)R * KP;
  float den[KP];
  load_vec<KP>(fsum + (int64_t)r * KP, den);
  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  for (int rr = 0; rr < rows_per_cta; ++rr) {
    const int64_t row = row0 + rr;
    if (row >= M) break;
    float* sp = state + row * rk + (int64_t)r * KP;
    float h[KP], num[KP];
    load_vec<KP>(sp, h);
#pragma unroll
    for (int k = 0; k < KP; ++k) num[k] = 0.f;
    const int beg = row_ptr[row], end = row_ptr[row + 1];
    int p = beg;
    // several non-zeros in flight per thread: the index -> gather -> dot -> divide chain of one non-zero is
    // long, and with one thread per (row, replicate) there is no other source of parallelism
    constexpr int U = (KP <= 12) ? 4 : 2;
    for (; p + U <= end; p += U) {
      int j[U];
      float x[U], w[U][KP], sa[U], sb[U], q[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { j[u] = col_idx[p + u]; x[u] = val[p + u]; }
#pragma unroll
      for (int u = 0; u < U; ++u) load_vec_h<KP>(F + (int64_t)j[u] * rk + (int64_t)r * KP, w[u]);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        sa[u] = 0.f; sb[u] = 0.f;
#pragma unroll
        for (int k = 0; k < KP; k += 2) { sa[u] = fmaf(h[k], w[u][k], sa[u]); sb[u] = fmaf(h[k + 1], w[u][k + 1], sb[u]); }
        q[u] = __fdividef(x[u], fmaxf(sa[u] + sb[u], eps));  // stored entries are > 0 (checked at ingest); 2-ulp divide
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int k = 0; k < KP; ++k) num[k] = fmaf(q[u], w[u][k], num[k]);
    }
    for (; p < end; ++p) {
      const int j0 = col_idx[p];
      const float x0 = val[p];
      float w0[KP];
      load_vec_h<KP>(F + (int64_t)j0 * rk + (int64_t)r * KP, w0);
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int k = 0; k < KP; k += 2) { s0 = fmaf(h[k], w0[k], s0); s1 = fmaf(h[k + 1], w0[k + 1], s1); }
      const float q0 = __fdividef(x0, fmaxf(s0 + s1, eps));
#pragma unroll
      for (int k = 0; k < KP; ++k) num[k] = fmaf(q0, w0[k], num[k]);
    }
#pragma unroll
    for (int k = 0; k < KP; ++k) {
      const float d = fmaf(l2, h[k], den[k]);
      const float n = l1 > 0.f ? fmaxf(num[k] - l1, 0.f) : num[k];
      const float rate = d < eps ? 0.f : __fdiv_rn(n, d);
      h[k] *= rate;
    }
#pragma unroll
    for (int c = 0; c < KP; c += 4) *reinterpret_cast<float4*>(sp + c) = make_float4(h[c], h[c + 1], h[c + 2], h[c + 3]);
    if (state_h != nullptr) {
      __half* hp = state_h + row * rk + (int64_t)r * KP;
#pragma unroll
      for (int c = 0; c < KP; c += 4) {
        uint2 t;
        *reinterpret_cast<__half2*>(&t.x) = __floats2half2_rn(fminf(h[c], 65504.f), fminf(h[c + 1], 65504.f));
        *reinterpret_cast<__half2*>(&t.y) = __floats2half2_rn(fminf(h[c + 2], 65504.f), fminf(h[c + 3], 65504.f));
        *reinterpret_cast<uint2*>(hp + c) = t;
      }
    }
  }
}

// sum over non-zeros of x * log(max(s, eps)) per replicate; per-CTA fp64 partials [gridDim.x x R]
constexpr int KLO_ROWS = 64;

template <int KP>
__global__ void __launch_bounds__(KL_THREADS) kl_xlogs_kernel(const int* __restrict__ row_ptr,
                                                              const int* __restrict__ col_idx,
                                                              const float* __restrict__ val, int64_t M,
                                                              const float* __restrict__ Hr, const float* __restrict__ Wt,
                                                              int R, float eps, double* __restrict__ partial) {
  const int r = blockIdx.y * KL_THREADS + threadIdx.x;
  if (r >= R) return;
  const int64_t rk = (int64_t)R * KP;
  double acc = 0.0;
  const int64_t row0 = (int64_t)blockIdx.x * KLO_ROWS;
  for (int rr = 0; rr < KLO_ROWS; ++rr) {
    const int64_t row = row0 + rr;
    if (row >= M) break;
    float h[KP];
    load_vec<KP>(Hr + row * rk + (int64_t)r * KP, h);
    const int beg = row_ptr[row], end = row_ptr[row + 1];
    float racc = 0.f;
    for (int p = beg; p < end; ++p) {
      const int j = col_idx[p];
      const float x = val[p];
      float w[KP];
      load_vec<KP>(Wt + (int64_t)j * rk + (int64_t)r * KP, w);
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < KP; ++k) s = fmaf(h[k], w[k], s);
      if (x > 0.f) racc = fmaf(x, logf(fmaxf(s, eps)), racc);
    }
    acc += (double)racc;
  }
  partial[(int64_t)blockIdx.x * R + r] = acc;
}

// column sums of a [L x C] row-major array: per-CTA partials over 256-row slabs, then a fixed-order sum
constexpr int CS_ROWS = 256;

__global__ void __launch_bounds__(128) colsum_partial_kernel(const float* __restrict__ A, int64_t L, int64_t C,
                                                             float* __restrict__ partial) {
  const int64_t c = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (c >= C) return;
  const int64_t l0 = (int64_t)blockIdx.y * CS_ROWS, l1 = min(L, l0 + (int64_t)CS_ROWS);
  float acc = 0.f;
  for (int64_t l = l0; l < l1; ++l) acc += A[l * C + c];
  partial[(int64_t)blockIdx.y * C + c] = acc;
}

__global__ void __launch_bounds__(128) colsum_final_kernel(const float* __restrict__ partial, int nslab, int64_t C,
                                                           float* __restrict__ out, double* __restrict__ out64) {
  const int64_t c = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (c >= C) return;
  double acc = 0.0;
  for (int s = 0; s < nslab; ++s) acc += (double)partial[(int64_t)s * C + c];
  out[c] = (float)acc;
  if (out64 != nullptr) out64[c] = acc;
}

// err2[r] = 2 * ( sum x log x - sum x log s - sum x + sum_k hsum_k wsum_k + regularisation )
__global__ void __launch_bounds__(32) kl_objective_kernel(double xlogx, double sumX, const double* __restrict__ partial,
                                                          int nblk, const double* __restrict__ hsum,
                                                          const double* __restrict__ wsum, int R, int Kp, float l1_H,
                                                          float l1_W, double* __restrict__ err2) {
  const int r = blockIdx.x, lane = threadIdx.x;
  double xls = 0.0;
  for (int b = lane; b < nblk; b += 32) xls += partial[(int64_t)b * R + r];
  double ss = 0.0, sh = 0.0, sw = 0.0;
  for (int k = lane; k < Kp; k += 32) {
    const double a = hsum[(int64_t)r * Kp + k], b = wsum[(int64_t)r * Kp + k];
    ss += a * b;
    sh += a;
    sw += b;
  }
  xls = warp_sum(xls); ss = warp_sum(ss); sh = warp_sum(sh); sw = warp_sum(sw);
  if (lane == 0) err2[r] = 2.0 * (xlogx - xls - sumX + ss + (double)l1_H * sh + (double)l1_W * sw);
}

// sum x and sum x log x over the stored values (constants of the KL objective), fp64, one CTA
__global__ void __launch_bounds__(1024) kl_constants_kernel(const float* __restrict__ val, int64_t nnz,
                                                            double* __restrict__ out) {
  __shared__ double r0[32], r1[32];
  double a = 0.0, b = 0.0;
  for (int64_t i = threadIdx.x; i < nnz; i += 1024) {
    const double x = val[i];
    if (x > 0.0) { a += x; b += x * log(x); }
  }
  a = warp_sum(a); b = warp_sum(b);
  if ((threadIdx.x & 31) == 0) { r0[threadIdx.x >> 5] = a; r1[threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ta = 0.0, tb = 0.0;
    for (int w = 0; w < 32; ++w) { ta += r0[w]; tb += r1[w]; }
    out[0] = ta;
    out[1] = tb;
  }
}

static int launch_colsum(const float* A, int64_t L, int64_t C, float* out, double* out64, float* partial,
                         cudaStream_t st) {
  const int nslab = (int)ceil_div(L, CS_ROWS);
  dim3 grid((unsigned)ceil_div(C, 128), nslab);
  colsum_partial_kernel<<<grid, 128, 0, st>>>(A, L, C, partial);
  CNMF_LAUNCH_CHECK();
  colsum_final_kernel<<<(unsigned)ceil_div(C, 128), 128, 0, st>>>(partial, nslab, C, out, out64);
  CNMF_LAUNCH_CHECK();
  return 0;
}

static int launch_to_half(const float* src, __half* dst, int64_t n, cudaStream_t st) {
  if (n == 0) return 0;
  int64_t blocks = ceil_div(n, 256 * 8);
  if (blocks > 148 * 16) blocks = 148 * 16;
  to_half_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, dst, n);
  CNMF_LAUNCH_CHECK();
  return 0;
}

static int launch_kl_half(const int* rp, const int* ci, const float* v, int64_t M, const __half* F, float* state,
                          __half* state_h, const float* fsum, const uint8_t* active, int R, int Kp, float l1, float l2,
                          float eps, cudaStream_t st) {
  // few long rows (the gene-major half-step): one row per CTA so that there are enough CTAs to fill the SMs
  const int rows = M >= 148 * 5 * 4 * KL_ROWS ? KL_ROWS : 1;
  dim3 grid((unsigned)ceil_div(M, rows), (unsigned)ceil_div(R, KL_THREADS));
  CNMF_DISPATCH_KP(Kp, (kl_half_step_kernel<KP><<<grid, KL_THREADS, 0, st>>>(rp, ci, v, M, F, state, state_h, fsum,
                                                                               active, R, l1, l2, eps, rows)));
  CNMF_LAUNCH_CHECK();
  return 0;
}

struct KlLayout {
  int64_t off_wsum, off_hsum, off_wsum64, off_hsum64, off_cpart, off_opart, off_wh, off_hh, total;
  int nblk_obj;
};

static KlLayout kl_layout(const cnmf_b200_kl_problem* p) {
  KlLayout l{};
  const int64_t C = (int64_t)p->R * p->Kp;
  const int64_t Lmax = p->n_cells > p->n_genes ? p->n_cells : p->n_genes;
  int64_t off = 0;
  l.off_wsum = off; off += align_up(C * 4, 256);
  l.off_hsum = off; off += align_up(C * 4, 256);
  l.off_wsum64 = off; off += align_up(C * 8, 256);
  l.off_hsum64 = off; off += align_up(C * 8, 256);
  l.off_cpart = off; off += align_up(ceil_div(Lmax, CS_ROWS) * C * 4, 256);
  l.nblk_obj = (int)ceil_div(p->n_cells, KLO_ROWS);
  l.off_opart = off; off += align_up((int64_t)l.nblk_obj * p->R * 8, 256);
  l.off_wh = off; off += align_up(p->n_genes * C * 2, 256);   // fp16 gather copies of Wt and Hr
  l.off_hh = off; off += align_up(p->n_cells * C * 2, 256);
  l.total = off;
  return l;
}

static int kl_check(const cnmf_b200_kl_problem* p, bool pointers) {
  CNMF_CHECK_ARG(p != nullptr, "kl: null problem");
  CNMF_CHECK_ARG(p->n_cells > 0 && p->n_genes > 0 && p->R > 0 && p->K > 0 && p->nnz >= 0, "kl: empty shape");
  CNMF_CHECK_ARG(p->nnz < (1ll << 31), "kl: nnz must fit int32 offsets");
  CNMF_CHECK_ARG(p->Kp >= p->K && p->Kp % 4 == 0 && p->Kp <= CNMF_B200_MAX_KP,
                 "kl: Kp=%d must be a multiple of 4, >= K=%d and <= %d", p->Kp, p->K, CNMF_B200_MAX_KP);
  if (pointers) {
    CNMF_CHECK_ARG(p->row_ptr && p->col_idx && p->val && p->t_row_ptr && p->t_col_idx && p->t_val && p->Hr && p->Wt &&
                       p->workspace,
                   "kl: null device pointer");
    CNMF_CHECK_ARG(aligned16(p->Hr) && aligned16(p->Wt) && aligned16(p->workspace), "kl: pointers must be 16-byte aligned");
    KlLayout l = kl_layout(p);
    if (p->workspace_bytes < l.total) {
      set_error("kl: workspace %lld < required %lld", (long long)p->workspace_bytes, (long long)l.total);
      return CNMF_B200_ERR_WORKSPACE;
    }
  }
  return 0;
}

static int kl_objective_impl(const cnmf_b200_kl_problem* p, const KlLayout& l, cudaStream_t st) {
  char* ws = reinterpret_cast<char*>(p->workspace);
  float* wsum = reinterpret_cast<float*>(ws + l.off_wsum);
  float* hsum = reinterpret_cast<float*>(ws + l.off_hsum);
  double* wsum64 = reinterpret_cast<double*>(ws + l.off_wsum64);
  double* hsum64 = reinterpret_cast<double*>(ws + l.off_hsum64);
  float* cpart = reinterpret_cast<float*>(ws + l.off_cpart);
  double* opart = reinterpret_cast<double*>(ws + l.off_opart);
  const int64_t C = (int64_t)p->R * p->Kp;
  CNMF_TRY(launch_colsum(p->Wt, p->n_genes, C, wsum, wsum64, cpart, st));
  CNMF_TRY(launch_colsum(p->Hr, p->n_cells, C, hsum, hsum64, cpart, st));
  dim3 grid((unsigned)l.nblk_obj, (unsigned)ceil_div(p->R, KL_THREADS));
  CNMF_DISPATCH_KP(p->Kp, (kl_xlogs_kernel<KP><<<grid, KL_THREADS, 0, st>>>(p->row_ptr, p->col_idx, p->val, p->n_cells,
                                                                              p->Hr, p->Wt, p->R, p->eps, opart)));
  CNMF_LAUNCH_CHECK();
  kl_objective_kernel<<<p->R, 32, 0, st>>>(p->xlogx, p->sumX, opart, l.nblk_obj, hsum64, wsum64, p->R, p->Kp, p->l1_H,
                                           p->l1_W, p->err2);
  CNMF_LAUNCH_CHECK();
  return 0;
}

}  // namespace cnmf

using namespace cnmf;

extern "C" {

int cnmf_b200_kl_constants(const float* val, int64_t nnz, double* out2, void* stream) {
  CNMF_CHECK_ARG(val && out2 && nnz >= 0, "kl_constants: bad argument");
  kl_constants_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(val, nnz, out2);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int64_t cnmf_b200_kl_workspace_bytes(const cnmf_b200_kl_problem* p) {
  if (kl_check(p, false) != 0) return CNMF_B200_ERR_INVALID;
  return kl_layout(p).total;
}

int cnmf_b200_to_half(const float* src, void* dst, int64_t n, void* stream) {
  CNMF_CHECK_ARG(src && dst && n >= 0, "to_half: bad argument");
  return launch_to_half(src, reinterpret_cast<__half*>(dst), n, (cudaStream_t)stream);
}

int cnmf_b200_kl_half_step(const int32_t* row_ptr, const int32_t* col_idx, const float* val, int64_t M,
                           const void* F_half, float* state, void* state_half, const float* fsum,
                           const uint8_t* active, int32_t R, int32_t Kp, float l1, float l2, float eps, void* stream) {
  CNMF_CHECK_ARG(row_ptr && col_idx && val && F_half && state && fsum, "kl_half_step: null pointer");
  CNMF_CHECK_ARG(M > 0 && R > 0, "kl_half_step: empty shape");
  CNMF_CHECK_ARG((reinterpret_cast<uintptr_t>(F_half) & 7u) == 0 && aligned16(state) && aligned16(fsum),
                 "kl_half_step: F_half must be 8-byte, state and fsum 16-byte aligned");
  return launch_kl_half(row_ptr, col_idx, val, M, reinterpret_cast<const __half*>(F_half), state,
                        reinterpret_cast<__half*>(state_half), fsum, active, R, Kp, l1, l2, eps, (cudaStream_t)stream);
}

int cnmf_b200_kl_iterate(const cnmf_b200_kl_problem* p, int32_t n_iters, void* stream) {
  reset_launch_count();
  CNMF_TRY(kl_check(p, true));
  CNMF_CHECK_ARG(n_iters >= 0, "kl: n_iters < 0");
  const KlLayout l = kl_layout(p);
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = reinterpret_cast<char*>(p->workspace);
  float* wsum = reinterpret_cast<float*>(ws + l.off_wsum);
  float* hsum = reinterpret_cast<float*>(ws + l.off_hsum);
  float* cpart = reinterpret_cast<float*>(ws + l.off_cpart);
  const int64_t C = (int64_t)p->R * p->Kp;
  __half* Wh = reinterpret_cast<__half*>(ws + l.off_wh);
  __half* Hh = reinterpret_cast<__half*>(ws + l.off_hh);
  if (n_iters > 0) CNMF_TRY(launch_to_half(p->Wt, Wh, p->n_genes * C, st));  // Hh is written by the first H half-step
  for (int it = 0; it < n_iters; ++it) {
    CNMF_TRY(launch_colsum(p->Wt, p->n_genes, C, wsum, nullptr, cpart, st));
    CNMF_TRY(launch_kl_half(p->row_ptr, p->col_idx, p->val, p->n_cells, Wh, p->Hr, Hh, wsum, p->active, p->R, p->Kp,
                            p->l1_H, p->l2_H, p->eps, st));
    if (it == 0 && p->active != nullptr)  // frozen replicates skip the write: their copy must exist too
      CNMF_TRY(launch_to_half(p->Hr, Hh, p->n_cells * C, st));
    CNMF_TRY(launch_colsum(p->Hr, p->n_cells, C, hsum, nullptr, cpart, st));
    CNMF_TRY(launch_kl_half(p->t_row_ptr, p->t_col_idx, p->t_val, p->n_genes, Hh, p->Wt, Wh, hsum, p->active, p->R,
                            p->Kp, p->l1_W, p->l2_W, p->eps, st));
  }
  if (p->err2 != nullptr && n_iters > 0) CNMF_TRY(kl_objective_impl(p, l, st));
  return 0;
}

int cnmf_b200_kl_objective(const cnmf_b200_kl_problem* p, void* stream) {
  reset_launch_count();
  CNMF_TRY(kl_check(p, true));
  CNMF_CHECK_ARG(p->err2 != nullptr, "kl_objective: err2 is NULL");
  return kl_objective_impl(p, kl_layout(p), (cudaStream_t)stream);
}

}  // extern "C"
