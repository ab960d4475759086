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
#include "common.cuh"

namespace cnmf {

constexpr int KL_THREADS = 128;  // replicates per CTA (one per thread)
constexpr int KL_ROWS = 4;       // CSR rows per CTA, processed one after the other

template <int KP>
__device__ __forceinline__ void load_vec(const float* __restrict__ p, float (&v)[KP]) {
#pragma unroll
  for (int c = 0; c < KP; c += 4) {
    const float4 t = *reinterpret_cast<const float4*>(p + c);
    v[c] = t.x; v[c + 1] = t.y; v[c + 2] = t.z; v[c + 3] = t.w;
  }
}

template <int KP>
__global__ void __launch_bounds__(KL_THREADS) kl_half_step_kernel(
    const int* __restrict__ row_ptr, const int* __restrict__ col_idx, const float* __restrict__ val, int64_t M,
    const float* __restrict__ F, float* __restrict__ state, const float* __restrict__ fsum,
    const uint8_t* __restrict__ active, int R, float l1, float l2, float eps) {
  const int r = blockIdx.y * KL_THREADS + threadIdx.x;
  if (r >= R) return;
  if (active != nullptr && active[r] == 0) return;
  const int64_t rk = (int64_t)R * KP;
  float den[KP];
  load_vec<KP>(fsum + (int64_t)r * KP, den);
  const int64_t row0 = (int64_t)blockIdx.x * KL_ROWS;
  for (int rr = 0; rr < KL_ROWS; ++rr) {
    const int64_t row = row0 + rr;
    if (row >= M) break;
    float* sp = state + row * rk + (int64_t)r * KP;
    float h[KP], num[KP];
    load_vec<KP>(sp, h);
#pragma unroll
    for (int k = 0; k < KP; ++k) num[k] = 0.f;
    const int beg = row_ptr[row], end = row_ptr[row + 1];
    int p = beg;
    // two non-zeros in flight per thread to overlap the dependent index -> gather latency
    for (; p + 1 < end; p += 2) {
      const int j0 = col_idx[p], j1 = col_idx[p + 1];
      const float x0 = val[p], x1 = val[p + 1];
      float w0[KP], w1[KP];
      load_vec<KP>(F + (int64_t)j0 * rk + (int64_t)r * KP, w0);
      load_vec<KP>(F + (int64_t)j1 * rk + (int64_t)r * KP, w1);
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int k = 0; k < KP; ++k) { s0 = fmaf(h[k], w0[k], s0); s1 = fmaf(h[k], w1[k], s1); }
      const float q0 = x0 > 0.f ? __fdiv_rn(x0, fmaxf(s0, eps)) : 0.f;
      const float q1 = x1 > 0.f ? __fdiv_rn(x1, fmaxf(s1, eps)) : 0.f;
#pragma unroll
      for (int k = 0; k < KP; ++k) num[k] = fmaf(q1, w1[k], fmaf(q0, w0[k], num[k]));
    }
    if (p < end) {
      const int j0 = col_idx[p];
      const float x0 = val[p];
      float w0[KP];
      load_vec<KP>(F + (int64_t)j0 * rk + (int64_t)r * KP, w0);
      float s0 = 0.f;
#pragma unroll
      for (int k = 0; k < KP; ++k) s0 = fmaf(h[k], w0[k], s0);
      const float q0 = x0 > 0.f ? __fdiv_rn(x0, fmaxf(s0, eps)) : 0.f;
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

static int launch_kl_half(const int* rp, const int* ci, const float* v, int64_t M, const float* F, float* state,
                          const float* fsum, const uint8_t* active, int R, int Kp, float l1, float l2, float eps,
                          cudaStream_t st) {
  dim3 grid((unsigned)ceil_div(M, KL_ROWS), (unsigned)ceil_div(R, KL_THREADS));
  CNMF_DISPATCH_KP(Kp, (kl_half_step_kernel<KP><<<grid, KL_THREADS, 0, st>>>(rp, ci, v, M, F, state, fsum, active, R,
                                                                               l1, l2, eps)));
  CNMF_LAUNCH_CHECK();
  return 0;
}

struct KlLayout {
  int64_t off_wsum, off_hsum, off_wsum64, off_hsum64, off_cpart, off_opart, total;
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

int cnmf_b200_kl_half_step(const int32_t* row_ptr, const int32_t* col_idx, const float* val, int64_t M, const float* F,
                           float* state, const float* fsum, const uint8_t* active, int32_t R, int32_t Kp, float l1,
                           float l2, float eps, void* stream) {
  CNMF_CHECK_ARG(row_ptr && col_idx && val && F && state && fsum, "kl_half_step: null pointer");
  CNMF_CHECK_ARG(M > 0 && R > 0, "kl_half_step: empty shape");
  CNMF_CHECK_ARG(aligned16(F) && aligned16(state) && aligned16(fsum), "kl_half_step: pointers must be 16-byte aligned");
  return launch_kl_half(row_ptr, col_idx, val, M, F, state, fsum, active, R, Kp, l1, l2, eps, (cudaStream_t)stream);
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
  for (int it = 0; it < n_iters; ++it) {
    CNMF_TRY(launch_colsum(p->Wt, p->n_genes, C, wsum, nullptr, cpart, st));
    CNMF_TRY(launch_kl_half(p->row_ptr, p->col_idx, p->val, p->n_cells, p->Wt, p->Hr, wsum, p->active, p->R, p->Kp,
                            p->l1_H, p->l2_H, p->eps, st));
    CNMF_TRY(launch_colsum(p->Hr, p->n_cells, C, hsum, nullptr, cpart, st));
    CNMF_TRY(launch_kl_half(p->t_row_ptr, p->t_col_idx, p->t_val, p->n_genes, p->Hr, p->Wt, hsum, p->active, p->R,
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
