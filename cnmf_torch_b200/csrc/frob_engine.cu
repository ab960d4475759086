// Replicate-batched dense Frobenius driver: enqueues n_iters x {H half-step, Gram(H), W half-step,
// Gram(W)} and the objective on one stream.  Replaces the batch loop inside run_nmf
// (/root/reference/src/cnmf/cnmf.py:819) for R replicates of one K at a time.
#include "common.cuh"

namespace cnmf {
int launch_gram(const float*, int64_t, int, int, int64_t, float*, void*, cudaStream_t);
int launch_gemm_nt(const float*, int64_t, int64_t, const float*, int64_t, int64_t, int64_t, float*,
                   int64_t, cudaStream_t);
int launch_row_update(const float*, int64_t, float*, int64_t, const float*, const uint8_t*, int64_t, int,
                      int, int, float, float, float, double*, cudaStream_t);
int launch_frob_objective(double, const double*, int, const double*, int, const float*, const float*, int,
                          int, float, float, float, float, double*, cudaStream_t);
// tcgen05 path (dense_tc.cu)
bool tc_supported(int64_t M, int64_t Kd, int R, int Kp);
int launch_tc_half_step(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                        float* state, float* state_op, int64_t lds, const float* gram, const uint8_t* active,
                        int R, int Kp, int algo, float l1, float l2, float eps, double* partials,
                        cudaStream_t st);
int launch_tc_half_step_g(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                          float* state, float* state_op, int64_t lds, const float* gram, int gram_ntile,
                          const uint8_t* active, int R, int Kp, int algo, float l1, float l2, float eps,
                          double* partials, cudaStream_t st, const uint8_t* rep_k = nullptr, const float* F2 = nullptr,
                          float* state_lo = nullptr);
int gram_num_tiles(int64_t L, int Kp);
int launch_gram_partial(const float* F, int64_t ld, int R, int Kp, int64_t L, void* ws, cudaStream_t st);
int launch_gram_final(const void* ws, int64_t L, int R, int Kp, float* gram, cudaStream_t st);
int launch_round_tf32(float* dst, const float* src, int64_t n, cudaStream_t st);
int launch_objective_sums(const double*, int, const double*, int, int, double*, cudaStream_t);
int launch_add_cross(double* part, const double* extra, int64_t n, cudaStream_t st);
int launch_residual_tf32(float* dst, const float* master, const float* rounded, int64_t rows, int64_t cols, int64_t ld,
                         cudaStream_t st);
int launch_objective_combine(double, const double*, const float*, const float*, int, int, float, float, float, float,
                             double*, cudaStream_t);

struct FrobLayout {
  int64_t off_pt, off_gram, off_gram2, off_partH, off_partW, off_zero, off_wop, off_hop, off_hlo, off_partW2, total;
  int nblkH, nblkW;
  int impl;
};

static int resolve_impl(const cnmf_b200_frob_problem* p) {
  if (p->impl == CNMF_B200_IMPL_SIMT) return CNMF_B200_IMPL_SIMT;
  const bool ok = tc_supported(p->n_cells, p->n_genes, p->R, p->Kp) &&
                  tc_supported(p->n_genes, p->n_cells, p->R, p->Kp);
  if (p->impl == CNMF_B200_IMPL_TCGEN05) return ok ? CNMF_B200_IMPL_TCGEN05 : -1;
  return ok ? CNMF_B200_IMPL_TCGEN05 : CNMF_B200_IMPL_SIMT;
}

static FrobLayout frob_layout(const cnmf_b200_frob_problem* p) {
  FrobLayout l{};
  l.impl = resolve_impl(p);
  const int64_t rk = (int64_t)p->R * p->Kp;
  const int64_t ldmax = align_up(p->n_cells > p->n_genes ? p->n_cells : p->n_genes, 4);
  l.nblkH = (int)cnmf_b200_row_update_blocks(p->n_cells);
  l.nblkW = (int)cnmf_b200_row_update_blocks(p->n_genes);
  int64_t off = 0;
  l.off_pt = off;
  if (l.impl == CNMF_B200_IMPL_SIMT) off += align_up(rk * ldmax * 4, 256);
  l.off_gram = off;
  int64_t gw = cnmf_b200_gram_workspace_bytes(p->R, p->Kp, p->n_cells);
  int64_t gg = cnmf_b200_gram_workspace_bytes(p->R, p->Kp, p->n_genes);
  off += align_up(gw > gg ? gw : gg, 256);
  l.off_gram2 = off;  // second partial buffer: both Grams' partials are live at once on the tcgen05 path
  off += align_up(gw > gg ? gw : gg, 256);
  l.off_partH = off;
  off += align_up((int64_t)l.nblkH * p->R * 2 * 8, 256);
  l.off_partW = off;
  off += align_up((int64_t)l.nblkW * p->R * 2 * 8, 256);
  l.off_zero = off;
  off += align_up(p->R, 256);
  // tcgen05 path: tf32-rounded operand copies of the two stacked factors (same leading dims)
  l.off_wop = off;
  if (l.impl == CNMF_B200_IMPL_TCGEN05) off += align_up(rk * p->ldw * 4, 256);
  l.off_hop = off;
  if (l.impl == CNMF_B200_IMPL_TCGEN05) off += align_up(rk * p->ldh * 4, 256);
  // ... and, for the objective only, the residual H - tf32(H) with its own partial sums
  l.off_hlo = off;
  if (l.impl == CNMF_B200_IMPL_TCGEN05) off += align_up(rk * p->ldh * 4, 256);
  l.off_partW2 = off;
  if (l.impl == CNMF_B200_IMPL_TCGEN05) off += align_up((int64_t)l.nblkW * p->R * 2 * 8, 256);
  l.total = off;
  return l;
}

// tcgen05 path: the cross term <W, H^T X> was formed from the tf32 copy of H.  Add <W, (H - tf32 H)^T X>
// (a second frozen pass on the residual) so that the objective is evaluated for the fp32 factors: without
// it the evaluation noise (~2.5e-5 relative) trips run_nmf's "prev <= cur" stop a check or two early.
static int tc_cross_residual(const cnmf_b200_frob_problem* p, const FrobLayout& l, cudaStream_t st) {
  char* ws = reinterpret_cast<char*>(p->workspace);
  float* Hop = reinterpret_cast<float*>(ws + l.off_hop);
  float* Hlo = reinterpret_cast<float*>(ws + l.off_hlo);
  double* partW = reinterpret_cast<double*>(ws + l.off_partW);
  double* partW2 = reinterpret_cast<double*>(ws + l.off_partW2);
  uint8_t* zero = reinterpret_cast<uint8_t*>(ws + l.off_zero);
  const int64_t rk = (int64_t)p->R * p->Kp;
  CNMF_CUDA(cudaMemsetAsync(zero, 0, p->R, st));
  CNMF_TRY(launch_residual_tf32(Hlo, p->Ht, Hop, rk, p->n_cells, p->ldh, st));
  CNMF_TRY(launch_tc_half_step(p->Xt, p->ldxt, p->n_genes, p->n_cells, Hlo, p->ldh, p->Wall, nullptr, p->ldw, p->gramH,
                               zero, p->R, p->Kp, p->algo, p->l1_W, p->l2_W, p->eps, partW2, st));
  return launch_add_cross(partW, partW2, (int64_t)l.nblkW * p->R, st);
}

static int check_problem(const cnmf_b200_frob_problem* p) {
  CNMF_CHECK_ARG(p != nullptr, "frob: null problem");
  CNMF_CHECK_ARG(p->n_cells > 0 && p->n_genes > 0 && p->R > 0 && p->K > 0, "frob: empty shape");
  CNMF_CHECK_ARG(p->Kp >= p->K && p->Kp % 4 == 0 && p->Kp <= CNMF_B200_MAX_KP,
                 "frob: Kp=%d must be a multiple of 4, >= K=%d and <= %d", p->Kp, p->K, CNMF_B200_MAX_KP);
  CNMF_CHECK_ARG(p->algo == CNMF_B200_ALGO_MU || p->algo == CNMF_B200_ALGO_HALS, "frob: unknown algo %d",
                 p->algo);
  return 0;
}

static int check_pointers(const cnmf_b200_frob_problem* p) {
  CNMF_CHECK_ARG(p->X && p->Xt && p->Wall && p->Ht && p->gramW && p->gramH && p->workspace,
                 "frob: null device pointer");
  CNMF_CHECK_ARG(p->ldx >= p->n_genes && p->ldxt >= p->n_cells && p->ldw >= p->n_genes &&
                     p->ldh >= p->n_cells,
                 "frob: leading dimension too small");
  CNMF_CHECK_ARG(p->ldx % 4 == 0 && p->ldxt % 4 == 0 && p->ldw % 4 == 0 && p->ldh % 4 == 0,
                 "frob: leading dimensions must be multiples of 4");
  CNMF_CHECK_ARG(aligned16(p->X) && aligned16(p->Xt) && aligned16(p->Wall) && aligned16(p->Ht) &&
                     aligned16(p->workspace),
                 "frob: pointers must be 16-byte aligned");
  return 0;
}

}  // namespace cnmf

using namespace cnmf;

extern "C" {

int64_t cnmf_b200_frob_workspace_bytes(const cnmf_b200_frob_problem* p) {
  if (check_problem(p) != 0) return CNMF_B200_ERR_INVALID;
  FrobLayout l = frob_layout(p);
  if (l.impl < 0) {
    set_error("frob: tcgen05 path does not support this shape");
    return CNMF_B200_ERR_UNSUPPORTED;
  }
  return l.total;
}

int cnmf_b200_frob_prepare(const cnmf_b200_frob_problem* p, void* stream) {
  CNMF_NVTX("cnmf_b200_frob_prepare");
  CNMF_TRY(check_problem(p));
  CNMF_TRY(check_pointers(p));
  FrobLayout l = frob_layout(p);
  CNMF_CHECK_ARG(l.impl >= 0, "frob: tcgen05 path does not support this shape");
  if (p->workspace_bytes < l.total) {
    set_error("frob: workspace %lld < required %lld", (long long)p->workspace_bytes, (long long)l.total);
    return CNMF_B200_ERR_WORKSPACE;
  }
  char* ws = reinterpret_cast<char*>(p->workspace);
  cudaStream_t st = (cudaStream_t)stream;
  if (l.impl == CNMF_B200_IMPL_TCGEN05) {
    const int64_t rk = (int64_t)p->R * p->Kp;
    CNMF_TRY(launch_round_tf32(reinterpret_cast<float*>(ws + l.off_wop), p->Wall, rk * p->ldw, st));
    CNMF_TRY(launch_round_tf32(reinterpret_cast<float*>(ws + l.off_hop), p->Ht, rk * p->ldh, st));
  }
  return launch_gram(p->Wall, p->ldw, p->R, p->Kp, p->n_genes, p->gramW, ws + l.off_gram, st);
}

int cnmf_b200_frob_iterate(const cnmf_b200_frob_problem* p, int32_t n_iters, void* stream) {
  CNMF_NVTX("cnmf_b200_frob_iterate");
  reset_launch_count();
  CNMF_TRY(check_problem(p));
  CNMF_TRY(check_pointers(p));
  CNMF_CHECK_ARG(n_iters >= 0, "frob: n_iters < 0");
  FrobLayout l = frob_layout(p);
  CNMF_CHECK_ARG(l.impl >= 0, "frob: tcgen05 path does not support this shape");
  if (p->workspace_bytes < l.total) {
    set_error("frob: workspace %lld < required %lld", (long long)p->workspace_bytes, (long long)l.total);
    return CNMF_B200_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = reinterpret_cast<char*>(p->workspace);
  float* Pt = reinterpret_cast<float*>(ws + l.off_pt);
  void* gws = ws + l.off_gram;
  double* partH = reinterpret_cast<double*>(ws + l.off_partH);
  double* partW = reinterpret_cast<double*>(ws + l.off_partW);
  float* Wop = reinterpret_cast<float*>(ws + l.off_wop);
  float* Hop = reinterpret_cast<float*>(ws + l.off_hop);
  const int64_t N = p->n_cells, G = p->n_genes;
  const int64_t rk = (int64_t)p->R * p->Kp;
  const int64_t ldpH = align_up(N, 4), ldpW = align_up(G, 4);

  void* gws2 = ws + l.off_gram2;
  const int ntH = gram_num_tiles(N, p->Kp), ntW = gram_num_tiles(G, p->Kp);
  for (int it = 0; it < n_iters; ++it) {
    // the objective only reads the partial sums of the last iteration (tcgen05 path skips them otherwise)
    const bool last = (it == n_iters - 1) && p->err2 != nullptr;
    double* pH = (last || l.impl == CNMF_B200_IMPL_SIMT) ? partH : nullptr;
    double* pW = (last || l.impl == CNMF_B200_IMPL_SIMT) ? partW : nullptr;
    if (l.impl == CNMF_B200_IMPL_SIMT) {
      // ---- H half-step: Pt[RKp x N] = Wall X^T ; Ht <- update(Ht; Pt, gramW)
      CNMF_TRY(launch_gemm_nt(p->Wall, p->ldw, rk, p->X, p->ldx, N, G, Pt, ldpH, st));
      CNMF_TRY(launch_row_update(Pt, ldpH, p->Ht, p->ldh, p->gramW, p->active, N, p->R, p->Kp, p->algo,
                                 p->l1_H, p->l2_H, p->eps, partH, st));
      CNMF_TRY(launch_gram(p->Ht, p->ldh, p->R, p->Kp, N, p->gramH, gws, st));
      // ---- W half-step: Pt[RKp x G] = Ht Xt^T ; Wall <- update(Wall; Pt, gramH)
      CNMF_TRY(launch_gemm_nt(p->Ht, p->ldh, rk, p->Xt, p->ldxt, G, N, Pt, ldpW, st));
      CNMF_TRY(launch_row_update(Pt, ldpW, p->Wall, p->ldw, p->gramH, p->active, G, p->R, p->Kp, p->algo,
                                 p->l1_W, p->l2_W, p->eps, partW, st));
      CNMF_TRY(launch_gram(p->Wall, p->ldw, p->R, p->Kp, G, p->gramW, gws, st));
    } else {
      // tcgen05 path: 4 launches per iteration.  The Gram kernels leave per-tile partials; the half-step
      // that consumes a Gram sums them in its (otherwise idle) epilogue warps.
      // On the iteration whose objective is read, the H half-step also leaves Hlo = tf32(H - tf32 H) and the W
      // half-step runs its k loop over both operand copies: Q = X^T (tf32 H + Hlo) is then X^T H to ~2^-22, so the
      // cross term <W, Q> of the objective is that of the fp32 factors (a tf32-only Q gives +-2.5e-5 evaluation
      // noise, enough to trip run_nmf's "prev <= cur" stop early) -- without the separate residual kernel and frozen
      // pass over X that round 1 spent on it (6 % of a 10-iteration step).
      float* Hlo = last ? reinterpret_cast<float*>(ws + l.off_hlo) : nullptr;
      if (it == 0)
        CNMF_TRY(launch_tc_half_step_g(p->X, p->ldx, N, G, Wop, p->ldw, p->Ht, Hop, p->ldh, p->gramW, 0, p->active,
                                       p->R, p->Kp, p->algo, p->l1_H, p->l2_H, p->eps, pH, st, p->rep_k, nullptr, Hlo));
      else
        CNMF_TRY(launch_tc_half_step_g(p->X, p->ldx, N, G, Wop, p->ldw, p->Ht, Hop, p->ldh,
                                       reinterpret_cast<const float*>(gws2), ntW, p->active, p->R, p->Kp, p->algo,
                                       p->l1_H, p->l2_H, p->eps, pH, st, p->rep_k, nullptr, Hlo));
      CNMF_TRY(launch_gram_partial(p->Ht, p->ldh, p->R, p->Kp, N, gws, st));
      CNMF_TRY(launch_tc_half_step_g(p->Xt, p->ldxt, G, N, Hop, p->ldh, p->Wall, Wop, p->ldw,
                                     reinterpret_cast<const float*>(gws), ntH, p->active, p->R, p->Kp, p->algo,
                                     p->l1_W, p->l2_W, p->eps, pW, st, p->rep_k, Hlo, nullptr));
      CNMF_TRY(launch_gram_partial(p->Wall, p->ldw, p->R, p->Kp, G, gws2, st));
      if (it == n_iters - 1) {  // leave both Grams in their final form (objective, next call, API contract)
        CNMF_TRY(launch_gram_final(gws, N, p->R, p->Kp, p->gramH, st));
        CNMF_TRY(launch_gram_final(gws2, G, p->R, p->Kp, p->gramW, st));
      }
    }
  }
  if (p->err2 != nullptr && n_iters > 0) {
    CNMF_TRY(launch_frob_objective(p->normX2, partW, l.nblkW, partH, l.nblkH, p->gramW, p->gramH, p->R,
                                   p->Kp, p->l1_H, p->l2_H, p->l1_W, p->l2_W, p->err2, st));
  }
  return 0;
}

/* Objective of the CURRENT state without changing it (used for the initial error that
 * normalises the convergence test, and after the last partial block). */
static int frob_objective_partials(const cnmf_b200_frob_problem* p, const FrobLayout& l, cudaStream_t st) {
  char* ws = reinterpret_cast<char*>(p->workspace);
  float* Pt = reinterpret_cast<float*>(ws + l.off_pt);
  void* gws = ws + l.off_gram;
  double* partH = reinterpret_cast<double*>(ws + l.off_partH);
  double* partW = reinterpret_cast<double*>(ws + l.off_partW);
  uint8_t* zero = reinterpret_cast<uint8_t*>(ws + l.off_zero);
  float* Wop = reinterpret_cast<float*>(ws + l.off_wop);
  float* Hop = reinterpret_cast<float*>(ws + l.off_hop);
  const int64_t N = p->n_cells, G = p->n_genes;
  const int64_t rk = (int64_t)p->R * p->Kp;
  const int64_t ldpH = align_up(N, 4), ldpW = align_up(G, 4);
  CNMF_CUDA(cudaMemsetAsync(zero, 0, p->R, st));
  CNMF_CUDA(cudaMemsetAsync(partH, 0, (size_t)l.nblkH * p->R * 2 * 8, st));
  CNMF_TRY(launch_gram(p->Ht, p->ldh, p->R, p->Kp, N, p->gramH, gws, st));
  CNMF_TRY(launch_gram(p->Wall, p->ldw, p->R, p->Kp, G, p->gramW, gws, st));
  // frozen W "half-step": only the partial sums (cross = <W, H^T X>, sum W) are produced
  if (l.impl == CNMF_B200_IMPL_SIMT) {
    CNMF_TRY(launch_gemm_nt(p->Ht, p->ldh, rk, p->Xt, p->ldxt, G, N, Pt, ldpW, st));
    CNMF_TRY(launch_row_update(Pt, ldpW, p->Wall, p->ldw, p->gramH, zero, G, p->R, p->Kp, p->algo, p->l1_W,
                               p->l2_W, p->eps, partW, st));
  } else {
    CNMF_TRY(launch_tc_half_step(p->Xt, p->ldxt, G, N, Hop, p->ldh, p->Wall, nullptr, p->ldw, p->gramH, zero,
                                 p->R, p->Kp, p->algo, p->l1_W, p->l2_W, p->eps, partW, st));
    CNMF_TRY(tc_cross_residual(p, l, st));
  }
  // sum(H) for the L1 term comes from partH; with l1_H == 0 the zeros written above are enough
  if (p->l1_H != 0.f) {
    if (l.impl == CNMF_B200_IMPL_SIMT) {
      CNMF_TRY(launch_gemm_nt(p->Wall, p->ldw, rk, p->X, p->ldx, N, G, Pt, ldpH, st));
      CNMF_TRY(launch_row_update(Pt, ldpH, p->Ht, p->ldh, p->gramW, zero, N, p->R, p->Kp, p->algo, p->l1_H,
                                 p->l2_H, p->eps, partH, st));
    } else {
      CNMF_TRY(launch_tc_half_step(p->X, p->ldx, N, G, Wop, p->ldw, p->Ht, nullptr, p->ldh, p->gramW, zero,
                                   p->R, p->Kp, p->algo, p->l1_H, p->l2_H, p->eps, partH, st));
    }
  }
  return 0;
}

static int frob_objective_common(const cnmf_b200_frob_problem* p, FrobLayout* l) {
  CNMF_TRY(check_problem(p));
  CNMF_TRY(check_pointers(p));
  *l = frob_layout(p);
  CNMF_CHECK_ARG(l->impl >= 0, "frob: tcgen05 path does not support this shape");
  if (p->workspace_bytes < l->total) {
    set_error("frob: workspace %lld < required %lld", (long long)p->workspace_bytes, (long long)l->total);
    return CNMF_B200_ERR_WORKSPACE;
  }
  return 0;
}

int cnmf_b200_frob_objective(const cnmf_b200_frob_problem* p, void* stream) {
  CNMF_NVTX("cnmf_b200_frob_objective");
  reset_launch_count();
  FrobLayout l;
  CNMF_TRY(frob_objective_common(p, &l));
  CNMF_CHECK_ARG(p->err2 != nullptr, "frob_objective: err2 is NULL");
  cudaStream_t st = (cudaStream_t)stream;
  CNMF_TRY(frob_objective_partials(p, l, st));
  char* ws = reinterpret_cast<char*>(p->workspace);
  return launch_frob_objective(p->normX2, reinterpret_cast<double*>(ws + l.off_partW), l.nblkW,
                               reinterpret_cast<double*>(ws + l.off_partH), l.nblkH, p->gramW, p->gramH, p->R, p->Kp,
                               p->l1_H, p->l2_H, p->l1_W, p->l2_W, p->err2, st);
}

/* Row-sharded form of the objective: the terms that are sums over this rank's cells.  sums[r*3 + 0] =
 * <W_r, H_r^T X> (cross term), sums[r*3 + 1] = sum(W_r), sums[r*3 + 2] = sum(H_r) over the local rows;
 * gramH = local H^T H, gramW = W W^T.  The caller all-reduces sums[:, 0], sums[:, 2] and gramH across
 * ranks (W is replicated) and then calls cnmf_b200_frob_objective_combine. */
int cnmf_b200_frob_objective_parts(const cnmf_b200_frob_problem* p, double* sums, void* stream) {
  CNMF_NVTX("cnmf_b200_frob_objective_parts");
  reset_launch_count();
  FrobLayout l;
  CNMF_TRY(frob_objective_common(p, &l));
  CNMF_CHECK_ARG(sums != nullptr, "frob_objective_parts: sums is NULL");
  cudaStream_t st = (cudaStream_t)stream;
  CNMF_TRY(frob_objective_partials(p, l, st));
  char* ws = reinterpret_cast<char*>(p->workspace);
  return launch_objective_sums(reinterpret_cast<double*>(ws + l.off_partW), l.nblkW,
                               reinterpret_cast<double*>(ws + l.off_partH), l.nblkH, p->R, sums, st);
}

int cnmf_b200_frob_objective_combine(double normX2, const double* sums, const float* gramW, const float* gramH,
                                     int32_t R, int32_t Kp, float l1_H, float l2_H, float l1_W, float l2_W,
                                     double* err2, void* stream) {
  CNMF_NVTX("cnmf_b200_frob_objective_combine");
  CNMF_CHECK_ARG(sums && gramW && gramH && err2 && R > 0 && Kp > 0, "frob_objective_combine: bad argument");
  return launch_objective_combine(normX2, sums, gramW, gramH, R, Kp, l1_H, l2_H, l1_W, l2_W, err2,
                                  (cudaStream_t)stream);
}

}  // extern "C"
