/*
 * cnmf_b200.h -- C ABI of libcnmf_b200.so: the B200 (sm_100a) kernels behind cNMF's
 * factorize + consensus hot path.
 *
 * The reference (aron0093/cNMF_torch) is pure Python; its "FFI" for this path is the single
 * Python call  run_nmf(X, **kwargs) -> (H, W, err)   at src/cnmf/cnmf.py:819 (external package
 * nmf-torch) and the sklearn/numpy calls at src/cnmf/cnmf.py:1056-1090.  Each entry point
 * below names the reference line(s) whose arithmetic it replaces.  The Python host layer
 * (cnmf_torch_b200/) binds these with ctypes; INTEGRATION.md shows the reference-side stub.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host; the caller owns all
 *     memory; nothing here allocates (workspaces are sized by the *_workspace_bytes queries);
 *   - `stream` is a cudaStream_t passed as void*; all work is asynchronous on that stream;
 *   - return value: 0 on success, negative error code otherwise; cnmf_b200_last_error()
 *     gives a thread-local message.  No C++ exceptions cross the boundary;
 *   - matrices are row-major with an explicit leading dimension (elements); all leading
 *     dimensions must be multiples of 4 and base pointers 16-byte aligned;
 *   - orientation (nmf-torch naming, cnmf.py:758): X (N x G) ~= H (N x K) * W (K x G).
 *     R replicates of one K are stacked: rows r*Kp .. r*Kp+K-1 of a [R*Kp x L] array belong
 *     to replicate r (Kp = K rounded up to a multiple of 4, padding rows are zero):
 *         W  stacked as  Wall [R*Kp x G]      (spectra, gene index contiguous)
 *         H  stacked as  Ht   [R*Kp x N]      (usages TRANSPOSED, cell index contiguous)
 */
#ifndef CNMF_B200_H_
#define CNMF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CNMF_B200_ABI_VERSION 1

#define CNMF_B200_OK 0
#define CNMF_B200_ERR_INVALID (-1)   /* bad argument (shape, alignment, enum)           */
#define CNMF_B200_ERR_CUDA (-2)      /* a CUDA runtime/driver call failed                */
#define CNMF_B200_ERR_WORKSPACE (-3) /* workspace too small                              */
#define CNMF_B200_ERR_UNSUPPORTED (-4)

#define CNMF_B200_ALGO_MU 0
#define CNMF_B200_ALGO_HALS 1

#define CNMF_B200_IMPL_AUTO 0
#define CNMF_B200_IMPL_SIMT 1    /* fp32 CUDA-core GEMM + separate row update (semantic baseline) */
#define CNMF_B200_IMPL_TCGEN05 2 /* TMA + tcgen05 TF32 GEMM, update fused in the TMEM epilogue     */

#define CNMF_B200_MAX_KP 32

int cnmf_b200_abi_version(void);
const char* cnmf_b200_last_error(void);

/* ------------------------------------------------------------------------------------------
 * Dense Frobenius engine (replicate-batched).  Replaces the arithmetic of run_nmf(beta_loss=
 * 'frobenius', algo in {'mu','hals'}) called at cnmf.py:819 for R replicates at once.
 * ---------------------------------------------------------------------------------------- */

/* out[0] = sum of squares of x (rows x cols, ld) accumulated in fp64.  ||X||_F^2 is the
 * constant term of the Frobenius objective (sklearn _nmf.py:113-117 restated by trace identity).
 * workspace: cnmf_b200_sumsq_workspace_bytes() bytes. */
int64_t cnmf_b200_sumsq_workspace_bytes(void);
int cnmf_b200_sumsq(const float* x, int64_t rows, int64_t cols, int64_t ld, double* out,
                    void* workspace, void* stream);

/* gram[r] = F_r F_r^T  (Kp x Kp, row-major) for every replicate block F_r = F[r*Kp : (r+1)*Kp, 0:L].
 * Replaces  W @ W.T  /  H.T @ H  (pattern in-repo at cnmf.py:346). Deterministic two-stage sum.
 * Kp: a multiple of 4 in (0, CNMF_B200_MAX_KP] (anything else is CNMF_B200_ERR_INVALID); for Kp > 12
 * ld must be a multiple of 4 and F 16-byte aligned (the tiled kernel reads F with 16-byte loads). */
int64_t cnmf_b200_gram_workspace_bytes(int32_t R, int32_t Kp, int64_t L);
int cnmf_b200_gram(const float* F, int64_t ld, int32_t R, int32_t Kp, int64_t L, float* gram,
                   void* workspace, void* stream);

/* cnmf_b200_gram for callers that cannot allocate device memory and hold no workspace (all of HBM
 * already owned by the host application's pool, or a loop that cannot afford cudaMalloc's implicit
 * device synchronisation).  No counterpart in the reference; same output as cnmf_b200_gram.
 * The split partials live in a 4 MiB array inside the library (resident on every GPU that loads it).
 * Refuses problems with R * ceil(L/256) * Kp*Kp > cnmf_b200_gram_static_max_floats() (2^20) with
 * CNMF_B200_ERR_WORKSPACE.  Allocates nothing.  The first call on a device resolves the array's
 * address: issue it before any stream capture, and before the host application takes the rest of
 * device memory -- with lazy module loading that first lookup may be what loads the library image
 * (including the array) onto the device, and it may synchronise.  The address is cached per device
 * ordinal for the life of the process: after cudaDeviceReset the library must be reloaded.  Calls
 * share the array: calls that overlap on the GPU race; nothing in it needs clearing between calls.
 * Same Kp / ld / alignment rules as cnmf_b200_gram. */
int64_t cnmf_b200_gram_static_max_floats(void);
int cnmf_b200_gram_static(const float* F, int64_t ld, int32_t R, int32_t Kp, int64_t L, float* gram,
                          void* stream);

/* C[M x N] = A[M x Kd] * B[N x Kd]^T, fp32 FFMA (CUDA cores).  Replaces  x @ W.T / H.T @ x
 * (cnmf.py:355) in the SIMT baseline. */
int cnmf_b200_gemm_nt(const float* A, int64_t lda, int64_t M, const float* B, int64_t ldb,
                      int64_t N, int64_t Kd, float* C, int64_t ldc, void* stream);

/* Row update of one stacked factor given Pt = F * A^T  ([R*Kp x M], ld = ldp):
 *   MU   (cnmf.py:358-372):  s <- s * max(p - l1, 0) / (s.Gram + l2*s), rate = 0 where denom < eps
 *   HALS (sklearn _cdnmf_fast.pyx restated): s_k <- max(0, s_k + (p_k - l1 - s.Gram_:k)/(Gram_kk + l2))
 * for every column m < M and every replicate r with active[r] != 0 (active may be NULL = all).
 * partials ([nblk x R x 2] doubles, nblk = cnmf_b200_row_update_blocks(M)) receives per-block
 * sums of (new_s * p) and (new_s) for the objective. */
int64_t cnmf_b200_row_update_blocks(int64_t M);
int cnmf_b200_row_update(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram,
                         const uint8_t* active, int64_t M, int32_t R, int32_t Kp, int32_t algo,
                         float l1, float l2, float eps, double* partials, void* stream);

/* In-place round-to-nearest fp32 -> tf32 (cvt.rna.tf32.f32) of n contiguous floats.  Applied once to
 * the static operands X and X^T before the tcgen05 path reads them: the tensor core ignores the low
 * 13 mantissa bits, and rounding first makes that error unbiased.  No reference counterpart (the
 * reference computes x @ W.T in fp32, cnmf.py:355). */
int cnmf_b200_round_tf32(float* x, int64_t n, void* stream);

/* Fused half-step on the tensor cores (TMA -> tcgen05.mma kind::tf32 -> TMEM epilogue):
 *   P = A[M x Kd] * F[R*Kp x Kd]^T is formed tile by tile in tensor memory and never stored;
 *   state[R*Kp x M] receives the MU / HALS row update of cnmf_b200_row_update, partials likewise
 *   (partials may be NULL: the objective's partial sums are then skipped);
 *   state_tf32 (same layout, may be NULL) receives the updated rows rounded to tf32 -- the operand
 *   F of the next half-step, so that operand rounding stays unbiased while the master stays fp32.
 * A and F should hold tf32-representable values (cnmf_b200_round_tf32); other bits are ignored.
 * Replaces  x @ W.T (+ update, cnmf.py:355-372)  or  H.T @ x (+ update)  for R replicates. */
int cnmf_b200_dense_half_step(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F,
                              int64_t ldf, float* state, float* state_tf32, int64_t lds,
                              const float* gram, const uint8_t* active, int32_t R, int32_t Kp,
                              int32_t algo, float l1, float l2, float eps, double* partials,
                              void* stream);

/* The same for a stack of several K padded to one Kp: rep_k[r] (device, [R]) is the true K of replicate r
 * (Kp-3 <= rep_k[r] <= Kp, the rows beyond are zero and stay zero); the epilogue works on rep_k[r] components only. */
int cnmf_b200_dense_half_step_k(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                                float* state, float* state_tf32, int64_t lds, const float* gram,
                                const uint8_t* active, const uint8_t* rep_k, int32_t R, int32_t Kp, int32_t algo,
                                float l1, float l2, float eps, double* partials, void* stream);

typedef struct cnmf_b200_frob_problem {
  int64_t n_cells;   /* N */
  int64_t n_genes;   /* G */
  const float* X;    /* [N x G]  */
  int64_t ldx;
  const float* Xt;   /* [G x N]  (X transposed, resident; X is static across replicates) */
  int64_t ldxt;
  int32_t R;         /* replicates in this batch */
  int32_t K;         /* n_components */
  int32_t Kp;        /* K rounded up to a multiple of 4 (<= CNMF_B200_MAX_KP) */
  int32_t algo;      /* CNMF_B200_ALGO_* */
  int32_t impl;      /* CNMF_B200_IMPL_* */
  int32_t reserved;
  float* Wall;       /* [R*Kp x G], updated in place */
  int64_t ldw;
  float* Ht;         /* [R*Kp x N], updated in place */
  int64_t ldh;
  float* gramW;      /* [R x Kp x Kp]  W_r W_r^T, must be current on entry, current on exit */
  float* gramH;      /* [R x Kp x Kp]  H_r^T H_r, written */
  const uint8_t* active; /* [R] 1 = update, 0 = frozen; NULL = all active */
  float l1_H, l2_H, l1_W, l2_W;
  float eps;         /* denominators below eps give a zero rate (cnmf.py:269, 370-371) */
  float pad0;
  double normX2;     /* ||X||_F^2 */
  double* err2;      /* [R] out: 2 * objective = ||X - H W||_F^2 + regularisation (may be NULL) */
  void* workspace;
  int64_t workspace_bytes;
  const uint8_t* rep_k;  /* [R] device: the true K of every replicate when a stack holds several K padded to one Kp
                            (Kp-3 <= rep_k[r] <= Kp; rows rep_k[r]..Kp-1 of the replicate are zero); NULL = all K.
                            The tensor-core epilogue then updates, loads and stores rep_k[r] components only. */
} cnmf_b200_frob_problem;

/* 'random' init of R replicates of one K, drawn on the device into the stacked layout:
 *   Ht[(r*Kp + k) * ldh + i] = |avg * z_H[i, k]|  (i = cell - row_offset, 0 <= i < n_local),
 *   Wall[(r*Kp + k) * ldw + j] = |avg * z_W[k, j]|,
 * where z_H [n_total x K] then z_W [K x n_genes] are what torch's CUDA generator seeded with seeds[r] yields
 * (torch.randn on this GPU: Philox4_32_10 via cuRAND's device API with ATen's grid geometry, restated here), i.e.
 * run_nmf(init='random', random_state=seeds[r]) of nmf-torch under use_gpu=True (called at cnmf.py:819; the seed
 * is the ledger's nmf_seed, cnmf.py:885).  avg = sqrt(mean(X) / K).  seeds: DEVICE array [R].  Rows k >= K and
 * padding columns are not written (the caller zero-fills).  A row-sharded caller passes its block of rows
 * (row_offset, n_local) of the n_total cells and gets its block of the global draw. */
int cnmf_b200_init_random(const int64_t* seeds, int32_t R, int32_t K, int32_t Kp, int64_t n_total, int64_t n_local,
                          int64_t row_offset, int64_t n_genes, float avg, float* Ht, int64_t ldh, float* Wall,
                          int64_t ldw, void* stream);

/* Bytes of workspace cnmf_b200_frob_iterate needs for this problem (only shapes/impl are read). */
int64_t cnmf_b200_frob_workspace_bytes(const cnmf_b200_frob_problem* p);

/* gramW <- Gram(Wall), and (tcgen05 path) the tf32 operand copies of Wall / Ht inside the workspace.
 * Call after the host has written Wall / Ht and before cnmf_b200_frob_objective / _iterate. */
int cnmf_b200_frob_prepare(const cnmf_b200_frob_problem* p, void* stream);

/* n_iters x { H half-step; gramH; W half-step; gramW }, then (if err2 != NULL) the objective of
 * the final state via  ||X||^2 - 2 tr(W X^T H) + tr(WW^T H^T H)  (+ regularisation), combined in
 * fp64.  One call enqueues everything; nothing synchronises.  Replaces the batch loop of
 * run_nmf (cnmf.py:819) for R replicates. */
int cnmf_b200_frob_iterate(const cnmf_b200_frob_problem* p, int32_t n_iters, void* stream);

/* err2[r] = objective (x2) of the CURRENT state; the factors are not modified (gramW/gramH are
 * refreshed).  Gives the initial error that normalises run_nmf's convergence test. */
int cnmf_b200_frob_objective(const cnmf_b200_frob_problem* p, void* stream);

/* Row-sharded objective (cells split over ranks, W replicated): _parts leaves the rank-local terms --
 * sums[r*3+0] = <W_r, H_r^T X> over the local cells, sums[r*3+1] = sum(W_r), sums[r*3+2] = local sum(H_r),
 * gramH = local H^T H, gramW = W W^T -- the caller adds sums[:,0], sums[:,2] and gramH over ranks
 * (ncclAllReduce, sum) and _combine forms err2 with the global ||X||^2. */
int cnmf_b200_frob_objective_parts(const cnmf_b200_frob_problem* p, double* sums, void* stream);
int cnmf_b200_frob_objective_combine(double normX2, const double* sums, const float* gramW, const float* gramH,
                                     int32_t R, int32_t Kp, float l1_H, float l2_H, float l1_W, float l2_W,
                                     double* err2, void* stream);

/* ------------------------------------------------------------------------------------------
 * Online mode building blocks (run_nmf(mode='online'), the mode cNMF itself configures at
 * cnmf.py:757-771).  Per chunk of online_chunk_size cells: P = x W^T once, inner MU loop on h until
 * its relative change < h_tol (the reference's own fit_H_online, cnmf.py:355-378), A += h^T h,
 * B += h^T x, inner MU loop on W against (A, B) until w_tol.
 * ---------------------------------------------------------------------------------------- */

/* out[R*Kp x M] (= or +=) F[R*Kp x Kd] * A[M x Kd]^T on the tensor cores (tf32 operands, fp32
 * accumulate); accumulate != 0 adds to out.  The GEMMs x @ W.T and h.T @ x of the online loop. */
int cnmf_b200_dense_gemm_tf32(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                              float* out, int64_t ldo, const uint8_t* active, int32_t R, int32_t Kp,
                              int32_t accumulate, void* stream);
/* dst = round-to-nearest tf32 of a [rows x cols] block with row pitch ld (both arrays). */
int cnmf_b200_round_tf32_2d(float* dst, const float* src, int64_t rows, int64_t cols, int64_t ld, void* stream);
/* y += x  (A += h^T h). */
int cnmf_b200_axpy(float* y, const float* x, int64_t n, void* stream);

/* n_iters x { s <- s * max(p - l1, 0) / (s G + l2 s) on every column of state [R*Kp x M];
 *             rel_r = ||new - old||_F / (||old||_F + eps);  iters[r] += 1;  done[r] = (rel_r < tol) }
 * for the replicates with done[r] == 0 (and active[r] != 0).  n_pending[0] = replicates still running
 * after the last iteration.  (cnmf.py:365-378; also the W loop on the sufficient statistics.) */
int64_t cnmf_b200_online_inner_workspace_bytes(int64_t M, int32_t R);
int cnmf_b200_online_inner(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram, uint8_t* done,
                           const uint8_t* active, int64_t M, int32_t R, int32_t Kp, float l1, float l2, float eps,
                           float tol, int32_t n_iters, int32_t* iters, int32_t* n_pending, void* workspace,
                           void* stream);

/* One chunk of an online pass in one call: what nmf-torch's online loop does per chunk of online_chunk_size cells
 * (inner loop on the chunk's usages against the cached x W^T, sufficient statistics A += h^T h, B += h^T x, inner
 * loop on W against (A, B); loop and stopping rule: cnmf.py:355-378).  Pointers are device pointers; x / xt / Ht
 * point at THIS chunk's rows (x: [c x ldx]) or columns (xt: [G x ldxt], Ht: [R*Kp x ldh]). */
#define CNMF_B200_ONLINE_REFRESH 0   /* P = x W^T (+ x (W - tf32 W)^T), inner loop on h (max_iter may be 0), cross += <h, P>,
                                        gramH_acc += h^T h: end-of-pass refresh of H and the objective's terms */
#define CNMF_B200_ONLINE_USAGE 1     /* P = x W^T, inner loop on h, Ac = h^T h, Bc = h^T x                          */
#define CNMF_B200_ONLINE_SPECTRA 2   /* A += Ac, B += Bc, inner loop on W, Wop = tf32(W), Wlo, gramW = W W^T        */
typedef struct cnmf_b200_online_chunk_args {
  const float* x;  int64_t ldx;
  const float* xt; int64_t ldxt;
  int64_t c, n_genes;
  int32_t R, Kp;
  float* Wall; float* Wop; int64_t ldw;
  float* Ht;   int64_t ldh;        /* this chunk's columns of the resident usages [R*Kp x ldh] */
  float* Hs;   float* Hopc;        /* compact scratch copies of the chunk's usages, fp32 / tf32: [R*Kp x ldp] */
  float* Pt;   int64_t ldp;        /* scratch [R*Kp x ldp], ldp >= c, a multiple of 4 */
  float* Wlo;                      /* [R*Kp x ldw] tf32(W - tf32 W), or NULL (then P carries tf32 rounding of W) */
  double* cross;                   /* [R] += <h, x W^T> in the refresh phase, or NULL */
  float* gramH_acc;                /* [R x Kp x Kp] += h^T h in the refresh phase, or NULL */
  float* gramW;                    /* [R x Kp x Kp] */
  float* A;  float* Ac;            /* [R x Kp x Kp] statistics of the pass / of this chunk */
  float* Bst; float* Bc;           /* [R*Kp x ldw] */
  const uint8_t* active;           /* [R] or NULL */
  float l1_H, l2_H, l1_W, l2_W, eps, h_tol, w_tol;
  int32_t max_iter;
  int32_t* iters_h; int32_t* iters_w;   /* [R], incremented */
  void* ws_loop;                   /* cnmf_b200_online_loop_workspace_bytes(R) */
  void* ws_gram;                   /* cnmf_b200_gram_workspace_bytes(R, Kp, max(c, n_genes)) */
} cnmf_b200_online_chunk_args;
int cnmf_b200_online_chunk(const cnmf_b200_online_chunk_args* a, int32_t phase, void* stream);

/* The same loop run TO CONVERGENCE in one launch, no host round trip: every replicate iterates until rel_r < tol or
 * max_iter, decided on the device (the replicate's CTAs meet at a counter in global memory after every iteration; a
 * cooperative launch keeps them co-resident).  iters[r] += iterations run; max_iters_seen (device, optional) is raised
 * to the largest of them.  Same arithmetic and stopping rule as cnmf_b200_online_inner (cnmf.py:365-378).
 * workspace: cnmf_b200_online_loop_workspace_bytes(R) bytes. */
int64_t cnmf_b200_online_loop_workspace_bytes(int32_t R);
int cnmf_b200_online_loop(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram,
                          const uint8_t* active, int64_t M, int32_t R, int32_t Kp, float l1, float l2, float eps,
                          float tol, int32_t max_iter, int32_t* iters, unsigned long long* max_iters_seen,
                          void* workspace, void* stream);

/* ------------------------------------------------------------------------------------------
 * Sparse Kullback-Leibler engine (replicate-batched, CSR X).  Replaces run_nmf(beta_loss=
 * 'kullback-leibler', algo='mu') called at cnmf.py:819; update rule sklearn _nmf.py:562-607, 640-678
 * restated with the reference's epsilon rule (cnmf.py:370-371).  Factors are replicate-interleaved:
 *     Hr [N x R*Kp]  (row = cell),   Wt [G x R*Kp]  (row = gene),   element (row, r*Kp + k).
 * ---------------------------------------------------------------------------------------- */

/* out2[0] = sum of the stored values, out2[1] = sum x log x  (fp64): the constants of the KL objective
 * (sklearn _nmf.py:113-157). */
int cnmf_b200_kl_constants(const float* val, int64_t nnz, double* out2, void* stream);

/* dst (fp16) = src rounded to nearest, clamped at 65504.  The gathered copy of a factor. */
int cnmf_b200_to_half(const float* src, void* dst, int64_t n, void* stream);

/* One half-step: for every CSR row m < M and replicate r,
 *   state[m, r, k] <- state[m, r, k] * max(sum_j q_mj F[j, r, k] - l1, 0) / (fsum[r, k] + l2 state[m, r, k]),
 *   q_mj = x_mj / max(sum_k state[m, r, k] F[j, r, k], eps) over the stored entries of row m,
 * rate 0 where the denominator is below eps.  fsum[r*Kp + k] = sum_j F[j, r, k] (from the fp32 master).
 * F_half is the fp16 copy of the fixed factor (the per-non-zero gather is what bounds this kernel);
 * state_half (may be NULL) receives the fp16 copy of the updated rows for the next half-step. */
int cnmf_b200_kl_half_step(const int32_t* row_ptr, const int32_t* col_idx, const float* val, int64_t M,
                           const void* F_half, float* state, void* state_half, const float* fsum,
                           const uint8_t* active, int32_t R, int32_t Kp, float l1, float l2, float eps,
                           void* stream);

typedef struct cnmf_b200_kl_problem {
  int64_t n_cells, n_genes, nnz;
  const int32_t* row_ptr;    /* CSR of X   : [N+1], [nnz], [nnz] */
  const int32_t* col_idx;
  const float* val;
  const int32_t* t_row_ptr;  /* CSR of X^T : [G+1], [nnz], [nnz]  (X is static: built once) */
  const int32_t* t_col_idx;
  const float* t_val;
  int32_t R, K, Kp;          /* Kp: K rounded up to an EVEN number here (<= CNMF_B200_MAX_KP); padding columns zero */
  int32_t impl;              /* CNMF_B200_KL_IMPL_GATHER (0) or CNMF_B200_KL_IMPL_SLAB (1, what the host layer selects) */
  float* Hr;                 /* [N x R*Kp] updated in place */
  float* Wt;                 /* [G x R*Kp] updated in place */
  const uint8_t* active;     /* [R] or NULL */
  float l1_H, l2_H, l1_W, l2_W;
  float eps;
  int32_t warm;              /* 0: cnmf_b200_kl_iterate rebuilds its tables (row segments, work split, packed entries)
                                and the fp16 copies of both factors from the fp32 masters -- stateless, the default.
                                1: the caller promises that the previous call on this problem was a cnmf_b200_kl_iterate
                                with the same X, R, Kp, impl and workspace, that the workspace was not touched since and
                                that nobody else wrote Hr / Wt: the tables and copies in the workspace are reused
                                (0.13 ms per call at 100 000 x 2000 -- 9 % of an iteration at R = 13). */
  double sumX, xlogx;        /* from cnmf_b200_kl_constants */
  double* err2;              /* [R] out: 2 * (generalised KL divergence + L1 and 0.5*l2*||.||^2 terms); may be NULL */
  void* workspace;
  int64_t workspace_bytes;
} cnmf_b200_kl_problem;

/* The two kernel families behind cnmf_b200_kl_iterate / _objective (same update rule, same results within the
 * tolerances of the tests):
 *   GATHER -- one thread per (row segment, replicate), lanes = consecutive replicates, the fp16 copy of the fixed
 *             factor gathered from L2 in coalesced runs (also what the standalone cnmf_b200_kl_half_step runs):
 *             7.45 ms per batched iteration at 100 000 x 2000, K = 9, R = 100; 1.92 ms at 20 000 cells;
 *   SLAB   -- the fixed factor's fp16 copy of one replicate group and one block of columns is staged in shared
 *             memory (bank-conflict-free line layout) and gathered from there; the (column, value) pairs of X are
 *             packed per row segment once and streamed by bulk copies; fp16 x fp16 + fp32 FMAs (FHFMA); 70 x less
 *             L2 traffic: 6.94 ms at 100 000 cells, 1.50 ms at 20 000 (R = 13: 1.37 vs 1.52 and 0.35 vs 0.48). */
#define CNMF_B200_KL_IMPL_GATHER 0
#define CNMF_B200_KL_IMPL_SLAB 1

int64_t cnmf_b200_kl_workspace_bytes(const cnmf_b200_kl_problem* p);
/* n_iters x { column sums of W; H half-step; column sums of H; W half-step }, then the objective. */
int cnmf_b200_kl_iterate(const cnmf_b200_kl_problem* p, int32_t n_iters, void* stream);
/* err2 of the current state; factors untouched. */
int cnmf_b200_kl_objective(const cnmf_b200_kl_problem* p, void* stream);
/* The gathered copies are fp16: a factor entry above 65504 (or a ratio x / (WH) that overflows fp16 on the slab
 * path) cannot be represented.  Nothing is clamped silently: the kernels raise a flag in the workspace instead, the
 * affected entries turn into inf/NaN, and this call reads the flag back (synchronises the stream): 0 = all values in
 * range since the last clear.  run_nmf has no counterpart (fp32 throughout); the host layer turns a raised flag into
 * an error that says so.  */
int cnmf_b200_kl_range_flag(const cnmf_b200_kl_problem* p, int32_t* flag_host, void* stream);
int cnmf_b200_kl_range_flag_clear(const cnmf_b200_kl_problem* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * Consensus kernel set.  Replaces the library calls of cNMF.consensus' front half
 * (cnmf.py:1056-1090) on the float32 dtype flow of this fork.
 * ---------------------------------------------------------------------------------------- */

/* out = rows of S divided by their L2 norm, evaluated exactly as the reference's float32 pandas
 * expression (cnmf.py:1056): fp32 squares summed in gene order, fp32 sqrt, fp32 division. */
int cnmf_b200_l2_normalize(const float* S, int64_t lds, int32_t M, int32_t G, float* out, int64_t ldo,
                           void* stream);

/* D[M x M] = euclidean_distances(X) for float32 X as sklearn computes it (cnmf.py:1065;
 * sklearn/metrics/pairwise.py:376-427, 567-638): -2 x.y + |x|^2 + |y|^2 accumulated in fp64,
 * cast to fp32, clamped at 0, diagonal set to 0, square root in fp32.  sqnorm: [M] doubles scratch. */
int cnmf_b200_pairwise_dist(const float* X, int64_t ld, int32_t M, int32_t G, float* D, int64_t ldd,
                            double* sqnorm, void* stream);

/* density[i] = (sum of the n_neighbors+1 smallest entries of row i of D) / n_neighbors
 * = np.argpartition(D, n+1)[:, :n+1] -> take -> sum(1)/n   (cnmf.py:1067-1072).
 * nbr (optional, [M x (n+1)] int32) receives the selected column indices (ties at the boundary
 * resolved towards the lower index).  M <= 51200. */
int cnmf_b200_topk_density(const float* D, int64_t ldd, int32_t M, int32_t n_neighbors, float* density,
                           int32_t* nbr, void* stream);

/* mean[g] = mean over rows (fp32 accumulation in row order, as numpy's axis-0 reduction);
 * X -= mean in place.  KMeans.fit centres its input this way (sklearn/cluster/_kmeans.py:1486-1491). */
int cnmf_b200_col_center(float* X, int64_t ld, int32_t M, int32_t G, float* mean, void* stream);

/* Lloyd iterations of KMeans(n_clusters=k, n_init, ...) (cnmf.py:1082-1083), all n_init restarts
 * batched: centres [n_init x k x G], labels [n_init x M].  k <= 32.
 *   assign : labels = argmin_j fl32(|c_j|^2 - 2 x.c_j), first minimum wins; n_changed[i] += #labels
 *            that changed (zero it before the call)       (_k_means_lloyd.pyx _update_chunk_dense)
 *   update : centres_new = per-cluster sum * fl32(1/count), center_shift[i*k+j] = |old - new|_2,
 *            counts[i*k+j] = members (0 => empty cluster, centre left at 0 for the caller to relocate)
 *   inertia: partials [n_init x cnmf_b200_kmeans_inertia_blocks(M)] doubles, sum_i |x_i - c_label(i)|^2
 * active ([n_init] bytes, may be NULL) switches converged restarts off. */
int64_t cnmf_b200_kmeans_workspace_bytes(int32_t n_init, int32_t k, int32_t G);
int cnmf_b200_kmeans_assign(const float* X, int64_t ld, int32_t M, int32_t G, const float* centers,
                            int32_t n_init, int32_t k, const uint8_t* active, int32_t* labels,
                            int32_t* n_changed, void* stream);
int cnmf_b200_kmeans_update(const float* X, int64_t ld, int32_t M, int32_t G, const int32_t* labels,
                            int32_t n_init, int32_t k, const uint8_t* active, const float* centers_old,
                            float* centers_new, float* center_shift, int32_t* counts, void* workspace,
                            void* stream);
int64_t cnmf_b200_kmeans_inertia_blocks(int32_t M);
int cnmf_b200_kmeans_inertia(const float* X, int64_t ld, int32_t M, int32_t G, const float* centers,
                             int32_t n_init, int32_t k, const int32_t* labels, double* partials,
                             void* stream);

/* k-means++ seeding of all restarts (sklearn/cluster/_kmeans.py:180-278, reached from cnmf.py:1083).
 * The host draws the random numbers in sklearn's order from RandomState(random_state): first_idx[i] =
 * choice(M) and rnd[i, c, :] = uniform(size=n_trials) for c = 1..k-1 (n_trials = 2 + int(log k)); the
 * device does the distance / potential / searchsorted arithmetic.  chosen [n_init x k] receives the
 * sample indices, centers [n_init x k x G] the rows.  n_init * n_trials <= 64, n_trials <= 8. */
int64_t cnmf_b200_kmeanspp_workspace_bytes(int32_t M, int32_t n_init, int32_t n_trials);
int cnmf_b200_kmeanspp(const float* Xc, int64_t ld, int32_t M, int32_t G, int32_t n_init, int32_t k,
                       int32_t n_trials, const int32_t* first_idx, const double* rnd, int32_t* chosen,
                       float* centers, void* workspace, void* stream);

/* median[j, g] = median over the members of cluster j (rows order[offsets[j] : offsets[j+1]]) of
 * X[:, g]  (= l2_spectra.groupby(labels).median(), cnmf.py:1087); normalized (may be NULL) = rows of
 * median divided by their sum (cnmf.py:1090).  Clusters of up to 3000 members. */
int cnmf_b200_cluster_median(const float* X, int64_t ld, int32_t G, const int32_t* order,
                             const int32_t* offsets, int32_t k, int32_t max_members, float* median,
                             float* normalized, void* stream);

/* ------------------------------------------------------------------------------------------
 * Rows after the hot path: k-selection statistics and the OLS accumulation.
 * ---------------------------------------------------------------------------------------- */

/* sil[i] = silhouette coefficient of sample i given the distance matrix D (cnmf_b200_pairwise_dist), labels in
 * 0..k-1 and the cluster sizes: silhouette_score(l2_spectra, labels, metric='euclidean') at cnmf.py:1097 is
 * their mean.  k <= 32. */
int cnmf_b200_silhouette(const float* D, int64_t ldd, int32_t M, const int32_t* labels, const int32_t* counts,
                         int32_t k, float* sil, void* stream);

/* The fp64 accumulations of efficient_ols_all_cols (cnmf.py:56-126): xty[k, j] = sum_i X[i, k] Y[i, j] over
 * the stored entries of the CSC target matrix Y (col_ptr int64 [n_targets+1], row_idx, val), with
 * ysum[j] = sum_i Y[i, j] and ysq[j] = sum_i Y[i, j]^2 for the global column mean / variance.
 * X: [n_samples x K] fp32 row-major, K <= 32. */
int cnmf_b200_ols_xty(const int64_t* col_ptr, const int32_t* row_idx, const float* val, int32_t n_targets,
                      const float* X, int32_t K, double* xty, double* ysum, double* ysq, void* stream);

/* Number of kernel launches the last cnmf_b200_frob_iterate on this thread enqueued. */
int64_t cnmf_b200_last_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Host-side helper of the on-disk path (no device work): n_files `.df.npz` archives as save_df_to_npz writes them
 * (/root/reference/src/cnmf/cnmf.py:32-34: np.savez_compressed(data=, index=, columns=)), compressed and written by up
 * to 16 host threads.  File i gets member data.npy = data_hdr[i] (the .npy header of its shape, made by the caller)
 * followed by data_bytes[i] bytes at data[i]; index.npy = index_npy[i] (a complete .npy image); columns.npy = the
 * caller's PRE-DEFLATED raw stream (the gene names: identical for every replicate file of a run, compressed once).
 * Members of 2 GiB or more are refused (no zip64).
 * ---------------------------------------------------------------------------------------- */
int cnmf_b200_write_npz_batch(int64_t n_files, const char* const* paths, const void* const* data,
                              const int64_t* data_bytes, const char* const* data_hdr, const int64_t* data_hdr_len,
                              const char* const* index_npy, const int64_t* index_npy_len, const char* columns_comp,
                              int64_t columns_comp_len, int64_t columns_raw_len, uint32_t columns_crc, int32_t level,
                              int32_t threads);

/* ------------------------------------------------------------------------------------------
 * Host-side helper of the on-disk path (no device work): a float table as tab-separated text, byte for byte what
 * save_df_to_text writes (DataFrame.to_csv(sep="\t"), /root/reference/src/cnmf/cnmf.py:44-45): numpy's shortest
 * round-trip repr of every value (positional for 1e-4 <= |x| < 1e16 -- 1e6 for float --, else scientific; NaN -> empty field), rows
 * formatted by up to 16 host threads.  header: the first line incl. its '\n' (index name, tab, column labels);
 * index_blob / index_off[n + 1]: the row labels, concatenated; data: [n x k] row-major with leading dimension ld,
 * float (is_f64 = 0) or double (1).  threads <= 0: hardware concurrency.
 * ---------------------------------------------------------------------------------------- */
int cnmf_b200_write_tsv(const char* path, const char* header, int64_t header_len, const char* index_blob,
                        const int64_t* index_off, const void* data, int32_t is_f64, int64_t n, int64_t k, int64_t ld,
                        int32_t threads);

#ifdef __cplusplus
}
#endif
#endif /* CNMF_B200_H_ */
