// Consensus kernel set: pairwise Euclidean distances (fp64 accumulate), top-(n+1) local density,
// column centring, batched Lloyd k-means (assign / update / inertia) and per-cluster medians.
// Replaces the sklearn / numpy / pandas calls at /root/reference/src/cnmf/cnmf.py:1065-1090; the
// numerical recipe of each kernel follows what those library calls do for float32 input
// (sklearn/metrics/pairwise.py:376-427, 567-638; sklearn/cluster/_k_means_lloyd.pyx:_update_chunk_dense;
// sklearn/cluster/_kmeans.py:_kmeans_single_lloyd) so that kNN sets and labels come out identical.
#include <float.h>

#include "common.cuh"

namespace cnmf {

// ---------------------------------------------------------------------------------------------
// K8: D = sqrt(max(0, -2 X X^T + xx + xx^T)) with fp64 accumulation, cast to fp32 BEFORE the clamp
// and the square root, diagonal forced to 0 (pairwise.py:404-425, 630-636).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) row_sqnorm64_kernel(const float* __restrict__ X, int64_t ld, int M,
                                                           int G, double* __restrict__ out) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  double acc = 0.0;
  for (int g = lane; g < G; g += 32) {
    double v = X[(int64_t)row * ld + g];
    acc = fma(v, v, acc);
  }
  acc = warp_sum(acc);
  if (lane == 0) out[row] = acc;
}

constexpr int DT = 64, DK = 16;

__global__ void __launch_bounds__(256) pairwise_dist_kernel(const float* __restrict__ X, int64_t ld, int M, int G,
                                                            const double* __restrict__ xx,
                                                            float* __restrict__ D, int64_t ldd, int nt) {
  // upper-triangular tile index -> (ti, tj), ti <= tj
  int rem = blockIdx.x, ti = 0;
  while (rem >= nt - ti) { rem -= nt - ti; ++ti; }
  const int tj = ti + rem;
  __shared__ __align__(16) double As[DK][DT + 2];
  __shared__ __align__(16) double Bs[DK][DT + 2];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int i0 = ti * DT, j0 = tj * DT;
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
  const int lrow = tid >> 2, lk = (tid & 3) * 4;
  for (int k0 = 0; k0 < G; k0 += DK) {
    float va[4] = {0.f, 0.f, 0.f, 0.f}, vb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + lk + q;
      if (k < G) {
        if (i0 + lrow < M) va[q] = X[(int64_t)(i0 + lrow) * ld + k];
        if (j0 + lrow < M) vb[q] = X[(int64_t)(j0 + lrow) * ld + k];
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      As[lk + q][lrow] = (double)va[q];
      Bs[lk + q][lrow] = (double)vb[q];
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < DK; ++kk) {
      double a[4], b[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { a[q] = As[kk][ty * 4 + q]; b[q] = Bs[kk][tx * 4 + q]; }
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[p][q] = fma(a[p], b[q], acc[p][q]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int i = i0 + ty * 4 + p;
    if (i >= M) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int j = j0 + tx * 4 + q;
      if (j >= M) continue;
      double d = -2.0 * acc[p][q];
      d += xx[i];
      d += xx[j];
      float f = (float)d;
      f = fmaxf(f, 0.f);
      f = (i == j) ? 0.f : sqrtf(f);
      D[(int64_t)i * ldd + j] = f;
      if (ti != tj) D[(int64_t)j * ldd + i] = f;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// K9: per row, the n+1 smallest distances (cnmf.py:1067) by radix select on the fp32 bit patterns
// (all values >= 0, so bit order == numeric order); density = sum / n (cnmf.py:1069-1072).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) topk_density_kernel(const float* __restrict__ D, int64_t ldd, int M, int kth,
                                                           int n_div, float* __restrict__ density,
                                                           int* __restrict__ nbr) {
  extern __shared__ uint32_t srow[];
  __shared__ int hist[256];
  __shared__ uint32_t s_prefix;
  __shared__ int s_remaining, s_less;
  __shared__ double sred[8];
  const int row = blockIdx.x;
  const float* drow = D + (int64_t)row * ldd;
  for (int j = threadIdx.x; j < M; j += 256) srow[j] = __float_as_uint(drow[j]);
  if (threadIdx.x == 0) { s_prefix = 0u; s_remaining = kth; s_less = 0; }
  __syncthreads();
  uint32_t mask = 0u;
  for (int pass = 3; pass >= 0; --pass) {
    hist[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    const int shift = pass * 8;
    for (int j = threadIdx.x; j < M; j += 256) {
      const uint32_t b = srow[j];
      if ((b & mask) == prefix) atomicAdd(&hist[(b >> shift) & 255u], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int cum = 0, b = 0;
      const int remaining = s_remaining;
      for (; b < 256; ++b) {
        if (cum + hist[b] >= remaining) break;
        cum += hist[b];
      }
      s_remaining = remaining - cum;
      s_less += cum;
      s_prefix = prefix | ((uint32_t)b << shift);
    }
    __syncthreads();
    mask |= 255u << shift;
  }
  const uint32_t T = s_prefix;     // bit pattern of the kth smallest value
  const int n_less = s_less;       // values strictly smaller
  const int ties = kth - n_less;   // copies of T that belong to the selection
  double acc = 0.0;
  for (int j = threadIdx.x; j < M; j += 256) {
    const uint32_t b = srow[j];
    if (b < T) acc += (double)__uint_as_float(b);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += sred[w];
    t += (double)ties * (double)__uint_as_float(T);
    density[row] = (float)t / (float)n_div;
  }
  if (nbr != nullptr && threadIdx.x < 32) {
    // ordered compaction by one warp: indices with value < T, then the first `ties` equal to T
    const int lane = threadIdx.x;
    int* out = nbr + (int64_t)row * kth;
    int n_out = 0, n_tie = 0;
    for (int j0 = 0; j0 < M; j0 += 32) {
      const int j = j0 + lane;
      const uint32_t b = j < M ? srow[j] : 0xFFFFFFFFu;
      const unsigned less = __ballot_sync(0xffffffffu, b < T);
      if (b < T) out[n_out + __popc(less & ((1u << lane) - 1u))] = j;
      n_out += __popc(less);
    }
    for (int j0 = 0; j0 < M && n_tie < ties; j0 += 32) {
      const int j = j0 + lane;
      const uint32_t b = j < M ? srow[j] : 0xFFFFFFFFu;
      const unsigned eq = __ballot_sync(0xffffffffu, b == T);
      const int pos = n_tie + __popc(eq & ((1u << lane) - 1u));
      if (b == T && pos < ties) out[n_less + pos] = j;
      n_tie += __popc(eq);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// K10 prologue: X -= mean over rows, the mean accumulated in fp32 in row order like numpy's
// axis-0 reduction (sklearn/cluster/_kmeans.py:1486-1491), divided in double then rounded.
// ---------------------------------------------------------------------------------------------
__global__ void col_mean_seq_kernel(const float* __restrict__ X, int64_t ld, int M, int G, float* __restrict__ mean) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  float acc = 0.f;
  for (int m = 0; m < M; ++m) acc += X[(int64_t)m * ld + g];
  mean[g] = (float)((double)acc / (double)M);
}

__global__ void sub_mean_kernel(float* __restrict__ X, int64_t ld, int M, int G, const float* __restrict__ mean) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)M * G) return;
  const int m = (int)(i / G), g = (int)(i - (int64_t)m * G);
  X[(int64_t)m * ld + g] -= mean[g];
}

// ---------------------------------------------------------------------------------------------
// K10 E-step, batched over inits: label = argmin_j fl32(||c_j||^2 - 2 x.c_j), first minimum wins
// (_k_means_lloyd.pyx:_update_chunk_dense).  CTA = 128 samples x (up to 32) centres of ONE init.
// ---------------------------------------------------------------------------------------------
constexpr int KA_S = 128, KA_C = 32, KA_G = 32;

__global__ void __launch_bounds__(256) kmeans_assign_kernel(const float* __restrict__ X, int64_t ld, int M, int G,
                                                            const float* __restrict__ centers, int k,
                                                            const uint8_t* __restrict__ active,
                                                            int* __restrict__ labels, int* __restrict__ n_changed) {
  const int init = blockIdx.y;
  if (active != nullptr && active[init] == 0) return;
  __shared__ float Xs[KA_S][KA_G + 1];
  __shared__ float Cs[KA_C][KA_G + 1];
  __shared__ float dist[KA_S][KA_C + 1];
  __shared__ float cn[KA_C];
  const float* C = centers + (int64_t)init * k * G;
  const int tid = threadIdx.x, sx = tid & 31, cy = tid >> 5;
  const int m0 = blockIdx.x * KA_S;
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
  float cnorm = 0.f;
  for (int g0 = 0; g0 < G; g0 += KA_G) {
    for (int i = tid; i < KA_S * KA_G; i += 256) {
      const int s = i / KA_G, g = i - s * KA_G;
      Xs[s][g] = (m0 + s < M && g0 + g < G) ? X[(int64_t)(m0 + s) * ld + g0 + g] : 0.f;
    }
    for (int i = tid; i < KA_C * KA_G; i += 256) {
      const int c = i / KA_G, g = i - c * KA_G;
      Cs[c][g] = (c < k && g0 + g < G) ? C[(int64_t)c * G + g0 + g] : 0.f;
    }
    __syncthreads();
    if (tid < KA_C) {
#pragma unroll 8
      for (int g = 0; g < KA_G; ++g) cnorm = fmaf(Cs[tid][g], Cs[tid][g], cnorm);
    }
#pragma unroll 8
    for (int g = 0; g < KA_G; ++g) {
      float xv[4], cv[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) { xv[a] = Xs[sx + 32 * a][g]; cv[a] = Cs[cy * 4 + a][g]; }
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(xv[a], cv[b], acc[a][b]);
    }
    __syncthreads();
  }
  if (tid < KA_C) cn[tid] = cnorm;
  __syncthreads();
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) dist[sx + 32 * a][cy * 4 + b] = fmaf(-2.f, acc[a][b], cn[cy * 4 + b]);
  __syncthreads();
  int changed = 0;
  if (tid < KA_S && m0 + tid < M) {
    float best = dist[tid][0];
    int lab = 0;
    for (int j = 1; j < k; ++j) {
      const float d = dist[tid][j];
      if (d < best) { best = d; lab = j; }
    }
    int* lp = labels + (int64_t)init * M + m0 + tid;
    changed = (*lp != lab);
    *lp = lab;
  }
  changed = __syncthreads_count(changed);
  if (tid == 0 && changed) atomicAdd(&n_changed[init], changed);
}

// ---------------------------------------------------------------------------------------------
// K10 M-step: per-cluster sums in sample order inside each segment, segments combined in order.
// ---------------------------------------------------------------------------------------------
constexpr int KU_G = 128;   // genes per CTA (one per thread)
constexpr int KU_SEG = 32;  // sample segments

__global__ void __launch_bounds__(KU_G) kmeans_update_kernel(const float* __restrict__ X, int64_t ld, int M, int G,
                                                             const int* __restrict__ labels, int k,
                                                             const uint8_t* __restrict__ active,
                                                             float* __restrict__ part, int* __restrict__ cnt_part) {
  const int init = blockIdx.z, seg = blockIdx.y;
  if (active != nullptr && active[init] == 0) return;
  __shared__ float acc[KA_C][KU_G];
  __shared__ int cnt[KA_C];
  const int t = threadIdx.x, g = blockIdx.x * KU_G + t;
  for (int j = 0; j < k; ++j) acc[j][t] = 0.f;
  if (t < KA_C) cnt[t] = 0;
  __syncthreads();
  const int per = (M + KU_SEG - 1) / KU_SEG;
  const int mb = seg * per, me = min(M, mb + per);
  const int* lab = labels + (int64_t)init * M;
  for (int m = mb; m < me; ++m) {
    const int l = lab[m];
    if (g < G) acc[l][t] += X[(int64_t)m * ld + g];
    if (t == 0) cnt[l] += 1;
  }
  __syncthreads();
  float* out = part + (((int64_t)init * KU_SEG + seg) * k) * G;
  if (g < G)
    for (int j = 0; j < k; ++j) out[(int64_t)j * G + g] = acc[j][t];
  if (blockIdx.x == 0 && t < k) cnt_part[((int64_t)init * KU_SEG + seg) * k + t] = cnt[t];
}

// one CTA per (centre, init): new centre = sum * fl(1/count) (_k_means_common.pyx:_average_centers),
// centre shift = sqrt(sum (old-new)^2) in fp32; count 0 leaves the sum (0) and reports the cluster empty.
__global__ void __launch_bounds__(128) kmeans_finalize_kernel(const float* __restrict__ part,
                                                              const int* __restrict__ cnt_part, int G, int k,
                                                              const uint8_t* __restrict__ active,
                                                              const float* __restrict__ c_old, float* __restrict__ c_new,
                                                              float* __restrict__ shift, int* __restrict__ counts) {
  const int j = blockIdx.x, init = blockIdx.y;
  if (active != nullptr && active[init] == 0) return;
  __shared__ float red[4];
  int count = 0;
  for (int s = 0; s < KU_SEG; ++s) count += cnt_part[((int64_t)init * KU_SEG + s) * k + j];
  const float alpha = count > 0 ? 1.0f / (float)count : 1.0f;
  float sh = 0.f;
  for (int g = threadIdx.x; g < G; g += 128) {
    float tot = 0.f;
    for (int s = 0; s < KU_SEG; ++s) tot += part[(((int64_t)init * KU_SEG + s) * k + j) * G + g];
    const float cnew = tot * alpha;
    const int64_t o = ((int64_t)init * k + j) * G + g;
    const float d = c_old[o] - cnew;
    sh = fmaf(d, d, sh);
    c_new[o] = cnew;
  }
  sh = warp_sum(sh);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sh;
  __syncthreads();
  if (threadIdx.x == 0) {
    shift[init * k + j] = sqrtf(red[0] + red[1] + red[2] + red[3]);
    counts[init * k + j] = count;
  }
}

__global__ void __launch_bounds__(256) kmeans_inertia_kernel(const float* __restrict__ X, int64_t ld, int M, int G,
                                                             const float* __restrict__ centers, int k,
                                                             const int* __restrict__ labels,
                                                             double* __restrict__ part) {
  // one warp per sample, 8 samples per CTA; partial per CTA
  const int init = blockIdx.y;
  __shared__ double red[8];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m = blockIdx.x * 8 + w;
  double acc = 0.0;
  if (m < M) {
    const int l = labels[(int64_t)init * M + m];
    const float* c = centers + ((int64_t)init * k + l) * G;
    const float* x = X + (int64_t)m * ld;
    for (int g = lane; g < G; g += 32) {
      const double d = (double)x[g] - (double)c[g];
      acc = fma(d, d, acc);
    }
  }
  acc = warp_sum(acc);
  if (lane == 0) red[w] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 8; ++i) t += red[i];
    part[(int64_t)init * gridDim.x + blockIdx.x] = t;
  }
}

// ---------------------------------------------------------------------------------------------
// K11: per-cluster per-gene median (cnmf.py:1087) by rank counting, then row-normalise (cnmf.py:1090)
// ---------------------------------------------------------------------------------------------
constexpr int MD_G = 16;  // genes per CTA

__global__ void __launch_bounds__(256) cluster_median_kernel(const float* __restrict__ X, int64_t ld, int G,
                                                             const int* __restrict__ order,
                                                             const int* __restrict__ offsets,
                                                             float* __restrict__ median) {
  extern __shared__ float vals[];  // [n][MD_G + 1]
  const int j = blockIdx.y, g0 = blockIdx.x * MD_G;
  const int beg = offsets[j], n = offsets[j + 1] - beg;
  if (n <= 0) {
    if (threadIdx.x < MD_G && g0 + threadIdx.x < G) median[(int64_t)j * G + g0 + threadIdx.x] = nanf("");
    return;
  }
  __shared__ float lo[MD_G], hi[MD_G];
  for (int i = threadIdx.x; i < n * MD_G; i += 256) {
    const int r = i / MD_G, c = i - r * MD_G;
    vals[r * (MD_G + 1) + c] = (g0 + c < G) ? X[(int64_t)order[beg + r] * ld + g0 + c] : 0.f;
  }
  __syncthreads();
  const int r_lo = (n - 1) / 2, r_hi = n / 2;
  for (int i = threadIdx.x; i < n * MD_G; i += 256) {
    const int c = i % MD_G, r = i / MD_G;
    const float v = vals[r * (MD_G + 1) + c];
    int rank = 0;
    for (int q = 0; q < n; ++q) {
      const float u = vals[q * (MD_G + 1) + c];
      rank += (u < v) || (u == v && q < r);
    }
    if (rank == r_lo) lo[c] = v;
    if (rank == r_hi) hi[c] = v;
  }
  __syncthreads();
  if (threadIdx.x < MD_G && g0 + threadIdx.x < G)
    median[(int64_t)j * G + g0 + threadIdx.x] = (float)(((double)lo[threadIdx.x] + (double)hi[threadIdx.x]) * 0.5);
}

__global__ void __launch_bounds__(256) row_normalize_kernel(const float* __restrict__ in, int G, float* __restrict__ out) {
  const int j = blockIdx.x;
  __shared__ double red[8];
  __shared__ float s_total;
  double acc = 0.0;
  for (int g = threadIdx.x; g < G; g += 256) acc += (double)in[(int64_t)j * G + g];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w];
    s_total = (float)t;
  }
  __syncthreads();
  for (int g = threadIdx.x; g < G; g += 256) out[(int64_t)j * G + g] = in[(int64_t)j * G + g] / s_total;
}


// ---------------------------------------------------------------------------------------------
// a8: rows / ||row||_2 exactly as the reference's pandas expression evaluates it for float32
// (cnmf.py:1056): squares rounded to fp32, summed in fp32 in column order (pandas reduces the
// transposed block, i.e. sequentially over genes -- checked bit-for-bit against pandas in
// tests/test_oracle_golden.py), fp32 sqrt, fp32 division.  No FMA contraction.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) l2_normalize_kernel(const float* __restrict__ S, int64_t lds, int M, int G,
                                                           float* __restrict__ out, int64_t ldo) {
  __shared__ float tile[128][33];
  __shared__ float snorm[128];
  const int m0 = blockIdx.x * 128, t = threadIdx.x;
  float acc = 0.f;
  for (int g0 = 0; g0 < G; g0 += 32) {
    for (int i = t; i < 128 * 32; i += 128) {
      const int r = i >> 5, c = i & 31;
      tile[r][c] = (m0 + r < M && g0 + c < G) ? S[(int64_t)(m0 + r) * lds + g0 + c] : 0.f;
    }
    __syncthreads();
    const int w = min(32, G - g0);
    for (int c = 0; c < w; ++c) acc = __fadd_rn(acc, __fmul_rn(tile[t][c], tile[t][c]));
    __syncthreads();
  }
  snorm[t] = __fsqrt_rn(acc);
  __syncthreads();
  for (int64_t i = t; i < (int64_t)128 * G; i += 128) {
    const int r = (int)(i / G), c = (int)(i - (int64_t)r * G);
    if (m0 + r < M) out[(int64_t)(m0 + r) * ldo + c] = __fdiv_rn(S[(int64_t)(m0 + r) * lds + c], snorm[r]);
  }
}

// ---------------------------------------------------------------------------------------------
// K10 seeding: k-means++ of all restarts side by side (sklearn/cluster/_kmeans.py:180-278).  The
// random numbers come from the host (numpy RandomState stream, drawn in sklearn's order); distances
// follow _euclidean_distances for float32 (fp64 accumulate, cast, clamp); the cumulative sum that
// searchsorted scans is the sequential fp32 np.cumsum.
// ---------------------------------------------------------------------------------------------
constexpr int KPP_NC = 64;   // max candidates (n_init * n_trials) handled per pass
constexpr int KPP_GC = 64;   // genes per chunk

// One warp per restart walks the fp32 running sum (np.cumsum is a sequential fp32 sum, so the additions
// cannot be re-associated) and records, for each trial value, the first index whose cumulative sum reaches
// it: np.searchsorted(np.cumsum(closest), rand * pot), side='left', clipped.  The 32 lanes load 32
// consecutive values at once (coalesced) and every lane replays the same sequential adds from shuffles, so
// the dependent chain is the 4-cycle FADD, not a global-memory round trip per element.
__global__ void __launch_bounds__(32) kpp_pick_kernel(const float* __restrict__ closest, const float* __restrict__ pot,
                                                      const double* __restrict__ rnd, int M, int T, int step,
                                                      int ksteps, int* __restrict__ cand) {
  const int init = blockIdx.x, lane = threadIdx.x;
  constexpr int TMAX = 8;
  // (double)acc >= val  <=>  acc >= (smallest float that is >= val): the thresholds are rounded UP to float
  // once, so the sequential loop compares floats and stays off the fp64 pipe
  float val[TMAX];
  int found[TMAX];
  const double p = (double)pot[init];
  const float kNever = __int_as_float(0x7f800000);  // +inf: unused / already found
#pragma unroll
  for (int t = 0; t < TMAX; ++t) {
    val[t] = t < T ? __double2float_ru(rnd[((int64_t)init * ksteps + step) * T + t] * p) : kNever;
    found[t] = M - 1;
  }
  const float* c = closest + (int64_t)init * M;
  float acc = 0.f;
  float vmin = kNever;  // smallest threshold still pending: one compare per element on the common path
#pragma unroll
  for (int t = 0; t < TMAX; ++t) vmin = fminf(vmin, val[t]);
  int n_pending = T;
  for (int j0 = 0; j0 < M && n_pending > 0; j0 += 32) {
    const float mine = (j0 + lane < M) ? c[j0 + lane] : 0.f;
    const int n = min(32, M - j0);
    for (int i = 0; i < n; ++i) {
      acc = __fadd_rn(acc, __shfl_sync(0xffffffffu, mine, i));
      if (acc >= vmin) {
        vmin = kNever;
#pragma unroll
        for (int t = 0; t < TMAX; ++t) {
          if (acc >= val[t]) { found[t] = j0 + i; val[t] = kNever; --n_pending; }
          vmin = fminf(vmin, val[t]);
        }
      }
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int t = 0; t < TMAX; ++t)
      if (t < T) cand[init * T + t] = found[t];
  }
}

// distances from NC (<= 64) candidate rows to every point, as a tiled fp64 GEMM (64 points x 64 candidates
// per CTA, 4x4 per thread): dmin = min(closest, d); per-CTA potential partials [gridDim.x x NC]
__global__ void __launch_bounds__(256) kpp_dist_kernel(const float* __restrict__ X, int64_t ld, int M, int G,
                                                       const double* __restrict__ xx, const int* __restrict__ cand,
                                                       int NC, int T, const float* __restrict__ closest,
                                                       float* __restrict__ dmin, double* __restrict__ potpart) {
  __shared__ __align__(16) double As[DK][DT + 2];   // points
  __shared__ __align__(16) double Bs[DK][DT + 2];   // candidates
  __shared__ int scand[KPP_NC];
  __shared__ float spot[KPP_NC][17];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int i0 = blockIdx.x * DT;
  if (tid < KPP_NC) scand[tid] = tid < NC ? cand[tid] : -1;
  __syncthreads();
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
  const int lrow = tid >> 2, lk = (tid & 3) * 4;
  const int crow = scand[lrow];
  for (int k0 = 0; k0 < G; k0 += DK) {
    float va[4] = {0.f, 0.f, 0.f, 0.f}, vb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + lk + q;
      if (k < G) {
        if (i0 + lrow < M) va[q] = X[(int64_t)(i0 + lrow) * ld + k];
        if (crow >= 0) vb[q] = X[(int64_t)crow * ld + k];
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      As[lk + q][lrow] = (double)va[q];
      Bs[lk + q][lrow] = (double)vb[q];
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < DK; ++kk) {
      double a[4], b[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { a[q] = As[kk][ty * 4 + q]; b[q] = Bs[kk][tx * 4 + q]; }
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[p][q] = fma(a[p], b[q], acc[p][q]);
    }
    __syncthreads();
  }
  // thread (ty, tx): points i0 + ty*4 + p, candidates tx*4 + q
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int c = tx * 4 + q;
    float psum = 0.f;
    if (c < NC) {
      const int init = c / T;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const int m = i0 + ty * 4 + p;
        if (m < M) {
          double d = -2.0 * acc[p][q];
          d += xx[scand[c]];
          d += xx[m];
          float f = fmaxf((float)d, 0.f);
          if (closest != nullptr) f = fminf(closest[(int64_t)init * M + m], f);
          dmin[(int64_t)c * M + m] = f;
          psum += f;   // 4 values per thread; the cross-thread sum below is fp64
        }
      }
      spot[c][ty] = psum;
    }
  }
  __syncthreads();
  if (tid < NC) {
    double t = 0.0;
    for (int y = 0; y < 16; ++y) t += (double)spot[tid][y];
    potpart[(int64_t)blockIdx.x * NC + tid] = t;
  }
}

// per restart: potentials of its T candidates (fixed-order sum), first minimum wins; adopt it
__global__ void __launch_bounds__(256) kpp_select_kernel(const double* __restrict__ potpart, int nblk, int NC, int T,
                                                         int M, const int* __restrict__ cand,
                                                         const float* __restrict__ dmin, float* __restrict__ closest,
                                                         float* __restrict__ pot, int* __restrict__ chosen, int k,
                                                         int step) {
  const int init = blockIdx.x;
  __shared__ double sp[8][8];
  __shared__ int s_best;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (w < T) {
    double t = 0.0;
    for (int b = lane; b < nblk; b += 32) t += potpart[(int64_t)b * NC + init * T + w];
    t = warp_sum(t);
    if (lane == 0) sp[w][0] = t;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int best = 0;
    float bv = (float)sp[0][0];
    for (int t = 1; t < T; ++t) {
      const float v = (float)sp[t][0];
      if (v < bv) { bv = v; best = t; }
    }
    s_best = best;
    pot[init] = bv;
    chosen[init * k + step] = cand[init * T + best];
  }
  __syncthreads();
  const float* src = dmin + ((int64_t)init * T + s_best) * M;
  float* dst = closest + (int64_t)init * M;
  for (int j = threadIdx.x; j < M; j += 256) dst[j] = src[j];
}

__global__ void kpp_gather_kernel(const float* __restrict__ X, int64_t ld, int G, const int* __restrict__ chosen,
                                  float* __restrict__ centers) {
  const int c = blockIdx.x;  // over n_init * k
  const float* src = X + (int64_t)chosen[c] * ld;
  for (int g = threadIdx.x; g < G; g += blockDim.x) centers[(int64_t)c * G + g] = src[g];
}


// ---------------------------------------------------------------------------------------------
// f3 (k-selection statistics, cnmf.py:1096-1110): silhouette of the KMeans labels on the L2 spectra.
// One CTA per sample: per-cluster sums of its row of the distance matrix, intra = sum_own / (n_own - 1),
// inter = min over other clusters of sum_c / n_c, s = (inter - intra) / max(intra, inter)
// (sklearn/metrics/cluster/_unsupervised.py:_silhouette_reduce, silhouette_samples).
// ---------------------------------------------------------------------------------------------
constexpr int SIL_THREADS = 128;

__global__ void __launch_bounds__(SIL_THREADS) silhouette_kernel(const float* __restrict__ D, int64_t ldd, int M,
                                                                 const int* __restrict__ labels,
                                                                 const int* __restrict__ counts, int k,
                                                                 float* __restrict__ sil) {
  __shared__ float part[KA_C][SIL_THREADS];
  __shared__ double csum[KA_C];
  const int i = blockIdx.x, t = threadIdx.x;
  for (int c = 0; c < k; ++c) part[c][t] = 0.f;
  const float* drow = D + (int64_t)i * ldd;
  for (int j = t; j < M; j += SIL_THREADS) part[labels[j]][t] += drow[j];   // each thread owns its column
  __syncthreads();
  if (t < k) {
    double acc = 0.0;
    for (int q = 0; q < SIL_THREADS; ++q) acc += (double)part[t][q];        // fixed order
    csum[t] = acc;
  }
  __syncthreads();
  if (t == 0) {
    const int own = labels[i];
    const int n_own = counts[own];
    float out = 0.f;
    if (n_own > 1) {
      const float intra = (float)csum[own] / (float)(n_own - 1);
      float inter = INFINITY;
      for (int c = 0; c < k; ++c)
        if (c != own && counts[c] > 0) inter = fminf(inter, (float)csum[c] / (float)counts[c]);
      const float den = fmaxf(intra, inter);
      out = (inter - intra) / den;
      if (!(out == out) || isinf(out)) out = 0.f;   // nan_to_num
    }
    sil[i] = out;
  }
}

// ---------------------------------------------------------------------------------------------
// f2 (cnmf.py:56-126, efficient_ols_all_cols): the accumulations X^T X and X^T Y in fp64.
//   xty[k, j] = sum_i x[i, k] * y[i, j]  over the stored entries of column j of Y (CSC), plus
//   ysum[j], ysq[j] (for the global mean / variance of Y's columns) and xsum[k].
// One warp per target column; K <= 32 predictors, lane = predictor.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ols_xty_kernel(const int64_t* __restrict__ col_ptr, const int* __restrict__ row_idx,
                                                      const float* __restrict__ val, int n_targets,
                                                      const float* __restrict__ Xp, int K,
                                                      double* __restrict__ xty, double* __restrict__ ysum,
                                                      double* __restrict__ ysq) {
  const int j = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (j >= n_targets) return;
  double acc = 0.0, s1 = 0.0, s2 = 0.0;
  for (int64_t p = col_ptr[j]; p < col_ptr[j + 1]; ++p) {
    const double y = (double)val[p];
    const int i = row_idx[p];
    if (lane < K) acc = fma((double)Xp[(int64_t)i * K + lane], y, acc);
    s1 += y;
    s2 = fma(y, y, s2);
  }
  if (lane < K) xty[(int64_t)lane * n_targets + j] = acc;
  if (lane == 0) { ysum[j] = s1; ysq[j] = s2; }
}

}  // namespace cnmf

using namespace cnmf;

extern "C" {

int cnmf_b200_pairwise_dist(const float* X, int64_t ld, int32_t M, int32_t G, float* D, int64_t ldd, double* sqnorm,
                            void* stream) {
  CNMF_NVTX("cnmf_b200_pairwise_dist");
  CNMF_CHECK_ARG(X && D && sqnorm, "pairwise_dist: null pointer");
  CNMF_CHECK_ARG(M > 0 && G > 0 && ld >= G && ldd >= M, "pairwise_dist: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  row_sqnorm64_kernel<<<(unsigned)ceil_div(M, 8), 256, 0, st>>>(X, ld, M, G, sqnorm);
  CNMF_LAUNCH_CHECK();
  const int nt = (int)ceil_div(M, DT);
  pairwise_dist_kernel<<<(unsigned)(nt * (nt + 1) / 2), 256, 0, st>>>(X, ld, M, G, sqnorm, D, ldd, nt);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int cnmf_b200_topk_density(const float* D, int64_t ldd, int32_t M, int32_t n_neighbors, float* density, int32_t* nbr,
                           void* stream) {
  CNMF_NVTX("cnmf_b200_topk_density");
  CNMF_CHECK_ARG(D && density, "topk_density: null pointer");
  CNMF_CHECK_ARG(M > 0 && ldd >= M, "topk_density: bad shape");
  CNMF_CHECK_ARG(n_neighbors >= 1 && n_neighbors + 1 <= M, "topk_density: n_neighbors=%d needs 1 <= n and n+1 <= rows=%d",
                 n_neighbors, M);
  const size_t smem = (size_t)M * 4;
  if (smem > 200 * 1024) {
    set_error("topk_density: %d rows exceed the shared-memory row buffer (51200)", M);
    return CNMF_B200_ERR_UNSUPPORTED;
  }
  static unsigned long long configured = 0;
  if (smem > 48 * 1024 && needs_func_attr(&configured))
    CNMF_CUDA(cudaFuncSetAttribute(topk_density_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  topk_density_kernel<<<M, 256, smem, (cudaStream_t)stream>>>(D, ldd, M, n_neighbors + 1, n_neighbors, density, nbr);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int cnmf_b200_col_center(float* X, int64_t ld, int32_t M, int32_t G, float* mean, void* stream) {
  CNMF_NVTX("cnmf_b200_col_center");
  CNMF_CHECK_ARG(X && mean && M > 0 && G > 0 && ld >= G, "col_center: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  col_mean_seq_kernel<<<(unsigned)ceil_div(G, 64), 64, 0, st>>>(X, ld, M, G, mean);
  CNMF_LAUNCH_CHECK();
  sub_mean_kernel<<<(unsigned)ceil_div((int64_t)M * G, 256), 256, 0, st>>>(X, ld, M, G, mean);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int64_t cnmf_b200_kmeans_workspace_bytes(int32_t n_init, int32_t k, int32_t G) {
  return (int64_t)n_init * KU_SEG * k * ((int64_t)G * 4 + 4) + 256;
}

int cnmf_b200_kmeans_assign(const float* X, int64_t ld, int32_t M, int32_t G, const float* centers, int32_t n_init,
                            int32_t k, const uint8_t* active, int32_t* labels, int32_t* n_changed, void* stream) {
  CNMF_NVTX("cnmf_b200_kmeans_assign");
  CNMF_CHECK_ARG(X && centers && labels && n_changed, "kmeans_assign: null pointer");
  CNMF_CHECK_ARG(M > 0 && G > 0 && n_init > 0 && k >= 1 && k <= KA_C, "kmeans_assign: need 1 <= k <= %d (k=%d)", KA_C, k);
  dim3 grid((unsigned)ceil_div(M, KA_S), n_init);
  kmeans_assign_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, ld, M, G, centers, k, active, labels, n_changed);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int cnmf_b200_kmeans_update(const float* X, int64_t ld, int32_t M, int32_t G, const int32_t* labels, int32_t n_init,
                            int32_t k, const uint8_t* active, const float* centers_old, float* centers_new,
                            float* center_shift, int32_t* counts, void* workspace, void* stream) {
  CNMF_NVTX("cnmf_b200_kmeans_update");
  CNMF_CHECK_ARG(X && labels && centers_old && centers_new && center_shift && counts && workspace,
                 "kmeans_update: null pointer");
  CNMF_CHECK_ARG(M > 0 && G > 0 && n_init > 0 && k >= 1 && k <= KA_C, "kmeans_update: need 1 <= k <= %d", KA_C);
  cudaStream_t st = (cudaStream_t)stream;
  float* part = reinterpret_cast<float*>(workspace);
  int* cnt_part = reinterpret_cast<int*>(part + (int64_t)n_init * KU_SEG * k * G);
  dim3 grid((unsigned)ceil_div(G, KU_G), KU_SEG, n_init);
  kmeans_update_kernel<<<grid, KU_G, 0, st>>>(X, ld, M, G, labels, k, active, part, cnt_part);
  CNMF_LAUNCH_CHECK();
  kmeans_finalize_kernel<<<dim3(k, n_init), 128, 0, st>>>(part, cnt_part, G, k, active, centers_old, centers_new,
                                                          center_shift, counts);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int64_t cnmf_b200_kmeans_inertia_blocks(int32_t M) { return ceil_div(M, 8); }

int cnmf_b200_kmeans_inertia(const float* X, int64_t ld, int32_t M, int32_t G, const float* centers, int32_t n_init,
                             int32_t k, const int32_t* labels, double* partials, void* stream) {
  CNMF_NVTX("cnmf_b200_kmeans_inertia");
  CNMF_CHECK_ARG(X && centers && labels && partials, "kmeans_inertia: null pointer");
  dim3 grid((unsigned)ceil_div(M, 8), n_init);
  kmeans_inertia_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, ld, M, G, centers, k, labels, partials);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int cnmf_b200_cluster_median(const float* X, int64_t ld, int32_t G, const int32_t* order, const int32_t* offsets,
                             int32_t k, int32_t max_members, float* median, float* normalized, void* stream) {
  CNMF_NVTX("cnmf_b200_cluster_median");
  CNMF_CHECK_ARG(X && order && offsets && median, "cluster_median: null pointer");
  CNMF_CHECK_ARG(k >= 1 && G > 0 && max_members >= 1, "cluster_median: bad shape");
  const size_t smem = (size_t)max_members * (MD_G + 1) * 4;
  if (smem > 200 * 1024) {
    set_error("cluster_median: cluster of %d members exceeds the shared-memory tile (3000)", max_members);
    return CNMF_B200_ERR_UNSUPPORTED;
  }
  static unsigned long long configured = 0;
  if (smem > 48 * 1024 && needs_func_attr(&configured))
    CNMF_CUDA(cudaFuncSetAttribute(cluster_median_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  cudaStream_t st = (cudaStream_t)stream;
  cluster_median_kernel<<<dim3((unsigned)ceil_div(G, MD_G), k), 256, smem, st>>>(X, ld, G, order, offsets, median);
  CNMF_LAUNCH_CHECK();
  if (normalized != nullptr) {
    row_normalize_kernel<<<k, 256, 0, st>>>(median, G, normalized);
    CNMF_LAUNCH_CHECK();
  }
  return 0;
}

int cnmf_b200_l2_normalize(const float* S, int64_t lds, int32_t M, int32_t G, float* out, int64_t ldo, void* stream) {
  CNMF_NVTX("cnmf_b200_l2_normalize");
  CNMF_CHECK_ARG(S && out && M > 0 && G > 0 && lds >= G && ldo >= G, "l2_normalize: bad argument");
  l2_normalize_kernel<<<(unsigned)ceil_div(M, 128), 128, 0, (cudaStream_t)stream>>>(S, lds, M, G, out, ldo);
  CNMF_LAUNCH_CHECK();
  return 0;
}

static int64_t kpp_layout(int32_t M, int32_t n_init, int32_t T, int64_t* o_xx, int64_t* o_closest, int64_t* o_dmin,
                          int64_t* o_cand, int64_t* o_pot, int64_t* o_part) {
  const int64_t nblk = ceil_div(M, DT), NC = (int64_t)n_init * T;
  int64_t off = 0;
  *o_xx = off; off += align_up((int64_t)M * 8, 256);
  *o_closest = off; off += align_up((int64_t)n_init * M * 4, 256);
  *o_dmin = off; off += align_up(NC * M * 4, 256);
  *o_cand = off; off += align_up(NC * 4, 256);
  *o_pot = off; off += align_up((int64_t)n_init * 4, 256);
  *o_part = off; off += align_up(nblk * NC * 8, 256);
  return off;
}

int64_t cnmf_b200_kmeanspp_workspace_bytes(int32_t M, int32_t n_init, int32_t n_trials) {
  int64_t a, b, c, d, e, f;
  return kpp_layout(M, n_init, n_trials, &a, &b, &c, &d, &e, &f);
}

int cnmf_b200_kmeanspp(const float* Xc, int64_t ld, int32_t M, int32_t G, int32_t n_init, int32_t k, int32_t n_trials,
                       const int32_t* first_idx, const double* rnd, int32_t* chosen, float* centers, void* workspace,
                       void* stream) {
  CNMF_NVTX("cnmf_b200_kmeanspp");
  CNMF_CHECK_ARG(Xc && first_idx && chosen && centers && workspace && (rnd || k == 1), "kmeanspp: null pointer");
  CNMF_CHECK_ARG(M > 0 && G > 0 && k >= 1 && k <= M && ld >= G, "kmeanspp: bad shape");
  CNMF_CHECK_ARG(n_trials >= 1 && n_trials <= 8 && n_init >= 1 && n_init * n_trials <= KPP_NC,
                 "kmeanspp: n_init * n_trials must be <= %d and n_trials <= 8", KPP_NC);
  cudaStream_t st = (cudaStream_t)stream;
  int64_t o_xx, o_closest, o_dmin, o_cand, o_pot, o_part;
  kpp_layout(M, n_init, n_trials, &o_xx, &o_closest, &o_dmin, &o_cand, &o_pot, &o_part);
  char* ws = reinterpret_cast<char*>(workspace);
  double* xx = reinterpret_cast<double*>(ws + o_xx);
  float* closest = reinterpret_cast<float*>(ws + o_closest);
  float* dmin = reinterpret_cast<float*>(ws + o_dmin);
  int* cand = reinterpret_cast<int*>(ws + o_cand);
  float* pot = reinterpret_cast<float*>(ws + o_pot);
  double* part = reinterpret_cast<double*>(ws + o_part);
  const int nblk = (int)ceil_div(M, DT);
  row_sqnorm64_kernel<<<(unsigned)ceil_div(M, 8), 256, 0, st>>>(Xc, ld, M, G, xx);
  CNMF_LAUNCH_CHECK();
  // first centre: the host-drawn index; "one candidate per restart", nothing to take the minimum with
  kpp_dist_kernel<<<nblk, 256, 0, st>>>(Xc, ld, M, G, xx, first_idx, n_init, 1, nullptr, dmin, part);
  CNMF_LAUNCH_CHECK();
  kpp_select_kernel<<<n_init, 256, 0, st>>>(part, nblk, n_init, 1, M, first_idx, dmin, closest, pot, chosen, k, 0);
  CNMF_LAUNCH_CHECK();
  const int NC = n_init * n_trials;
  for (int c = 1; c < k; ++c) {
    kpp_pick_kernel<<<n_init, 32, 0, st>>>(closest, pot, rnd, M, n_trials, c - 1, k - 1, cand);
    CNMF_LAUNCH_CHECK();
    kpp_dist_kernel<<<nblk, 256, 0, st>>>(Xc, ld, M, G, xx, cand, NC, n_trials, closest, dmin, part);
    CNMF_LAUNCH_CHECK();
    kpp_select_kernel<<<n_init, 256, 0, st>>>(part, nblk, NC, n_trials, M, cand, dmin, closest, pot, chosen, k, c);
    CNMF_LAUNCH_CHECK();
  }
  kpp_gather_kernel<<<n_init * k, 128, 0, st>>>(Xc, ld, G, chosen, centers);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int cnmf_b200_silhouette(const float* D, int64_t ldd, int32_t M, const int32_t* labels, const int32_t* counts, int32_t k,
                         float* sil, void* stream) {
  CNMF_NVTX("cnmf_b200_silhouette");
  CNMF_CHECK_ARG(D && labels && counts && sil, "silhouette: null pointer");
  CNMF_CHECK_ARG(M > 0 && ldd >= M && k >= 1 && k <= KA_C, "silhouette: need 1 <= k <= %d", KA_C);
  silhouette_kernel<<<M, SIL_THREADS, 0, (cudaStream_t)stream>>>(D, ldd, M, labels, counts, k, sil);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int cnmf_b200_ols_xty(const int64_t* col_ptr, const int32_t* row_idx, const float* val, int32_t n_targets, const float* X,
                      int32_t K, double* xty, double* ysum, double* ysq, void* stream) {
  CNMF_NVTX("cnmf_b200_ols_xty");
  CNMF_CHECK_ARG(col_ptr && row_idx && val && X && xty && ysum && ysq, "ols_xty: null pointer");
  CNMF_CHECK_ARG(n_targets > 0 && K >= 1 && K <= 32, "ols_xty: need 1 <= K <= 32 predictors");
  ols_xty_kernel<<<(unsigned)ceil_div(n_targets, 8), 256, 0, (cudaStream_t)stream>>>(col_ptr, row_idx, val, n_targets, X, K,
                                                                                   xty, ysum, ysq);
  CNMF_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
