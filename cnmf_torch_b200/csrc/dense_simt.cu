// Dense Frobenius engine, CUDA-core (fp32 FFMA) kernels: the semantic baseline that the
// tcgen05 path is checked against, plus the small kernels both paths share (Gram, sum of
// squares, objective combine).  See include/cnmf_b200.h for the reference lines replaced.
#include <mutex>

#include "common.cuh"

namespace cnmf {

// ---------------------------------------------------------------------------------------------
// sum of squares (fp64 accumulate), two stages, deterministic
// ---------------------------------------------------------------------------------------------
constexpr int SUMSQ_BLOCKS = 1024;

__global__ void __launch_bounds__(256) sumsq_partial_kernel(const float* __restrict__ x, int64_t rows,
                                                            int64_t cols, int64_t ld,
                                                            double* __restrict__ partial) {
  double acc = 0.0;
  const int64_t total = rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / cols, c = i - r * cols;
    float v = x[r * ld + c];
    acc += (double)v * (double)v;
  }
  __shared__ double red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w];
    partial[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(32) sumsq_final_kernel(const double* __restrict__ partial, int n,
                                                         double* __restrict__ out) {
  // fixed-order tree: lane-strided partial sums then a shuffle tree
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += 32) acc += partial[i];
  acc = warp_sum(acc);
  if (threadIdx.x == 0) out[0] = acc;
}

// ---------------------------------------------------------------------------------------------
// Gram of every replicate block: gram[r] = F_r F_r^T
// ---------------------------------------------------------------------------------------------
constexpr int GRAM_TILE = 256;       // columns staged per step
constexpr int GRAM_THREADS = 128;

static inline int64_t gram_tiles(int64_t L) { return ceil_div(L, GRAM_TILE); }

// One CTA = one replicate x one 256-column tile.  The tile [KP x 256] is staged in shared memory
// with 128-bit loads; each thread owns one 4x4 block of the KP x KP output and a slice of the
// columns (16 FMA per 8 shared loads), slices are combined through shared memory in fixed order.
// Output: one partial per (replicate, tile), [R][ntile][KP*KP]; gram_final_kernel sums the tiles.
template <int KP>
__global__ void __launch_bounds__(GRAM_THREADS) gram_partial_kernel(const float* __restrict__ F,
                                                                    int64_t ld, int64_t L,
                                                                    float* __restrict__ partial,
                                                                    int ntile) {
  constexpr int NB1 = KP / 4, NB = NB1 * NB1;                 // 4x4 output blocks
  constexpr int NSL = GRAM_THREADS / NB > 0 ? GRAM_THREADS / NB : 1;  // column slices
  static_assert(NB <= GRAM_THREADS, "one thread per output block at least");
  __shared__ __align__(16) float tile[KP][GRAM_TILE];
  __shared__ float red[NSL][KP * KP + 1];
  const int r = blockIdx.y, tl = blockIdx.x;
  const float* Fr = F + (int64_t)r * KP * ld;
  const int64_t l0 = (int64_t)tl * GRAM_TILE;
  const int width = (int)min((int64_t)GRAM_TILE, L - l0);

  for (int i = threadIdx.x; i < KP * (GRAM_TILE / 4); i += GRAM_THREADS) {
    const int a = i / (GRAM_TILE / 4), c = (i - a * (GRAM_TILE / 4)) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* src = Fr + (int64_t)a * ld + l0 + c;
    if (c + 3 < width) v = *reinterpret_cast<const float4*>(src);
    else {
      if (c < width) v.x = src[0];
      if (c + 1 < width) v.y = src[1];
      if (c + 2 < width) v.z = src[2];
    }
    *reinterpret_cast<float4*>(&tile[a][c]) = v;
  }
  __syncthreads();

  const int sl = threadIdx.x % NSL, blk = threadIdx.x / NSL;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  if (blk < NB) {
    const int ba = (blk / NB1) * 4, bb = (blk % NB1) * 4;
    for (int c = sl; c < GRAM_TILE; c += NSL) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = tile[ba + i][c]; b[i] = tile[bb + i][c]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) red[sl][(ba + i) * KP + bb + j] = acc[i][j];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < KP * KP; o += GRAM_THREADS) {
    float t = 0.f;
#pragma unroll 4
    for (int s2 = 0; s2 < NSL; ++s2) t += red[s2][o];
    partial[((int64_t)r * ntile + tl) * (KP * KP) + o] = t;
  }
}

// Kp <= 12: the upper triangle of the Gram (Kp(Kp+1)/2 <= 78 values) fits in registers.  One CTA = one
// replicate x one column split; lane = column inside a 32-column group (coalesced loads, no shared-memory
// staging, no barrier in the main loop), warps stride over the groups, two groups in flight per warp.
// Far fewer and longer-lived CTAs than the tiled kernel: with thousands of tiny CTAs (R = 400) the tiled
// kernel spent its time in CTA scheduling, not in arithmetic.
constexpr int GW_SPLIT = 8192;  // columns per CTA (smaller splits measured slower: the per-CTA reduction dominates)

// Sum N per-lane values over the 32 lanes of a warp by recursive halving: at every step a lane keeps one half of
// its vector and hands the other half to its partner (lane ^ O), so the five steps cost N/2 + N/4 + ... ~ N
// shuffles instead of 5 N for N independent butterflies.  Afterwards lane L holds the totals of at most
// ceil(N/32) original indices; `put(i, v)` is called for each of them.  Fixed tree: deterministic.
template <int N, int O>
struct WarpHalving {
  static constexpr int H = (N + 1) / 2;
  // a[0..cnt) are the values of original indices base .. base+cnt-1 (a[j] = 0 for j >= cnt)
  template <typename Put>
  static __device__ __forceinline__ void run(const float (&a)[N], int lane, int base, int cnt, Put put) {
    const bool up = (lane & O) != 0;
    float k[H];
#pragma unroll
    for (int j = 0; j < H; ++j) {
      const float lo = a[j];
      const float hi = (j + H < N) ? a[j + H] : 0.f;
      k[j] = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, O);
    }
    const int cnt_lo = cnt < H ? cnt : H;
    WarpHalving<H, O / 2>::run(k, lane, up ? base + H : base, up ? cnt - cnt_lo : cnt_lo, put);
  }
};
template <int N>
struct WarpHalving<N, 0> {
  template <typename Put>
  static __device__ __forceinline__ void run(const float (&a)[N], int, int base, int cnt, Put put) {
#pragma unroll
    for (int j = 0; j < N; ++j)
      if (j < cnt) put(base + j, a[j]);
  }
};

template <int KP, int GW_THREADS>
__global__ void __launch_bounds__(GW_THREADS) gram_warp_kernel(const float* __restrict__ F, int64_t ld, int64_t L,
                                                               float* __restrict__ partial, int nsplit, int vec) {
  constexpr int NT = KP * (KP + 1) / 2;
  __shared__ float red[GW_THREADS / 32][NT];
  const int r = blockIdx.y, split = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* Fr = F + (int64_t)r * KP * ld;
  const int64_t l_begin = (int64_t)split * GW_SPLIT;
  const int64_t l_end = min(L, l_begin + (int64_t)GW_SPLIT);
  float acc[NT];
#pragma unroll
  for (int i = 0; i < NT; ++i) acc[i] = 0.f;
  // A lane takes four consecutive columns per step (one 16-byte load per row: 12 loads keep 6 KB per warp in flight;
  // the scalar form with two columns per lane kept 3 KB and ran at 2.7 TB/s, latency-bound).  vec = 0 (ld not a
  // multiple of 4 or F not 16-byte aligned: only a caller of the public entry point can do that): scalar loads.
  for (int64_t l0 = l_begin + warp * 128; l0 < l_end; l0 += (GW_THREADS / 32) * 128) {
    const int64_t c = l0 + lane * 4;
    float4 f[KP];
    if (vec && c + 3 < l_end) {
#pragma unroll
      for (int a = 0; a < KP; ++a) f[a] = *reinterpret_cast<const float4*>(Fr + (int64_t)a * ld + c);
    } else {
#pragma unroll
      for (int a = 0; a < KP; ++a) {
        const float* q = Fr + (int64_t)a * ld + c;
        f[a] = make_float4(c < l_end ? q[0] : 0.f, c + 1 < l_end ? q[1] : 0.f, c + 2 < l_end ? q[2] : 0.f,
                           c + 3 < l_end ? q[3] : 0.f);
      }
    }
    int t = 0;
#pragma unroll
    for (int a = 0; a < KP; ++a)
#pragma unroll
      for (int b = a; b < KP; ++b, ++t)
        acc[t] = fmaf(f[a].w, f[b].w, fmaf(f[a].z, f[b].z, fmaf(f[a].y, f[b].y, fmaf(f[a].x, f[b].x, acc[t]))));
  }
  WarpHalving<NT, 16>::run(acc, lane, 0, NT, [&](int i, float v) { red[warp][i] = v; });
  __syncthreads();
  for (int i = threadIdx.x; i < KP * KP; i += GW_THREADS) {
    const int a = i / KP, b = i - a * KP;
    const int lo = a < b ? a : b, hi = a < b ? b : a;
    const int t = lo * KP - lo * (lo - 1) / 2 + (hi - lo);  // index of (lo, hi) in the row-major upper triangle
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < GW_THREADS / 32; ++w) v += red[w][t];
    partial[((int64_t)r * nsplit + split) * (KP * KP) + i] = v;
  }
}

__global__ void gram_final_kernel(const float* __restrict__ partial, int nsplit, int np,
                                  float* __restrict__ gram, int64_t total) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // over R*np
  if (i >= total) return;
  int64_t r = i / np;
  int pair = (int)(i - r * np);
  // fp64 combine of the split partials, fixed order
  double t = 0.0;
  for (int s = 0; s < nsplit; ++s) t += (double)partial[(r * nsplit + s) * np + pair];
  gram[i] = (float)t;
}

// ---------------------------------------------------------------------------------------------
// C = A B^T, fp32, 128x128x16 tiles, 8x8 per thread
// ---------------------------------------------------------------------------------------------
constexpr int GM = 128, GN = 128, GK = 16;

__global__ void __launch_bounds__(256) gemm_nt_kernel(const float* __restrict__ A, int64_t lda,
                                                      int64_t M, const float* __restrict__ B,
                                                      int64_t ldb, int64_t N, int64_t Kd,
                                                      float* __restrict__ C, int64_t ldc) {
  __shared__ __align__(16) float As[GK][GM + 4];
  __shared__ __align__(16) float Bs[GK][GN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, 8x8 each
  const int64_t m0 = (int64_t)blockIdx.y * GM, n0 = (int64_t)blockIdx.x * GN;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  // each thread stages 2 float4 of A and 2 of B per k-step: row = idx/4, k-quad = idx%4
  for (int64_t k0 = 0; k0 < Kd; k0 += GK) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int idx = tid + h * 256;
      int row = idx >> 2, kq = (idx & 3) * 4;
      int64_t k = k0 + kq;
      float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
      if (m0 + row < M) {
        const float* pa = A + (m0 + row) * lda + k;
        if (k + 3 < Kd) va = *reinterpret_cast<const float4*>(pa);
        else {
          if (k < Kd) va.x = pa[0];
          if (k + 1 < Kd) va.y = pa[1];
          if (k + 2 < Kd) va.z = pa[2];
        }
      }
      if (n0 + row < N) {
        const float* pb = B + (n0 + row) * ldb + k;
        if (k + 3 < Kd) vb = *reinterpret_cast<const float4*>(pb);
        else {
          if (k < Kd) vb.x = pb[0];
          if (k + 1 < Kd) vb.y = pb[1];
          if (k + 2 < Kd) vb.z = pb[2];
        }
      }
      As[kq + 0][row] = va.x; As[kq + 1][row] = va.y; As[kq + 2][row] = va.z; As[kq + 3][row] = va.w;
      Bs[kq + 0][row] = vb.x; Bs[kq + 1][row] = vb.y; Bs[kq + 2][row] = vb.z; Bs[kq + 3][row] = vb.w;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < GK; ++kk) {
      float a[8], b[8];
      *reinterpret_cast<float4*>(&a[0]) = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      *reinterpret_cast<float4*>(&a[4]) = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
      *reinterpret_cast<float4*>(&b[0]) = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      *reinterpret_cast<float4*>(&b[4]) = *reinterpret_cast<const float4*>(&Bs[kk][64 + tx * 4]);
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int64_t m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= M) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      int64_t n = n0 + jh * 64 + tx * 4;
      float* pc = C + m * ldc + n;
      if (n + 3 < N) {
        *reinterpret_cast<float4*>(pc) =
            make_float4(acc[i][jh * 4 + 0], acc[i][jh * 4 + 1], acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (n + j < N) pc[j] = acc[i][jh * 4 + j];
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Row update (state and Pt both [R*KP x M]: consecutive threads = consecutive m => coalesced)
// ---------------------------------------------------------------------------------------------
constexpr int RU_THREADS = 128;
constexpr int RU_REPS = 4;  // replicates per CTA

template <int KP, int ALGO>
__global__ void __launch_bounds__(RU_THREADS) row_update_kernel(
    const float* __restrict__ Pt, int64_t ldp, float* __restrict__ state, int64_t lds,
    const float* __restrict__ gram, const uint8_t* __restrict__ active, int64_t M, int R, float l1,
    float l2, float eps, double* __restrict__ partials) {
  __shared__ float sgram[RU_REPS][KP * KP];
  __shared__ double red[RU_THREADS / 32][2];
  const int r0 = blockIdx.y * RU_REPS;
  for (int i = threadIdx.x; i < RU_REPS * KP * KP; i += RU_THREADS) {
    int rr = i / (KP * KP);
    sgram[rr][i - rr * KP * KP] = (r0 + rr < R) ? gram[(int64_t)(r0 + rr) * KP * KP + (i - rr * KP * KP)] : 0.f;
  }
  __syncthreads();
  const int64_t m = (int64_t)blockIdx.x * RU_THREADS + threadIdx.x;
  const bool in = m < M;
  for (int rr = 0; rr < RU_REPS; ++rr) {
    const int r = r0 + rr;
    if (r >= R) break;
    const bool act = active == nullptr || active[r] != 0;
    float cross = 0.f, ssum = 0.f;
    if (in) {
      float p[KP], s[KP];
      const int64_t base = (int64_t)r * KP;
#pragma unroll
      for (int k = 0; k < KP; ++k) {
        p[k] = Pt[(base + k) * ldp + m];
        s[k] = state[(base + k) * lds + m];
      }
      if (act) {
        update_row<KP, ALGO>(p, s, sgram[rr], l1, l2, eps);
#pragma unroll
        for (int k = 0; k < KP; ++k) state[(base + k) * lds + m] = s[k];
      }
#pragma unroll
      for (int k = 0; k < KP; ++k) {
        cross = fmaf(s[k], p[k], cross);
        ssum += s[k];
      }
    }
    double dc = warp_sum((double)cross), dsum = warp_sum((double)ssum);
    if ((threadIdx.x & 31) == 0) {
      red[threadIdx.x >> 5][0] = dc;
      red[threadIdx.x >> 5][1] = dsum;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double c = 0.0, s2 = 0.0;
      for (int w = 0; w < RU_THREADS / 32; ++w) {
        c += red[w][0];
        s2 += red[w][1];
      }
      partials[((int64_t)blockIdx.x * R + r) * 2 + 0] = c;
      partials[((int64_t)blockIdx.x * R + r) * 2 + 1] = s2;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// Objective combine: err2[r] = ||X||^2 - 2 cross_r + <gramW_r, gramH_r> + regularisation, fp64
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) frob_objective_kernel(
    double normX2, const double* __restrict__ partW, int nblkW, const double* __restrict__ partH,
    int nblkH, const float* __restrict__ gramW, const float* __restrict__ gramH, int R, int Kp,
    float l1_H, float l2_H, float l1_W, float l2_W, double* __restrict__ err2) {
  const int r = blockIdx.x;
  const int lane = threadIdx.x;
  double cross = 0.0, sumW = 0.0, sumH = 0.0;
  for (int b = lane; b < nblkW; b += 32) {
    cross += partW[((int64_t)b * R + r) * 2 + 0];
    sumW += partW[((int64_t)b * R + r) * 2 + 1];
  }
  for (int b = lane; b < nblkH; b += 32) sumH += partH[((int64_t)b * R + r) * 2 + 1];
  double tr = 0.0, trW = 0.0, trH = 0.0;
  const int np = Kp * Kp;
  for (int i = lane; i < np; i += 32) {
    double gw = gramW[(int64_t)r * np + i], gh = gramH[(int64_t)r * np + i];
    tr += gw * gh;
    if (i / Kp == i % Kp) {
      trW += gw;
      trH += gh;
    }
  }
  cross = warp_sum(cross); sumW = warp_sum(sumW); sumH = warp_sum(sumH);
  tr = warp_sum(tr); trW = warp_sum(trW); trH = warp_sum(trH);
  if (lane == 0) {
    double v = normX2 - 2.0 * cross + tr;
    v += 2.0 * ((double)l1_H * sumH + (double)l1_W * sumW) + (double)l2_H * trH + (double)l2_W * trW;
    err2[r] = v;
  }
}

// row-sharded variant: first reduce the per-block partials to [R x 3] sums (cross, sum W, sum H) ...
__global__ void __launch_bounds__(32) objective_sums_kernel(const double* __restrict__ partW, int nblkW,
                                                            const double* __restrict__ partH, int nblkH, int R,
                                                            double* __restrict__ sums) {
  const int r = blockIdx.x, lane = threadIdx.x;
  double cross = 0.0, sumW = 0.0, sumH = 0.0;
  for (int b = lane; b < nblkW; b += 32) {
    cross += partW[((int64_t)b * R + r) * 2 + 0];
    sumW += partW[((int64_t)b * R + r) * 2 + 1];
  }
  for (int b = lane; b < nblkH; b += 32) sumH += partH[((int64_t)b * R + r) * 2 + 1];
  cross = warp_sum(cross); sumW = warp_sum(sumW); sumH = warp_sum(sumH);
  if (lane == 0) { sums[r * 3 + 0] = cross; sums[r * 3 + 1] = sumW; sums[r * 3 + 2] = sumH; }
}

// ... then, after the caller has summed them over ranks, combine with the Grams
__global__ void __launch_bounds__(32) objective_combine_kernel(double normX2, const double* __restrict__ sums,
                                                               const float* __restrict__ gramW,
                                                               const float* __restrict__ gramH, int R, int Kp,
                                                               float l1_H, float l2_H, float l1_W, float l2_W,
                                                               double* __restrict__ err2) {
  const int r = blockIdx.x, lane = threadIdx.x;
  double tr = 0.0, trW = 0.0, trH = 0.0;
  const int np = Kp * Kp;
  for (int i = lane; i < np; i += 32) {
    const double gw = gramW[(int64_t)r * np + i], gh = gramH[(int64_t)r * np + i];
    tr += gw * gh;
    if (i / Kp == i % Kp) { trW += gw; trH += gh; }
  }
  tr = warp_sum(tr); trW = warp_sum(trW); trH = warp_sum(trH);
  if (lane == 0) {
    double v = normX2 - 2.0 * sums[r * 3 + 0] + tr;
    v += 2.0 * ((double)l1_H * sums[r * 3 + 2] + (double)l1_W * sums[r * 3 + 1]) + (double)l2_H * trH +
         (double)l2_W * trW;
    err2[r] = v;
  }
}

// ---------------------------------------------------------------------------------------------
// host-side launchers
// ---------------------------------------------------------------------------------------------
// number of partial Grams per replicate that launch_gram_partial leaves in the workspace
int gram_num_tiles(int64_t L, int Kp) { return Kp <= 12 ? (int)ceil_div(L, GW_SPLIT) : (int)gram_tiles(L); }

int launch_gram_partial(const float* F, int64_t ld, int R, int Kp, int64_t L, void* ws, cudaStream_t st) {
  CNMF_NVTX("frob: Gram partials");
  float* partial = reinterpret_cast<float*>(ws);
  if (Kp <= 12) {
    const int ns = (int)ceil_div(L, GW_SPLIT);
    dim3 grid(ns, R);
    // one CTA per (replicate, column split).  With 256 threads two CTAs fit an SM: beyond 296 CTAs a second, mostly
    // empty wave starts, and 128-thread CTAs (four per SM, twice the columns per warp) finish sooner
    // (R=400, Kp=12: 18.5 -> 14.4 us; R=200, Kp=8: 10.1 us with 256 threads, 11.4 with 128).
    const bool narrow = static_cast<int64_t>(ns) * R > 296;
    const int vec = (ld % 4 == 0 && aligned16(F)) ? 1 : 0;
    switch (Kp) {
      case 4:
        if (narrow) gram_warp_kernel<4, 128><<<grid, 128, 0, st>>>(F, ld, L, partial, ns, vec);
        else gram_warp_kernel<4, 256><<<grid, 256, 0, st>>>(F, ld, L, partial, ns, vec);
        break;
      case 8:
        if (narrow) gram_warp_kernel<8, 128><<<grid, 128, 0, st>>>(F, ld, L, partial, ns, vec);
        else gram_warp_kernel<8, 256><<<grid, 256, 0, st>>>(F, ld, L, partial, ns, vec);
        break;
      case 12:
        if (narrow) gram_warp_kernel<12, 128><<<grid, 128, 0, st>>>(F, ld, L, partial, ns, vec);
        else gram_warp_kernel<12, 256><<<grid, 256, 0, st>>>(F, ld, L, partial, ns, vec);
        break;
      default:
        set_error("gram: Kp=%d unsupported (multiples of 4 up to %d)", Kp, CNMF_B200_MAX_KP);
        return CNMF_B200_ERR_INVALID;
    }
    CNMF_LAUNCH_CHECK();
    return 0;
  }
  const int ntile = (int)gram_tiles(L);
  dim3 grid(ntile, R);
  CNMF_DISPATCH_KP(Kp, (gram_partial_kernel<KP><<<grid, GRAM_THREADS, 0, st>>>(F, ld, L, partial, ntile)));
  CNMF_LAUNCH_CHECK();
  return 0;
}

int launch_gram_final(const void* ws, int64_t L, int R, int Kp, float* gram, cudaStream_t st) {
  const int ntile = gram_num_tiles(L, Kp);
  const int64_t total = (int64_t)R * Kp * Kp;
  gram_final_kernel<<<(unsigned)ceil_div(total, 128), 128, 0, st>>>(reinterpret_cast<const float*>(ws), ntile, Kp * Kp,
                                                                    gram, total);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int launch_gram(const float* F, int64_t ld, int R, int Kp, int64_t L, float* gram, void* ws,
                cudaStream_t st) {
  CNMF_NVTX("frob: Gram");
  CNMF_TRY(launch_gram_partial(F, ld, R, Kp, L, ws, st));
  return launch_gram_final(ws, L, R, Kp, gram, st);
}

int launch_gemm_nt(const float* A, int64_t lda, int64_t M, const float* B, int64_t ldb, int64_t N,
                   int64_t Kd, float* C, int64_t ldc, cudaStream_t st) {
  dim3 grid((unsigned)ceil_div(N, GN), (unsigned)ceil_div(M, GM));
  gemm_nt_kernel<<<grid, 256, 0, st>>>(A, lda, M, B, ldb, N, Kd, C, ldc);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int launch_row_update(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram,
                      const uint8_t* active, int64_t M, int R, int Kp, int algo, float l1, float l2,
                      float eps, double* partials, cudaStream_t st) {
  dim3 grid((unsigned)ceil_div(M, RU_THREADS), (unsigned)ceil_div(R, RU_REPS));
  if (algo == CNMF_B200_ALGO_MU) {
    CNMF_DISPATCH_KP(Kp, (row_update_kernel<KP, CNMF_B200_ALGO_MU><<<grid, RU_THREADS, 0, st>>>(
                             Pt, ldp, state, lds, gram, active, M, R, l1, l2, eps, partials)));
  } else if (algo == CNMF_B200_ALGO_HALS) {
    CNMF_DISPATCH_KP(Kp, (row_update_kernel<KP, CNMF_B200_ALGO_HALS><<<grid, RU_THREADS, 0, st>>>(
                             Pt, ldp, state, lds, gram, active, M, R, l1, l2, eps, partials)));
  } else {
    set_error("unknown algo %d", algo);
    return CNMF_B200_ERR_INVALID;
  }
  CNMF_LAUNCH_CHECK();
  return 0;
}

int launch_frob_objective(double normX2, const double* partW, int nblkW, const double* partH,
                          int nblkH, const float* gramW, const float* gramH, int R, int Kp, float l1_H,
                          float l2_H, float l1_W, float l2_W, double* err2, cudaStream_t st) {
  frob_objective_kernel<<<R, 32, 0, st>>>(normX2, partW, nblkW, partH, nblkH, gramW, gramH, R, Kp,
                                          l1_H, l2_H, l1_W, l2_W, err2);
  CNMF_LAUNCH_CHECK();
  return 0;
}

__global__ void add_cross_kernel(double* __restrict__ part, const double* __restrict__ extra, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) part[i * 2] += extra[i * 2];   // entry 0 of every (block, replicate) pair is the cross term
}

int launch_add_cross(double* part, const double* extra, int64_t n, cudaStream_t st) {
  add_cross_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(part, extra, n);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int launch_objective_sums(const double* partW, int nblkW, const double* partH, int nblkH, int R, double* sums,
                          cudaStream_t st) {
  objective_sums_kernel<<<R, 32, 0, st>>>(partW, nblkW, partH, nblkH, R, sums);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int launch_objective_combine(double normX2, const double* sums, const float* gramW, const float* gramH, int R, int Kp,
                             float l1_H, float l2_H, float l1_W, float l2_W, double* err2, cudaStream_t st) {
  objective_combine_kernel<<<R, 32, 0, st>>>(normX2, sums, gramW, gramH, R, Kp, l1_H, l2_H, l1_W, l2_W, err2);
  CNMF_LAUNCH_CHECK();
  return 0;
}

}  // namespace cnmf

using namespace cnmf;

extern "C" {

int64_t cnmf_b200_sumsq_workspace_bytes(void) { return SUMSQ_BLOCKS * (int64_t)sizeof(double); }

int cnmf_b200_sumsq(const float* x, int64_t rows, int64_t cols, int64_t ld, double* out,
                    void* workspace, void* stream) {
  CNMF_NVTX("cnmf_b200_sumsq");
  CNMF_CHECK_ARG(x && out && workspace, "sumsq: null pointer");
  CNMF_CHECK_ARG(rows >= 0 && cols >= 0 && ld >= cols, "sumsq: bad shape %lld x %lld ld %lld",
                 (long long)rows, (long long)cols, (long long)ld);
  cudaStream_t st = (cudaStream_t)stream;
  double* partial = reinterpret_cast<double*>(workspace);
  sumsq_partial_kernel<<<SUMSQ_BLOCKS, 256, 0, st>>>(x, rows, cols, ld, partial);
  CNMF_LAUNCH_CHECK();
  sumsq_final_kernel<<<1, 32, 0, st>>>(partial, SUMSQ_BLOCKS, out);
  CNMF_LAUNCH_CHECK();
  return 0;
}

// Kp a multiple of 4 in (0, CNMF_B200_MAX_KP]; the tiled kernel (Kp >= 16) reads F with 16-byte loads
static int gram_check_layout(const char* who, const float* F, int64_t ld, int Kp) {
  CNMF_CHECK_ARG(Kp > 0 && Kp % 4 == 0 && Kp <= CNMF_B200_MAX_KP,
                 "%s: Kp=%d unsupported (multiples of 4 up to %d)", who, Kp, CNMF_B200_MAX_KP);
  CNMF_CHECK_ARG(Kp <= 12 || (ld % 4 == 0 && aligned16(F)),
                 "%s: for Kp > 12 ld must be a multiple of 4 and F 16-byte aligned", who);
  return 0;
}

int64_t cnmf_b200_gram_workspace_bytes(int32_t R, int32_t Kp, int64_t L) {
  return (int64_t)R * gram_tiles(L) * Kp * Kp * (int64_t)sizeof(float);
}

int cnmf_b200_gram(const float* F, int64_t ld, int32_t R, int32_t Kp, int64_t L, float* gram,
                   void* workspace, void* stream) {
  CNMF_NVTX("cnmf_b200_gram");
  CNMF_CHECK_ARG(F && gram && workspace, "gram: null pointer");
  CNMF_CHECK_ARG(R > 0 && L > 0 && ld >= L, "gram: bad shape R=%d L=%lld ld=%lld", R, (long long)L,
                 (long long)ld);
  CNMF_TRY(gram_check_layout("gram", F, ld, Kp));
  return launch_gram(F, ld, R, Kp, L, gram, workspace, (cudaStream_t)stream);
}

int cnmf_b200_gemm_nt(const float* A, int64_t lda, int64_t M, const float* B, int64_t ldb, int64_t N,
                      int64_t Kd, float* C, int64_t ldc, void* stream) {
  CNMF_NVTX("cnmf_b200_gemm_nt");
  CNMF_CHECK_ARG(A && B && C, "gemm_nt: null pointer");
  CNMF_CHECK_ARG(M > 0 && N > 0 && Kd > 0, "gemm_nt: empty shape");
  CNMF_CHECK_ARG(lda % 4 == 0 && ldb % 4 == 0 && ldc % 4 == 0 && aligned16(A) && aligned16(B) &&
                     aligned16(C),
                 "gemm_nt: leading dimensions must be multiples of 4 and pointers 16B aligned");
  CNMF_CHECK_ARG(lda >= Kd && ldb >= Kd && ldc >= N, "gemm_nt: leading dimension too small");
  return launch_gemm_nt(A, lda, M, B, ldb, N, Kd, C, ldc, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// Gram for callers that cannot allocate.
//
// Some hosts of this library have no way to hand it a workspace: an application that has given all
// of HBM to its own pool allocator (a later cudaMalloc would simply fail), or a latency-critical
// loop where the implicit device-wide synchronisation of cudaMalloc / cudaFree is not acceptable.
// For them the split partials of the Gram kernel live in the fixed array below instead of a
// caller-provided workspace.  The array is part of the library image: every GPU that loads
// libcnmf_b200.so carries these 4 MiB until the process exits.
// ---------------------------------------------------------------------------------------------
constexpr int64_t GRAM_STATIC_FLOATS = 1 << 20;  // 4 MiB: e.g. R=100, Kp=12 up to L = 18 432 columns
static __device__ __align__(256) float g_gram_static_scratch[GRAM_STATIC_FLOATS];

constexpr int GRAM_STATIC_MAX_DEVICES = 64;
static float* g_gram_static_ptr[GRAM_STATIC_MAX_DEVICES] = {nullptr};
static std::mutex g_gram_static_mutex;

// `g_gram_static_scratch` used from host code is only a handle for the runtime, not a device
// address; the address valid on the CURRENT device comes from cudaGetSymbolAddress.  Looked up once
// per device (first call on that device) and cached.
static int gram_static_scratch(float** out) {
  int dev = -1;
  CNMF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= GRAM_STATIC_MAX_DEVICES) {
    set_error("gram_static: device ordinal %d outside the supported range [0, %d)", dev,
              GRAM_STATIC_MAX_DEVICES);
    return CNMF_B200_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_gram_static_mutex);
  if (g_gram_static_ptr[dev] == nullptr) {
    float* p = nullptr;
    CNMF_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gram_static_scratch));
    if (p == nullptr) {
      set_error("gram_static: cudaGetSymbolAddress returned a null address on device %d", dev);
      return CNMF_B200_ERR_CUDA;
    }
    g_gram_static_ptr[dev] = p;
  }
  *out = g_gram_static_ptr[dev];
  return 0;
}

int64_t cnmf_b200_gram_static_max_floats(void) { return GRAM_STATIC_FLOATS; }

/* Same result as cnmf_b200_gram, but needs no workspace and allocates nothing (no cudaMalloc /
 * cudaFree, no library allocator): the split partials go to a 4 MiB array built into the library.
 *   - limit: R * ceil(L / 256) * Kp * Kp  must not exceed cnmf_b200_gram_static_max_floats();
 *     larger problems are refused with CNMF_B200_ERR_WORKSPACE (use cnmf_b200_gram with your own
 *     workspace instead);
 *   - the first call on each device looks up the array's device address; make that first call
 *     before starting any stream capture on that device;
 *   - all calls on a device share the one array and it keeps whatever the previous call left in
 *     it.  Nothing in it has to start at zero -- every partial that is read is written first by the
 *     same call -- but two calls that overlap on the GPU (different streams, or different host
 *     threads on the same device) race on it.  Keep them on one stream or serialise them. */
int cnmf_b200_gram_static(const float* F, int64_t ld, int32_t R, int32_t Kp, int64_t L, float* gram,
                          void* stream) {
  CNMF_NVTX("cnmf_b200_gram_static");
  CNMF_CHECK_ARG(F && gram, "gram_static: null pointer");
  CNMF_CHECK_ARG(R > 0 && L > 0 && ld >= L, "gram_static: bad shape R=%d L=%lld ld=%lld", R, (long long)L,
                 (long long)ld);
  CNMF_TRY(gram_check_layout("gram_static", F, ld, Kp));
  const int64_t need = (int64_t)R * gram_tiles(L) * Kp * Kp;  // one partial per 256-column tile
  if (need > GRAM_STATIC_FLOATS) {
    set_error("gram_static: needs %lld scratch floats, the built-in array holds %lld "
              "(R=%d, Kp=%d, L=%lld); use cnmf_b200_gram with a caller workspace",
              (long long)need, (long long)GRAM_STATIC_FLOATS, R, Kp, (long long)L);
    return CNMF_B200_ERR_WORKSPACE;
  }
  float* scratch = nullptr;
  CNMF_TRY(gram_static_scratch(&scratch));
  return launch_gram(F, ld, R, Kp, L, gram, scratch, (cudaStream_t)stream);
}

int64_t cnmf_b200_row_update_blocks(int64_t M) { return ceil_div(M, RU_THREADS); }

int cnmf_b200_row_update(const float* Pt, int64_t ldp, float* state, int64_t lds, const float* gram,
                         const uint8_t* active, int64_t M, int32_t R, int32_t Kp, int32_t algo, float l1,
                         float l2, float eps, double* partials, void* stream) {
  CNMF_NVTX("cnmf_b200_row_update");
  CNMF_CHECK_ARG(Pt && state && gram && partials, "row_update: null pointer");
  CNMF_CHECK_ARG(M > 0 && R > 0 && ldp >= M && lds >= M, "row_update: bad shape");
  return launch_row_update(Pt, ldp, state, lds, gram, active, M, R, Kp, algo, l1, l2, eps, partials,
                           (cudaStream_t)stream);
}

}  // extern "C"
