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
#include "kl_slab.cuh"
#include "kl_tables.cuh"

namespace cnmf {

// The KL path pads K to an EVEN Kp (the dense path needs multiples of 4 for its MMA shapes; here every padded
// component is a gathered value, a conversion and two FMAs per stored entry and replicate: K = 9 or 10 at Kp = 10
// instead of 12 is 17 % less of everything).
#define KL_DISPATCH_KP(kp, ...)                                                                       \
  switch (kp) {                                                                                       \
    case 2:  { constexpr int KP = 2;  __VA_ARGS__; } break;                                           \
    case 4:  { constexpr int KP = 4;  __VA_ARGS__; } break;                                           \
    case 6:  { constexpr int KP = 6;  __VA_ARGS__; } break;                                           \
    case 8:  { constexpr int KP = 8;  __VA_ARGS__; } break;                                           \
    case 10: { constexpr int KP = 10; __VA_ARGS__; } break;                                           \
    case 12: { constexpr int KP = 12; __VA_ARGS__; } break;                                           \
    case 14: { constexpr int KP = 14; __VA_ARGS__; } break;                                           \
    case 16: { constexpr int KP = 16; __VA_ARGS__; } break;                                           \
    case 18: { constexpr int KP = 18; __VA_ARGS__; } break;                                           \
    case 20: { constexpr int KP = 20; __VA_ARGS__; } break;                                           \
    case 22: { constexpr int KP = 22; __VA_ARGS__; } break;                                           \
    case 24: { constexpr int KP = 24; __VA_ARGS__; } break;                                           \
    case 26: { constexpr int KP = 26; __VA_ARGS__; } break;                                           \
    case 28: { constexpr int KP = 28; __VA_ARGS__; } break;                                           \
    case 30: { constexpr int KP = 30; __VA_ARGS__; } break;                                           \
    case 32: { constexpr int KP = 32; __VA_ARGS__; } break;                                           \
    default:                                                                                          \
      cnmf::set_error("kl: Kp=%d unsupported (even, <= 32)", (int)(kp));                              \
      return CNMF_B200_ERR_INVALID;                                                                   \
  }

constexpr int KL_THREADS = 128;  // (row, replicate) pairs per CTA, one per thread
constexpr int KL_SEG_MAX = 2048;  // engine path: rows longer than the segment length (256 .. 2048, kl_seg_len) are cut
constexpr int KL_SEG_MIN = 256;

// KP is even.  A multiple of 4: 16-byte (fp32) / 8-byte (fp16) accesses; otherwise 8-byte / 4-byte ones (a replicate's
// KP values start at r * KP elements, which is then only 8- / 4-byte aligned).
template <int KP>
__device__ __forceinline__ void load_vec(const float* __restrict__ p, float (&v)[KP]) {
  if constexpr (KP % 4 == 0) {
#pragma unroll
    for (int c = 0; c < KP; c += 4) {
      const float4 t = *reinterpret_cast<const float4*>(p + c);
      v[c] = t.x; v[c + 1] = t.y; v[c + 2] = t.z; v[c + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int c = 0; c < KP; c += 2) {
      const float2 t = *reinterpret_cast<const float2*>(p + c);
      v[c] = t.x; v[c + 1] = t.y;
    }
  }
}
template <int KP>
__device__ __forceinline__ void store_vec(float* __restrict__ p, const float (&v)[KP]) {
  if constexpr (KP % 4 == 0) {
#pragma unroll
    for (int c = 0; c < KP; c += 4) *reinterpret_cast<float4*>(p + c) = make_float4(v[c], v[c + 1], v[c + 2], v[c + 3]);
  } else {
#pragma unroll
    for (int c = 0; c < KP; c += 2) *reinterpret_cast<float2*>(p + c) = make_float2(v[c], v[c + 1]);
  }
}

// The fixed factor is gathered once per non-zero and replicate: that gather (nnz * R * Kp values per
// half-step) is what bounds the kernel, so the gathered copy is kept in fp16 (11-bit significand, the same
// operand precision as the tf32 copies of the dense path; values below 6e-5 become subnormal, an absolute
// error < 3e-8 that is irrelevant inside s = h.w).  Masters stay fp32; every half-step writes the fp16
// copy of the factor it updates next to the fp32 state.
template <int KP>
__device__ __forceinline__ void load_vec_h(const __half* __restrict__ p, float (&v)[KP]) {
  if constexpr (KP % 4 == 0) {
#pragma unroll
    for (int c = 0; c < KP; c += 4) {
      const uint2 t = *reinterpret_cast<const uint2*>(p + c);
      const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&t.x));
      const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&t.y));
      v[c] = a.x; v[c + 1] = a.y; v[c + 2] = b.x; v[c + 3] = b.y;
    }
  } else {
#pragma unroll
    for (int c = 0; c < KP; c += 2) {
      const float2 a = __half22float2(*reinterpret_cast<const __half2*>(p + c));
      v[c] = a.x; v[c + 1] = a.y;
    }
  }
}
template <int KP>
__device__ __forceinline__ void store_vec_h(__half* __restrict__ p, const float (&v)[KP]) {
  if constexpr (KP % 4 == 0) {
#pragma unroll
    for (int c = 0; c < KP; c += 4) {
      uint2 t;
      *reinterpret_cast<__half2*>(&t.x) = __floats2half2_rn(fminf(v[c], 65504.f), fminf(v[c + 1], 65504.f));
      *reinterpret_cast<__half2*>(&t.y) = __floats2half2_rn(fminf(v[c + 2], 65504.f), fminf(v[c + 3], 65504.f));
      *reinterpret_cast<uint2*>(p + c) = t;
    }
  } else {
#pragma unroll
    for (int c = 0; c < KP; c += 2)
      *reinterpret_cast<__half2*>(p + c) = __floats2half2_rn(fminf(v[c], 65504.f), fminf(v[c + 1], 65504.f));
  }
}

__global__ void to_half_kernel(const float* __restrict__ src, __half* __restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = __float2half_rn(fminf(src[i], 65504.f));
}

// One thread = one (row, replicate) pair, pairs numbered row-major with the replicate fastest and packed densely
// over the threads of the grid: a CTA is 128 consecutive pairs whatever R is (R = 13 replicates per GPU in an 8-way
// sharded run fills the lanes as well as R = 128; with one CTA per row only 13 of 128 lanes worked).  Lanes of a
// warp may sit in two or three consecutive rows: their index/value loads coalesce per row and the loop runs to the
// longest of those rows.
template <int KP>
__device__ __forceinline__ void kl_apply_update(float (&h)[KP], const float (&num)[KP], const float* __restrict__ fsum_r,
                                                float l1, float l2, float eps, float* __restrict__ sp,
                                                __half* __restrict__ hp) {
  float den[KP];
  load_vec<KP>(fsum_r, den);
#pragma unroll
  for (int k = 0; k < KP; ++k) {
    const float d = fmaf(l2, h[k], den[k]);
    const float n = l1 > 0.f ? fmaxf(num[k] - l1, 0.f) : num[k];
    const float rate = d < eps ? 0.f : __fdiv_rn(n, d);
    h[k] *= rate;
  }
  store_vec<KP>(sp, h);
  if (hp != nullptr) store_vec_h<KP>(hp, h);
}

// Segment table of one CSR matrix (engine path, built on the device by kl_build_segments).
//  * Rows longer than the segment length are cut into segments of that length.  A gene expressed in every cell is
//    a row of N entries in CSR(X^T); one thread walking it alone (20 000 entries ~ 1.6 ms) set the duration of the
//    whole gene-major half-step whatever R.  A thread then is a (segment, replicate) pair; segments of a cut row leave
//    their KP numerators in `part` and a finish pass adds them in segment order (deterministic) and applies the update.
//  * Column blocks (nb > 1, the gene-major half-step of a large matrix): a row is first cut at multiples of `block`
//    COLUMNS (cells), and segments are numbered block-major -- all rows' entries in cells [0, block) first, then
//    [block, 2 block), ... -- so that the threads resident at one time gather rows of the fixed factor from one or two
//    cell blocks (tens of MB, L2-resident) instead of from all cells (240 MB of fp16 H at 100 000 cells, R = 100:
//    every gather a DRAM access).  "Virtual row" v = b * M + row; with nb = 1 it is the row itself.

template <int KP, bool SEG>
__global__ void __launch_bounds__(KL_THREADS) kl_half_step_kernel(
    const int* __restrict__ row_ptr, const int* __restrict__ col_idx, const float* __restrict__ val, int64_t M,
    const __half* __restrict__ F, float* __restrict__ state, __half* __restrict__ state_h,
    const float* __restrict__ fsum, const uint8_t* __restrict__ active, int R, float l1, float l2, float eps,
    const KlSegs sg, const int gr) {
  const int64_t g = (int64_t)blockIdx.x * KL_THREADS + threadIdx.x;
  int64_t unit;                           // row, or segment
  int r;
  if (gr > 0) {
    // replicate-group-major numbering: all units for replicates [0, gr), then [gr, 2 gr), ...  The CTAs resident at
    // one time then gather the fixed factor of ONE group of gr replicates -- n_cols * gr * KP * 2 bytes, which the
    // L1 holds when gr is small -- instead of all R replicates' (L2 only).
    const int64_t units = SEG ? (int64_t)sg.vseg_ptr[M * sg.nb] : M;
    const int64_t per_group = units * gr;
    const int64_t grp = g / per_group;
    const int64_t within = g - grp * per_group;
    unit = within / gr;
    r = (int)(grp * gr + (within - unit * gr));
    if (r >= R) return;
  } else {
    unit = g / R;
    r = (int)(g - unit * R);
  }
  int64_t row = unit;
  int beg, end, slot = -1;
  if (SEG) {
    if (unit >= sg.vseg_ptr[M * sg.nb]) return;
    const int v = sg.seg_vrow[unit];
    row = sg.nb > 1 ? v % M : v;
    beg = sg.seg_beg[unit];
    const int v_end = sg.vend[v];
    end = beg + sg.seg_len < v_end ? beg + sg.seg_len : v_end;
    if (sg.nb > 1) {
      slot = (int)unit;   // every row is finished by kl_finish_all_kernel
    } else {
      const int s0 = sg.vseg_ptr[v];
      if (sg.vseg_ptr[v + 1] - s0 > 1) slot = sg.row_slot_ptr[v] + (int)(unit - s0);
    }
  } else {
    if (row >= M) return;
    beg = row_ptr[row];
    end = row_ptr[row + 1];
  }
  if (active != nullptr && active[r] == 0) return;
  const int64_t rk = (int64_t)R * KP;
  float* sp = state + row * rk + (int64_t)r * KP;
  float h[KP], num[KP];
  load_vec<KP>(sp, h);
#pragma unroll
  for (int k = 0; k < KP; ++k) num[k] = 0.f;
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
  if (SEG && slot >= 0) {   // a cut row: numerators of this segment only
    float* pp = sg.part + (int64_t)slot * rk + (int64_t)r * KP;
    store_vec<KP>(pp, num);
    return;
  }
  kl_apply_update<KP>(h, num, fsum + (int64_t)r * KP, l1, l2, eps, sp,
                      state_h != nullptr ? state_h + row * rk + (int64_t)r * KP : nullptr);
}

// rows that were cut into segments (nb == 1): numerators = sum of the segments' partials in segment order
template <int KP>
__global__ void __launch_bounds__(KL_THREADS) kl_finish_rows_kernel(float* __restrict__ state, __half* __restrict__ state_h,
                                                                     const float* __restrict__ fsum,
                                                                     const uint8_t* __restrict__ active, int R, float l1,
                                                                     float l2, float eps, const KlSegs sg) {
  const int64_t g = (int64_t)blockIdx.x * KL_THREADS + threadIdx.x;
  const int64_t mi = g / R;
  if (mi >= *sg.n_multi) return;
  const int r = (int)(g - mi * R);
  if (active != nullptr && active[r] == 0) return;
  const int64_t row = sg.multi_rows[mi];
  const int64_t rk = (int64_t)R * KP;
  const int nseg = sg.vseg_ptr[row + 1] - sg.vseg_ptr[row];
  const float* pp = sg.part + (int64_t)sg.row_slot_ptr[row] * rk + (int64_t)r * KP;
  float h[KP], num[KP], t[KP];
  load_vec<KP>(pp, num);
  for (int s2 = 1; s2 < nseg; ++s2) {
    load_vec<KP>(pp + (int64_t)s2 * rk, t);
#pragma unroll
    for (int k = 0; k < KP; ++k) num[k] += t[k];
  }
  float* sp = state + row * rk + (int64_t)r * KP;
  load_vec<KP>(sp, h);
  kl_apply_update<KP>(h, num, fsum + (int64_t)r * KP, l1, l2, eps, sp,
                      state_h != nullptr ? state_h + row * rk + (int64_t)r * KP : nullptr);
}

// column-blocked table (nb > 1): every row is finished here -- its segments of block 0, then block 1, ... in order
template <int KP>
__global__ void __launch_bounds__(KL_THREADS) kl_finish_all_kernel(int64_t M, float* __restrict__ state,
                                                                    __half* __restrict__ state_h,
                                                                    const float* __restrict__ fsum,
                                                                    const uint8_t* __restrict__ active, int R, float l1,
                                                                    float l2, float eps, const KlSegs sg) {
  const int64_t g = (int64_t)blockIdx.x * KL_THREADS + threadIdx.x;
  const int64_t row = g / R;
  if (row >= M) return;
  const int r = (int)(g - row * R);
  if (active != nullptr && active[r] == 0) return;
  const int64_t rk = (int64_t)R * KP;
  float h[KP], num[KP], t[KP];
#pragma unroll
  for (int k = 0; k < KP; ++k) num[k] = 0.f;
  for (int b = 0; b < sg.nb; ++b) {
    const int64_t v = (int64_t)b * M + row;
    for (int sgi = sg.vseg_ptr[v]; sgi < sg.vseg_ptr[v + 1]; ++sgi) {
      load_vec<KP>(sg.part + (int64_t)sgi * rk + (int64_t)r * KP, t);
#pragma unroll
      for (int k = 0; k < KP; ++k) num[k] += t[k];
    }
  }
  float* sp = state + row * rk + (int64_t)r * KP;
  load_vec<KP>(sp, h);
  kl_apply_update<KP>(h, num, fsum + (int64_t)r * KP, l1, l2, eps, sp,
                      state_h != nullptr ? state_h + row * rk + (int64_t)r * KP : nullptr);
}

// ---- segment table construction -------------------------------------------------------------------------------
// entries of row j with column in [b * block, (b+1) * block): vbeg / vend of virtual row b * M + j (columns sorted)
__global__ void __launch_bounds__(256) kl_block_bounds_kernel(const int* __restrict__ row_ptr,
                                                              const int* __restrict__ col_idx, int64_t M, int nb,
                                                              int block, int* __restrict__ vbeg, int* __restrict__ vend) {
  const int64_t g = (int64_t)blockIdx.x * 256 + threadIdx.x;   // (row, boundary b = 0 .. nb)
  const int64_t row = g / (nb + 1);
  if (row >= M) return;
  const int b = (int)(g - row * (nb + 1));
  int lo = row_ptr[row], hi = row_ptr[row + 1];
  const int64_t key = (int64_t)b * block;
  while (lo < hi) {   // first entry with column >= key
    const int mid = (lo + hi) >> 1;
    if (col_idx[mid] < key) lo = mid + 1; else hi = mid;
  }
  if (b < nb) vbeg[(int64_t)b * M + row] = lo;
  if (b > 0) vend[(int64_t)(b - 1) * M + row] = lo;
}

// one CTA of 1024 threads scans the V virtual rows (a few million at most).  min_one: a row without entries still
// gets a segment (nb == 1: its update, a zero rate, is applied by the half-step kernel itself)
__global__ void __launch_bounds__(1024) kl_seg_scan_kernel(const int* __restrict__ vbeg, const int* __restrict__ vend,
                                                           int64_t V, int seg_len, int min_one,
                                                           int* __restrict__ vseg_ptr, int* __restrict__ row_slot_ptr) {
  __shared__ int s_seg[1024], s_slot[1024];
  const int t = threadIdx.x;
  const int64_t chunk = (V + 1023) / 1024;
  const int64_t lo = t * chunk, hi = lo + chunk < V ? lo + chunk : V;
  int a = 0, b = 0;
  for (int64_t i = lo; i < hi; ++i) {
    const int len = vend[i] - vbeg[i];
    const int c = len > seg_len ? (len + seg_len - 1) / seg_len : (len > 0 || min_one ? 1 : 0);
    a += c;
    b += c > 1 ? c : 0;
  }
  s_seg[t] = a; s_slot[t] = b;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {   // Hillis-Steele inclusive scan
    const int va = t >= off ? s_seg[t - off] : 0, vb = t >= off ? s_slot[t - off] : 0;
    __syncthreads();
    s_seg[t] += va; s_slot[t] += vb;
    __syncthreads();
  }
  a = s_seg[t] - a; b = s_slot[t] - b;       // exclusive prefix of this thread's chunk
  for (int64_t i = lo; i < hi; ++i) {
    vseg_ptr[i] = a;
    if (row_slot_ptr != nullptr) row_slot_ptr[i] = b;
    const int len = vend[i] - vbeg[i];
    const int c = len > seg_len ? (len + seg_len - 1) / seg_len : (len > 0 || min_one ? 1 : 0);
    a += c;
    b += c > 1 ? c : 0;
  }
  if (t == 1023) {
    vseg_ptr[V] = s_seg[1023];
    if (row_slot_ptr != nullptr) row_slot_ptr[V] = s_slot[1023];
  }
}

__global__ void __launch_bounds__(256) kl_seg_fill_kernel(const int* __restrict__ vbeg, int64_t V, int seg_len,
                                                          const int* __restrict__ vseg_ptr, int* __restrict__ seg_vrow,
                                                          int* __restrict__ seg_beg, int* __restrict__ multi_rows,
                                                          int* __restrict__ n_multi) {
  const int64_t v = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (v >= V) return;
  const int s0 = vseg_ptr[v], n = vseg_ptr[v + 1] - s0, beg = vbeg[v];
  for (int i = 0; i < n; ++i) { seg_vrow[s0 + i] = (int)v; seg_beg[s0 + i] = beg + i * seg_len; }
  if (multi_rows != nullptr && n > 1) multi_rows[atomicAdd(n_multi, 1)] = (int)v;   // list order is irrelevant
}

// sum over non-zeros of x * log(max(s, eps)) per replicate; fp64 partials per block of KLO_ROWS rows [nblk x R]
constexpr int KLO_ROWS = 8;

template <int KP>
__global__ void __launch_bounds__(KL_THREADS) kl_xlogs_kernel(const int* __restrict__ row_ptr,
                                                              const int* __restrict__ col_idx,
                                                              const float* __restrict__ val, int64_t M,
                                                              const float* __restrict__ Hr, const float* __restrict__ Wt,
                                                              int R, float eps, double* __restrict__ partial) {
  // one thread = (block of KLO_ROWS rows, replicate), pairs packed densely over the grid like the half-step
  const int64_t g = (int64_t)blockIdx.x * KL_THREADS + threadIdx.x;
  const int64_t blk = g / R;
  const int64_t row0 = blk * KLO_ROWS;
  if (row0 >= M) return;
  const int r = (int)(g - blk * R);
  const int64_t rk = (int64_t)R * KP;
  double acc = 0.0;
  for (int rr = 0; rr < KLO_ROWS; ++rr) {
    const int64_t row = row0 + rr;
    if (row >= M) break;
    float h[KP];
    load_vec<KP>(Hr + row * rk + (int64_t)r * KP, h);
    const int beg = row_ptr[row], end = row_ptr[row + 1];
    float racc = 0.f;
    int p = beg;
    for (; p + 2 <= end; p += 2) {   // two entries in flight
      const int j0 = col_idx[p], j1 = col_idx[p + 1];
      const float x0 = val[p], x1 = val[p + 1];
      float w0[KP], w1[KP];
      load_vec<KP>(Wt + (int64_t)j0 * rk + (int64_t)r * KP, w0);
      load_vec<KP>(Wt + (int64_t)j1 * rk + (int64_t)r * KP, w1);
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int k = 0; k < KP; ++k) { s0 = fmaf(h[k], w0[k], s0); s1 = fmaf(h[k], w1[k], s1); }
      if (x0 > 0.f) racc = fmaf(x0, logf(fmaxf(s0, eps)), racc);
      if (x1 > 0.f) racc = fmaf(x1, logf(fmaxf(s1, eps)), racc);
    }
    if (p < end) {
      const int j = col_idx[p];
      const float x = val[p];
      float w[KP];
      load_vec<KP>(Wt + (int64_t)j * rk + (int64_t)r * KP, w);
      float s0 = 0.f;
#pragma unroll
      for (int k = 0; k < KP; ++k) s0 = fmaf(h[k], w[k], s0);
      if (x > 0.f) racc = fmaf(x, logf(fmaxf(s0, eps)), racc);
    }
    acc += (double)racc;
  }
  partial[blk * R + r] = acc;
}

// column sums of a [L x C] row-major array: per-CTA partials over 256-row slabs, then a fixed-order sum
constexpr int CS_ROWS = 256;

template <bool SQ>
__global__ void __launch_bounds__(128) colsum_partial_kernel(const float* __restrict__ A, int64_t L, int64_t C,
                                                             float* __restrict__ partial) {
  const int64_t c = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (c >= C) return;
  const int64_t l0 = (int64_t)blockIdx.y * CS_ROWS, l1 = min(L, l0 + (int64_t)CS_ROWS);
  float acc = 0.f;
  for (int64_t l = l0; l < l1; ++l) {
    const float a = A[l * C + c];
    acc = SQ ? fmaf(a, a, acc) : acc + a;   // SQ: column sums of squares (the L2 term of the objective)
  }
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
__global__ void __launch_bounds__(256) kl_objective_kernel(double xlogx, double sumX, const double* __restrict__ partial,
                                                           int nblk, const double* __restrict__ hsum,
                                                           const double* __restrict__ wsum, int R, int Kp, float l1_H,
                                                           float l1_W, const double* __restrict__ hsq,
                                                           const double* __restrict__ wsq, float l2_H, float l2_W,
                                                           double* __restrict__ err2) {
  // one CTA per replicate; fixed assignment of partials to threads and a fixed reduction tree: deterministic
  __shared__ double sx[8];
  const int r = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double xls = 0.0;
  for (int b = threadIdx.x; b < nblk; b += 256) xls += partial[(int64_t)b * R + r];
  xls = warp_sum(xls);
  if (lane == 0) sx[warp] = xls;
  __syncthreads();
  if (warp != 0) return;
  xls = 0.0;
  for (int w = 0; w < 8; ++w) xls += sx[w];
  double ss = 0.0, sh = 0.0, sw = 0.0;
  for (int k = lane; k < Kp; k += 32) {
    const double a = hsum[(int64_t)r * Kp + k], b = wsum[(int64_t)r * Kp + k];
    ss += a * b;
    sh += a;
    sw += b;
  }
  ss = warp_sum(ss); sh = warp_sum(sh); sw = warp_sum(sw);
  double q = 0.0;   // l2_H ||H_r||^2 + l2_W ||W_r||^2 (= 2 x the 0.5 l2 ||.||^2 terms)
  if (hsq != nullptr) {
    for (int k = lane; k < Kp; k += 32)
      q += (double)l2_H * hsq[(int64_t)r * Kp + k] + (double)l2_W * wsq[(int64_t)r * Kp + k];
    q = warp_sum(q);
  }
  if (lane == 0) err2[r] = 2.0 * (xlogx - xls - sumX + ss + (double)l1_H * sh + (double)l1_W * sw) + q;
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
                         cudaStream_t st, bool squares = false) {
  CNMF_NVTX("kl: column sums");
  const int nslab = (int)ceil_div(L, CS_ROWS);
  dim3 grid((unsigned)ceil_div(C, 128), nslab);
  if (squares) colsum_partial_kernel<true><<<grid, 128, 0, st>>>(A, L, C, partial);
  else colsum_partial_kernel<false><<<grid, 128, 0, st>>>(A, L, C, partial);
  CNMF_LAUNCH_CHECK();
  colsum_final_kernel<<<(unsigned)ceil_div(C, 128), 128, 0, st>>>(partial, nslab, C, out, out64);
  CNMF_LAUNCH_CHECK();
  return 0;
}

// gather path: its fp16 copies saturate at 65504; an entry beyond that raises the range flag (checked on the masters
// at the start and at the end of every cnmf_b200_kl_iterate call) so that the host can refuse instead of going on
__global__ void __launch_bounds__(256) kl_range_check_kernel(const float* __restrict__ a, int64_t n, int* __restrict__ flag) {
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) bad |= !(a[i] <= 65504.f);
  if (bad) atomicOr(flag, 1);
}
static int launch_range_check(const float* a, int64_t n, int* flag, cudaStream_t st) {
  int64_t blocks = ceil_div(n, 256 * 8);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  kl_range_check_kernel<<<(unsigned)blocks, 256, 0, st>>>(a, n, flag);
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

// sg == nullptr: one thread per (row, replicate) (the standalone entry point, which has no workspace for a
// segment table); else one thread per (segment, replicate) + the pass that finishes the cut rows (nb == 1) or
// every row (column-blocked table).
static int launch_kl_half(const int* rp, const int* ci, const float* v, int64_t M, const __half* F, float* state,
                          __half* state_h, const float* fsum, const uint8_t* active, int R, int Kp, float l1, float l2,
                          float eps, cudaStream_t st, const KlSegs* sg = nullptr, int64_t nseg_max = 0,
                          int64_t multi_max = 0, int gr = 0) {
  CNMF_NVTX("kl: gather half-step");
  if (gr >= R) gr = 0;
  const int64_t Rpad = gr > 0 ? ceil_div(R, gr) * gr : R;   // group-major numbering: the last group is padded
  const int64_t blocks = ceil_div((sg ? nseg_max : M) * Rpad, KL_THREADS);
  CNMF_CHECK_ARG(blocks < (1ll << 31), "kl: rows * replicates too large for one launch");
  if (sg == nullptr) {
    KlSegs none{};
    KL_DISPATCH_KP(Kp, (kl_half_step_kernel<KP, false><<<(unsigned)blocks, KL_THREADS, 0, st>>>(
                             rp, ci, v, M, F, state, state_h, fsum, active, R, l1, l2, eps, none, gr)));
    CNMF_LAUNCH_CHECK();
    return 0;
  }
  KL_DISPATCH_KP(Kp, (kl_half_step_kernel<KP, true><<<(unsigned)blocks, KL_THREADS, 0, st>>>(
                           rp, ci, v, M, F, state, state_h, fsum, active, R, l1, l2, eps, *sg, gr)));
  CNMF_LAUNCH_CHECK();
  if (sg->nb > 1) {
    const int64_t fb = ceil_div(M * R, KL_THREADS);
    KL_DISPATCH_KP(Kp, (kl_finish_all_kernel<KP><<<(unsigned)fb, KL_THREADS, 0, st>>>(M, state, state_h, fsum, active,
                                                                                         R, l1, l2, eps, *sg)));
    CNMF_LAUNCH_CHECK();
  } else if (multi_max > 0) {
    const int64_t fb = ceil_div(multi_max * R, KL_THREADS);
    KL_DISPATCH_KP(Kp, (kl_finish_rows_kernel<KP><<<(unsigned)fb, KL_THREADS, 0, st>>>(state, state_h, fsum, active, R,
                                                                                          l1, l2, eps, *sg)));
    CNMF_LAUNCH_CHECK();
  }
  return 0;
}

// Segment length for a matrix of M rows and nnz entries at R replicates: long enough that whole rows stay whole
// where there is parallelism anyway, short enough that (segments x R) threads fill the GPU when R is small (an
// 8-way sharded run has R = 12 or 13 per K) -- a thread walks its entries one gather latency after the other.
static int kl_seg_len(int64_t M, int64_t nnz, int R) {
  const int64_t want = 148ll * 2048;                     // resident threads of the GPU
  const int64_t spare = want - M * R;
  int64_t s = spare > 0 ? (nnz * R) / spare : KL_SEG_MAX;
  int len = KL_SEG_MIN;
  while (len < KL_SEG_MAX && len * 2 <= s) len *= 2;
  return len;
}
// Column block (in columns) of a matrix whose half-step gathers rows of a fixed factor with n_cols rows of R*Kp fp16
// values: 0 (no blocking) while that factor fits the L2 budget, else the largest power of two such that two blocks
// of it do (the resident threads straddle at most two blocks).
constexpr int64_t KL_L2_BUDGET = 40ll << 20;
static int kl_col_block(int64_t n_cols, int R, int Kp) {
  static const int mode = [] { const char* e = getenv("CNMF_B200_KL_BLOCK"); return e ? atoi(e) : -1; }();  // dev knob
  if (mode == 0) return 0;
  const int64_t row_bytes = (int64_t)R * Kp * 2;
  if (mode > 0) return mode < n_cols ? mode : 0;
  if (n_cols * row_bytes <= KL_L2_BUDGET) return 0;
  int64_t b = 1024;
  while (2 * (b * 2) * row_bytes <= KL_L2_BUDGET) b *= 2;
  return b < n_cols ? (int)b : 0;
}
// Replicates per group of the cell-major half-step (gathers rows of W): as many as keep one group's fp16 slice of W,
// n_genes * gr * Kp * 2 bytes, within an L1 budget.  0 = no grouping.  (CNMF_B200_KL_GR overrides: dev knob.)
static int kl_rep_group(int64_t n_genes, int Kp) {
  static const int forced = [] { const char* e = getenv("CNMF_B200_KL_GR"); return e ? atoi(e) : -1; }();
  if (forced >= 0) return forced;
  return 0;
}
// bounds of the table
static int64_t kl_multi_max(int64_t nnz, int seg) { return nnz / seg; }            // rows longer than seg
static int64_t kl_slot_max(int64_t V, int64_t nnz, int seg, int nb) {              // partial slots
  return nb > 1 ? kl_nseg_max(V, nnz, seg) : 2 * (nnz / seg) + 2;
}


static int64_t kl_seg_layout(int64_t M, int64_t n_cols, int64_t nnz, int R, int Kp, bool allow_blocks, int64_t off,
                             KlSegLayout* o) {
  o->seg = kl_seg_len(M, nnz, R);
  o->block = allow_blocks ? kl_col_block(n_cols, R, Kp) : 0;
  o->nb = o->block > 0 ? (int)ceil_div(n_cols, o->block) : 1;
  o->V = M * o->nb;
  o->off_vsp = off; off += align_up((o->V + 1) * 4, 256);
  o->off_rslot = off; off += align_up((o->V + 1) * 4, 256);
  o->off_svrow = off; off += align_up(kl_nseg_max(o->V, nnz, o->seg) * 4, 256);
  o->off_sbeg = off; off += align_up(kl_nseg_max(o->V, nnz, o->seg) * 4, 256);
  o->off_multi = off; off += align_up((kl_multi_max(nnz, o->seg) + 1) * 4, 256);
  o->off_nmulti = off; off += 256;
  o->off_vbeg = off; if (o->nb > 1) off += align_up(o->V * 4, 256);
  o->off_vend = off; if (o->nb > 1) off += align_up(o->V * 4, 256);
  o->all_partial = false;
  return off;
}

int64_t kl_seg_layout_explicit(int64_t M, int64_t n_cols, int64_t nnz, int seg, int block, int64_t off, KlSegLayout* o) {
  o->seg = seg;
  o->block = block;
  o->nb = block > 0 ? (int)ceil_div(n_cols, block) : 1;
  o->V = M * o->nb;
  o->all_partial = true;
  o->off_vsp = off; off += align_up((o->V + 1) * 4, 256);
  o->off_rslot = off;
  o->off_svrow = off; off += align_up(kl_nseg_max(o->V, nnz, o->seg) * 4, 256);
  o->off_sbeg = off; off += align_up(kl_nseg_max(o->V, nnz, o->seg) * 4, 256);
  o->off_multi = off;
  o->off_nmulti = off; off += 256;
  o->off_vbeg = off; if (o->nb > 1) off += align_up(o->V * 4, 256);
  o->off_vend = off; if (o->nb > 1) off += align_up(o->V * 4, 256);
  return off;
}

int kl_build_segments(const int* row_ptr, const int* col_idx, int64_t M, char* ws, const KlSegLayout& o, float* part,
                      KlSegs* out, cudaStream_t st, bool launch) {
  int* vsp = reinterpret_cast<int*>(ws + o.off_vsp);
  int* rslot = reinterpret_cast<int*>(ws + o.off_rslot);
  int* svrow = reinterpret_cast<int*>(ws + o.off_svrow);
  int* sbeg = reinterpret_cast<int*>(ws + o.off_sbeg);
  int* multi = reinterpret_cast<int*>(ws + o.off_multi);
  int* nmulti = reinterpret_cast<int*>(ws + o.off_nmulti);
  const int* vbeg = row_ptr;
  const int* vend = row_ptr + 1;
  if (o.nb > 1) {
    int* vb = reinterpret_cast<int*>(ws + o.off_vbeg);
    int* ve = reinterpret_cast<int*>(ws + o.off_vend);
    if (launch) {
      kl_block_bounds_kernel<<<(unsigned)ceil_div(M * (o.nb + 1), 256), 256, 0, st>>>(row_ptr, col_idx, M, o.nb, o.block,
                                                                                       vb, ve);
      CNMF_LAUNCH_CHECK();
    }
    vbeg = vb; vend = ve;
  }
  if (launch) {
    const bool partial = o.nb > 1 || o.all_partial;   // one partial slot per segment, empty virtual rows own none
    CNMF_CUDA(cudaMemsetAsync(nmulti, 0, 4, st));
    kl_seg_scan_kernel<<<1, 1024, 0, st>>>(vbeg, vend, o.V, o.seg, partial ? 0 : 1, vsp, partial ? nullptr : rslot);
    CNMF_LAUNCH_CHECK();
    kl_seg_fill_kernel<<<(unsigned)ceil_div(o.V, 256), 256, 0, st>>>(vbeg, o.V, o.seg, vsp, svrow, sbeg,
                                                                     partial ? nullptr : multi, nmulti);
    CNMF_LAUNCH_CHECK();
  }
  out->vseg_ptr = vsp; out->row_slot_ptr = rslot; out->seg_vrow = svrow; out->seg_beg = sbeg; out->vend = vend;
  out->multi_rows = multi; out->n_multi = nmulti; out->part = part; out->seg_len = o.seg; out->nb = o.nb;
  return 0;
}

struct KlLayout {
  int64_t off_wsum, off_hsum, off_wsum64, off_hsum64, off_wsq64, off_hsq64, off_cpart, off_opart, off_wh, off_hh, off_part, total;
  int64_t off_flag, off_wcumX, off_wcumXt;   // slab path: range flag, static work split of the two matrices
  int64_t off_pairsX, off_pairsXt;           // slab path: (column, value) pairs per segment, padded to whole steps
  KlSegLayout segX, segXt;  // CSR(X): never column-blocked (gathers W, G rows); CSR(X^T): blocked over cells when H is large
  int nblk_obj;
  bool slab;
};

static bool kl_use_slab(const cnmf_b200_kl_problem* p) { return p->impl == CNMF_B200_KL_IMPL_SLAB; }

static KlLayout kl_layout(const cnmf_b200_kl_problem* p) {
  KlLayout l{};
  l.slab = kl_use_slab(p);
  const int64_t C = (int64_t)p->R * p->Kp;
  const int64_t Lmax = p->n_cells > p->n_genes ? p->n_cells : p->n_genes;
  int64_t off = 0;
  l.off_wsum = off; off += align_up(C * 4, 256);
  l.off_hsum = off; off += align_up(C * 4, 256);
  l.off_wsum64 = off; off += align_up(C * 8, 256);
  l.off_hsum64 = off; off += align_up(C * 8, 256);
  l.off_wsq64 = off; off += align_up(C * 8, 256);
  l.off_hsq64 = off; off += align_up(C * 8, 256);
  l.off_cpart = off; off += align_up(ceil_div(Lmax, CS_ROWS) * C * 4, 256);
  l.off_flag = off; off += 256;
  if (l.slab) {
    // fp16 copies in the line layout [rows][groups] x 128 bytes; objective partials [SMs][R]
    const int64_t C16 = (int64_t)ks_num_groups(p->R, p->Kp) * KS_LINE;   // one 128-byte line per (row, replicate group)
    l.nblk_obj = ks_sm_count();
    l.off_opart = off; off += align_up((int64_t)l.nblk_obj * p->R * 8, 256);
    l.off_wh = off; off += align_up(p->n_genes * C16, 256);
    l.off_hh = off; off += align_up(p->n_cells * C16, 256);
    off = kl_seg_layout_explicit(p->n_cells, p->n_genes, p->nnz, ks_seg_len(p->Kp), ks_block_cols(p->n_genes), off, &l.segX);
    off = kl_seg_layout_explicit(p->n_genes, p->n_cells, p->nnz, ks_seg_len(p->Kp), ks_block_cols(p->n_cells), off, &l.segXt);
    const int64_t sx = kl_nseg_max(l.segX.V, p->nnz, l.segX.seg), sxt = kl_nseg_max(l.segXt.V, p->nnz, l.segXt.seg);
    l.off_wcumX = off; off += align_up((sx + 1) * 4, 256);
    l.off_wcumXt = off; off += align_up((sxt + 1) * 4, 256);
    l.off_pairsX = off; off += align_up(ks_pairs_max(p->nnz, sx, p->Kp) * 8, 256);
    l.off_pairsXt = off; off += align_up(ks_pairs_max(p->nnz, sxt, p->Kp) * 8, 256);
    l.off_part = off; off += align_up((sx > sxt ? sx : sxt) * C * 4, 256);
    l.total = off;
    return l;
  }
  l.nblk_obj = (int)ceil_div(p->n_cells, KLO_ROWS);
  l.off_opart = off; off += align_up((int64_t)l.nblk_obj * p->R * 8, 256);
  l.off_wh = off; off += align_up(p->n_genes * C * 2, 256);   // fp16 gather copies of Wt and Hr
  l.off_hh = off; off += align_up(p->n_cells * C * 2, 256);
  // segment tables of CSR(X) and CSR(X^T) and the partial numerators of cut rows (one buffer: the two half-steps
  // of an iteration run one after the other)
  off = kl_seg_layout(p->n_cells, p->n_genes, p->nnz, p->R, p->Kp, false, off, &l.segX);
  off = kl_seg_layout(p->n_genes, p->n_cells, p->nnz, p->R, p->Kp, true, off, &l.segXt);
  const int64_t sx = kl_slot_max(l.segX.V, p->nnz, l.segX.seg, l.segX.nb);
  const int64_t sxt = kl_slot_max(l.segXt.V, p->nnz, l.segXt.seg, l.segXt.nb);
  l.off_part = off; off += align_up((sx > sxt ? sx : sxt) * C * 4, 256);
  l.total = off;
  return l;
}

static int kl_check(const cnmf_b200_kl_problem* p, bool pointers) {
  CNMF_CHECK_ARG(p != nullptr, "kl: null problem");
  CNMF_CHECK_ARG(p->n_cells > 0 && p->n_genes > 0 && p->R > 0 && p->K > 0 && p->nnz >= 0, "kl: empty shape");
  CNMF_CHECK_ARG(p->nnz < (1ll << 31), "kl: nnz must fit int32 offsets");
  CNMF_CHECK_ARG(p->impl == CNMF_B200_KL_IMPL_SLAB || p->impl == CNMF_B200_KL_IMPL_GATHER, "kl: unknown impl %d", p->impl);
  CNMF_CHECK_ARG(p->n_cells < (1ll << 31) / 64 && p->n_genes < (1ll << 31) / 64, "kl: more than 2^25 rows or columns");
  CNMF_CHECK_ARG(p->Kp >= p->K && p->Kp % 2 == 0 && p->Kp <= CNMF_B200_MAX_KP,
                 "kl: Kp=%d must be even, >= K=%d and <= %d", p->Kp, p->K, CNMF_B200_MAX_KP);
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

// arguments of the slab kernel for CSR(X) (rows = cells, fixed factor W) or CSR(X^T) (rows = genes, fixed factor H)
static KlSlabArgs kl_slab_args(const cnmf_b200_kl_problem* p, const KlLayout& l, bool transposed, const KlSegs& sg,
                               int grid) {
  char* ws = reinterpret_cast<char*>(p->workspace);
  const uint2* W16 = reinterpret_cast<const uint2*>(ws + l.off_wh);
  const uint2* H16 = reinterpret_cast<const uint2*>(ws + l.off_hh);
  const KlSegLayout& o = transposed ? l.segXt : l.segX;
  KlSlabArgs a{};
  a.pairs = reinterpret_cast<const uint2*>(ws + (transposed ? l.off_pairsXt : l.off_pairsX));
  a.M = (int)(transposed ? p->n_genes : p->n_cells);
  a.C = (int)(transposed ? p->n_cells : p->n_genes);
  a.F16 = transposed ? H16 : W16;
  a.S16 = transposed ? W16 : H16;
  a.R = p->R;
  a.RP = (p->R + 1) / 2;
  a.ngroups = ks_num_groups(p->R, p->Kp);
  a.active = p->active;
  a.eps = p->eps;
  a.vseg_ptr = sg.vseg_ptr; a.seg_vrow = sg.seg_vrow; a.seg_beg = sg.seg_beg; a.vend = sg.vend;
  a.wcum = reinterpret_cast<const int*>(ws + (transposed ? l.off_wcumXt : l.off_wcumX));
  a.nb = o.nb; a.block = o.block; a.seg_len = o.seg;
  a.part = sg.part;
  a.opart = nullptr;
  a.grid = grid;
  return a;
}

// slab path: fp16 copies of both factors from the fp32 masters, segment table + work split of CSR(X) (and CSR(X^T))
static int kl_slab_prepare(const cnmf_b200_kl_problem* p, const KlLayout& l, bool both, KlSegs* sgX, KlSegs* sgXt,
                           cudaStream_t st) {
  char* ws = reinterpret_cast<char*>(p->workspace);
  int* flag = reinterpret_cast<int*>(ws + l.off_flag);
  float* part = reinterpret_cast<float*>(ws + l.off_part);
  if (p->warm != 0) {   // tables, pair tables and fp16 copies are what the previous call left: bind only
    CNMF_TRY(kl_build_segments(p->row_ptr, p->col_idx, p->n_cells, ws, l.segX, part, sgX, st, false));
    CNMF_TRY(kl_build_segments(p->t_row_ptr, p->t_col_idx, p->n_genes, ws, l.segXt, part, sgXt, st, false));
    return 0;
  }
  CNMF_TRY(launch_kl_to_half_pairs(p->Wt, ws + l.off_wh, p->n_genes, p->R, p->Kp, flag, st));
  CNMF_TRY(launch_kl_to_half_pairs(p->Hr, ws + l.off_hh, p->n_cells, p->R, p->Kp, flag, st));
  CNMF_TRY(kl_build_segments(p->row_ptr, p->col_idx, p->n_cells, ws, l.segX, part, sgX, st));
  CNMF_TRY(launch_kl_work_scan(*sgX, l.segX.V, p->Kp, reinterpret_cast<int*>(ws + l.off_wcumX), st));
  CNMF_TRY(launch_kl_pack_pairs(*sgX, l.segX.V, p->n_cells, l.segX.block, p->Kp, p->col_idx, p->val,
                                reinterpret_cast<const int*>(ws + l.off_wcumX), ws + l.off_pairsX, st));
  if (both) {
    CNMF_TRY(kl_build_segments(p->t_row_ptr, p->t_col_idx, p->n_genes, ws, l.segXt, part, sgXt, st));
    CNMF_TRY(launch_kl_work_scan(*sgXt, l.segXt.V, p->Kp, reinterpret_cast<int*>(ws + l.off_wcumXt), st));
    CNMF_TRY(launch_kl_pack_pairs(*sgXt, l.segXt.V, p->n_genes, l.segXt.block, p->Kp, p->t_col_idx, p->t_val,
                                  reinterpret_cast<const int*>(ws + l.off_wcumXt), ws + l.off_pairsXt, st));
  }
  return 0;
}

static int kl_objective_impl(const cnmf_b200_kl_problem* p, const KlLayout& l, cudaStream_t st,
                             const KlSegs* sgX = nullptr) {
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
  if (l.slab) {
    // sum x log(s) by the slab kernel on the fp16 copies (the caller has made them current): [SMs][R] partials
    CNMF_CUDA(cudaMemsetAsync(opart, 0, (size_t)l.nblk_obj * p->R * 8, st));
    KlSlabArgs a = kl_slab_args(p, l, /*transposed=*/false, *sgX, l.nblk_obj);
    a.opart = opart;
    CNMF_TRY(launch_kl_slab(a, p->Kp, 1, st));
  } else {
    const unsigned grid = (unsigned)ceil_div((int64_t)l.nblk_obj * p->R, KL_THREADS);
    KL_DISPATCH_KP(p->Kp, (kl_xlogs_kernel<KP><<<grid, KL_THREADS, 0, st>>>(p->row_ptr, p->col_idx, p->val, p->n_cells,
                                                                                p->Hr, p->Wt, p->R, p->eps, opart)));
    CNMF_LAUNCH_CHECK();
  }
  double* wsq64 = nullptr;
  double* hsq64 = nullptr;
  if (p->l2_H != 0.f || p->l2_W != 0.f) {   // hsum / wsum (fp32 outputs) are scratch here: the next iteration recomputes them
    wsq64 = reinterpret_cast<double*>(ws + l.off_wsq64);
    hsq64 = reinterpret_cast<double*>(ws + l.off_hsq64);
    CNMF_TRY(launch_colsum(p->Wt, p->n_genes, C, wsum, wsq64, cpart, st, true));
    CNMF_TRY(launch_colsum(p->Hr, p->n_cells, C, hsum, hsq64, cpart, st, true));
  }
  kl_objective_kernel<<<p->R, 256, 0, st>>>(p->xlogx, p->sumX, opart, l.nblk_obj, hsum64, wsum64, p->R, p->Kp, p->l1_H,
                                           p->l1_W, hsq64, wsq64, p->l2_H, p->l2_W, p->err2);
  CNMF_LAUNCH_CHECK();
  return 0;
}

}  // namespace cnmf

using namespace cnmf;

extern "C" {

int cnmf_b200_kl_constants(const float* val, int64_t nnz, double* out2, void* stream) {
  CNMF_NVTX("cnmf_b200_kl_constants");
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
  CNMF_NVTX("cnmf_b200_to_half");
  CNMF_CHECK_ARG(src && dst && n >= 0, "to_half: bad argument");
  return launch_to_half(src, reinterpret_cast<__half*>(dst), n, (cudaStream_t)stream);
}

int cnmf_b200_kl_half_step(const int32_t* row_ptr, const int32_t* col_idx, const float* val, int64_t M,
                           const void* F_half, float* state, void* state_half, const float* fsum,
                           const uint8_t* active, int32_t R, int32_t Kp, float l1, float l2, float eps, void* stream) {
  CNMF_NVTX("cnmf_b200_kl_half_step");
  CNMF_CHECK_ARG(row_ptr && col_idx && val && F_half && state && fsum, "kl_half_step: null pointer");
  CNMF_CHECK_ARG(M > 0 && R > 0, "kl_half_step: empty shape");
  CNMF_CHECK_ARG((reinterpret_cast<uintptr_t>(F_half) & 7u) == 0 && aligned16(state) && aligned16(fsum),
                 "kl_half_step: F_half must be 8-byte, state and fsum 16-byte aligned");
  return launch_kl_half(row_ptr, col_idx, val, M, reinterpret_cast<const __half*>(F_half), state,
                        reinterpret_cast<__half*>(state_half), fsum, active, R, Kp, l1, l2, eps, (cudaStream_t)stream);
}

int cnmf_b200_kl_iterate(const cnmf_b200_kl_problem* p, int32_t n_iters, void* stream) {
  CNMF_NVTX("cnmf_b200_kl_iterate");
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
  if (n_iters == 0) return 0;
  if (l.slab) {
    KlSegs sgX{}, sgXt{};
    int* flag = reinterpret_cast<int*>(ws + l.off_flag);
    CNMF_TRY(kl_slab_prepare(p, l, true, &sgX, &sgXt, st));   // cold: copies and tables rebuilt; warm: bound only
    const KlSlabArgs aX = kl_slab_args(p, l, false, sgX, 0), aXt = kl_slab_args(p, l, true, sgXt, 0);
    for (int it = 0; it < n_iters; ++it) {
      CNMF_TRY(launch_colsum(p->Wt, p->n_genes, C, wsum, nullptr, cpart, st));
      CNMF_TRY(launch_kl_slab(aX, p->Kp, 0, st));
      CNMF_TRY(launch_kl_finish_slab(p->n_cells, p->Hr, ws + l.off_hh, wsum, p->active, p->R, p->Kp, p->l1_H, p->l2_H,
                                     p->eps, sgX, flag, st));
      CNMF_TRY(launch_colsum(p->Hr, p->n_cells, C, hsum, nullptr, cpart, st));
      CNMF_TRY(launch_kl_slab(aXt, p->Kp, 0, st));
      CNMF_TRY(launch_kl_finish_slab(p->n_genes, p->Wt, ws + l.off_wh, hsum, p->active, p->R, p->Kp, p->l1_W, p->l2_W,
                                     p->eps, sgXt, flag, st));
    }
    if (p->err2 != nullptr) CNMF_TRY(kl_objective_impl(p, l, st, &sgX));
    return 0;
  }
  __half* Wh = reinterpret_cast<__half*>(ws + l.off_wh);
  __half* Hh = reinterpret_cast<__half*>(ws + l.off_hh);
  int* gflag = reinterpret_cast<int*>(ws + l.off_flag);
  const bool cold = p->warm == 0;   // warm: tables and the fp16 copy of W are what the previous iterate call left
  if (cold) {
    CNMF_TRY(launch_range_check(p->Hr, p->n_cells * C, gflag, st));
    CNMF_TRY(launch_range_check(p->Wt, p->n_genes * C, gflag, st));
    CNMF_TRY(launch_to_half(p->Wt, Wh, p->n_genes * C, st));  // Hh is written by the first H half-step
  }
  KlSegs sgX{}, sgXt{};
  float* part = reinterpret_cast<float*>(ws + l.off_part);
  CNMF_TRY(kl_build_segments(p->row_ptr, p->col_idx, p->n_cells, ws, l.segX, part, &sgX, st, cold));
  CNMF_TRY(kl_build_segments(p->t_row_ptr, p->t_col_idx, p->n_genes, ws, l.segXt, part, &sgXt, st, cold));
  const int64_t nsX = kl_nseg_max(l.segX.V, p->nnz, l.segX.seg), nsXt = kl_nseg_max(l.segXt.V, p->nnz, l.segXt.seg);
  const int64_t mmX = kl_multi_max(p->nnz, l.segX.seg), mmXt = kl_multi_max(p->nnz, l.segXt.seg);
  for (int it = 0; it < n_iters; ++it) {
    CNMF_TRY(launch_colsum(p->Wt, p->n_genes, C, wsum, nullptr, cpart, st));
    CNMF_TRY(launch_kl_half(p->row_ptr, p->col_idx, p->val, p->n_cells, Wh, p->Hr, Hh, wsum, p->active, p->R, p->Kp,
                            p->l1_H, p->l2_H, p->eps, st, &sgX, nsX, mmX, kl_rep_group(p->n_genes, p->Kp)));
    if (it == 0 && cold && p->active != nullptr)  // frozen replicates skip the write: their copy must exist too
      CNMF_TRY(launch_to_half(p->Hr, Hh, p->n_cells * C, st));
    CNMF_TRY(launch_colsum(p->Hr, p->n_cells, C, hsum, nullptr, cpart, st));
    CNMF_TRY(launch_kl_half(p->t_row_ptr, p->t_col_idx, p->t_val, p->n_genes, Hh, p->Wt, Wh, hsum, p->active, p->R,
                            p->Kp, p->l1_W, p->l2_W, p->eps, st, &sgXt, nsXt, mmXt));
  }
  CNMF_TRY(launch_range_check(p->Hr, p->n_cells * C, gflag, st));
  CNMF_TRY(launch_range_check(p->Wt, p->n_genes * C, gflag, st));
  if (p->err2 != nullptr) CNMF_TRY(kl_objective_impl(p, l, st));
  return 0;
}

int cnmf_b200_kl_objective(const cnmf_b200_kl_problem* p, void* stream) {
  CNMF_NVTX("cnmf_b200_kl_objective");
  reset_launch_count();
  CNMF_TRY(kl_check(p, true));
  CNMF_CHECK_ARG(p->err2 != nullptr, "kl_objective: err2 is NULL");
  const KlLayout l = kl_layout(p);
  if (l.slab) {
    KlSegs sgX{}, sgXt{};
    CNMF_TRY(kl_slab_prepare(p, l, true, &sgX, &sgXt, (cudaStream_t)stream));   // (both tables: the next call may be warm)
    return kl_objective_impl(p, l, (cudaStream_t)stream, &sgX);
  }
  return kl_objective_impl(p, l, (cudaStream_t)stream);
}

int cnmf_b200_kl_range_flag(const cnmf_b200_kl_problem* p, int32_t* flag_host, void* stream) {
  CNMF_NVTX("cnmf_b200_kl_range_flag");
  CNMF_TRY(kl_check(p, true));
  CNMF_CHECK_ARG(flag_host != nullptr, "kl_range_flag: null output");
  const KlLayout l = kl_layout(p);
  cudaStream_t st = (cudaStream_t)stream;
  CNMF_CUDA(cudaMemcpyAsync(flag_host, reinterpret_cast<char*>(p->workspace) + l.off_flag, 4, cudaMemcpyDeviceToHost, st));
  CNMF_CUDA(cudaStreamSynchronize(st));
  return 0;
}

int cnmf_b200_kl_range_flag_clear(const cnmf_b200_kl_problem* p, void* stream) {
  CNMF_NVTX("cnmf_b200_kl_range_flag_clear");
  CNMF_TRY(kl_check(p, true));
  const KlLayout l = kl_layout(p);
  CNMF_CUDA(cudaMemsetAsync(reinterpret_cast<char*>(p->workspace) + l.off_flag, 0, 4, (cudaStream_t)stream));
  return 0;
}

}  // extern "C"
