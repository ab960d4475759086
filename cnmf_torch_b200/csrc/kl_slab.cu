// Sparse KL half-step around a shared-memory slab of the fixed factor.
//
// The first sparse kernel (kl_sparse.cu: one thread per (row, replicate), the fixed factor's fp16 copy gathered
// from L2 for every stored entry and replicate) moved 70 x its algorithmic bytes from L2 to L1 -- 7.3 GB per
// half-step at 20 000 cells, 8-9 TB/s of L2 bandwidth -- and spent a quarter of its instructions unpacking fp16.
// Here
//   * a CTA owns a GROUP of replicates and a block of columns of the CSR matrix; the fixed factor's fp16 values
//     for that (block, group) are staged ONCE in shared memory, one 128-byte line per column, and every gather
//     is a shared-memory load;
//   * the line layout makes those loads bank-conflict-free for ARBITRARY column indices: a lane of a half-warp is
//     a (entry slot a, replicate pair rp) and reads, at load m, the 8-byte word rp*P + ((m + a) mod P) of its
//     column's line -- the 8-byte bank it hits depends on the lane alone (P = Kp/2 words per replicate pair).
//     Each lane therefore holds its operands ROTATED by a; the dot product does not care, and the numerators are
//     put back in order by a shuffle round per register at the end of a row segment;
//   * a lane serves TWO replicates per stored entry (the index/value load and the address arithmetic are shared),
//     a warp 2 x P entries per step (one set per half-warp), all of ONE row segment: no divergence on row lengths,
//     index loads are contiguous;
//   * arithmetic: fma.rn.f32.f16 (SASS FHFMA: fp16 x fp16 + fp32 -> fp32, half selectors in the operand, full
//     FFMA rate on sm_100a -- profiles/r02_simt_rate.txt) on the packed words as they come out of shared memory:
//     no unpacking.  s = h.w with both operands fp16 (11-bit significands, like the tf32 operands of the dense
//     path), the ratio q = x / s rounded to fp16 for the second FMA chain, accumulation fp32.  The staged values
//     are x * 2^-8 (KS_XSCALE; undone exactly where the partials are summed), so that fp16 holds ratios up to
//     1.6e7 instead of 65504 -- badly fitted entries of a long run do reach 1e5 -- at full precision down to 0.016
//     (smaller ratios, which weigh nothing in a numerator of ratios near 1, lose bits gradually).  An overflow
//     beyond that turns into inf and is caught where the factor is written.
// Rows are cut into segments (kl_tables.cuh) at column-block boundaries and at `seg_len` entries; every segment
// leaves its Kp numerators per replicate in `part`, and kl_finish_slab_kernel adds a row's segments in order,
// applies the multiplicative update to the fp32 master and writes the fp16 copy (deterministic, no atomics).
// Work is split statically: the (group, segment) space, weighted by steps per segment, is cut into gridDim.x equal
// contiguous ranges, so a CTA reloads its slab only when its range crosses a column block or a group.
//
// Replaces run_nmf(beta_loss='kullback-leibler') (external nmf-torch, called at
// /root/reference/src/cnmf/cnmf.py:819); update rule sklearn/decomposition/_nmf.py:562-607 with the reference's
// epsilon rule (cnmf.py:370-371).
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"
#include "kl_fhfma.cuh"
#include "kl_slab.cuh"
#include "kl_tables.cuh"

namespace cnmf {

template <int KP>
struct SlabCfg {
  static constexpr int P = KP / 2;        // 8-byte words per replicate pair and column: {pair of halves of rep 0, of rep 1}
  static constexpr int RGP = 16 / P;      // replicate pairs per 128-byte line
  static constexpr int A = P;             // entry slots per half-warp (needs A * RGP <= 16: RGP = floor(16 / P))
  static constexpr int LANES = A * RGP;   // working lanes per half-warp
  static_assert(KP % 2 == 0 && KP >= 2 && KP <= 32, "even Kp up to 32");
  static_assert(RGP >= 1 && LANES <= 16, "a line is 16 words");
};

__host__ __device__ static inline int ks_rgp(int Kp) { return 16 / (Kp / 2); }

__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ float fast_rcp(float v) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}
// mbarrier + bulk-copy wrappers (shared::cta addresses)
__device__ __forceinline__ void ks_bar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void ks_bar_expect(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ks_bar_wait(uint32_t bar, uint32_t parity) {   // bounded: a protocol bug traps
  uint32_t ok = 0, spins = 0;
  while (true) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) break;
    if (++spins > (1u << 28)) __trap();
  }
}
__device__ __forceinline__ void ks_bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// first index i in [0, n] with a[i] >= key (a non-decreasing, a[n] = total)
__device__ __forceinline__ int ks_lower_bound(const int* __restrict__ a, int n, int64_t key) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((int64_t)a[mid] < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// MODE 0: numerators of the multiplicative update per segment -> part.  MODE 1: sum x log(s) per replicate -> opart.
template <int KP, int MODE>
__global__ void __launch_bounds__(KS_THREADS, 1) kl_slab_kernel(const KlSlabArgs a) {
  using Cfg = SlabCfg<KP>;
  constexpr int P = Cfg::P, RGP = Cfg::RGP, A = Cfg::A, LANES = Cfg::LANES;
  constexpr int STEP_BYTES = 2 * A * 8;                      // (column, value) pairs of one warp step
  constexpr int CHS = KS_CHUNK_BYTES / STEP_BYTES;           // steps per staged chunk
  static_assert(CHS >= 1 && STEP_BYTES % 16 == 0, "a chunk holds whole steps; bulk copies move multiples of 16 bytes");
  extern __shared__ __align__(128) unsigned char slab[];
  __shared__ double s_obj[KS_WARPS][2 * RGP];
  __shared__ __align__(8) uint64_t s_bar[KS_WARPS][KS_RING + 2];   // staging ring, then the two line buffers of h

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, hw = lane >> 4, li = lane & 15;
  const bool lane_on = li < LANES;
  const int aslot = lane_on ? li / RGP : 0, rp = lane_on ? li % RGP : 0;
  const int M = a.M;
  const int nseg = a.vseg_ptr[(int64_t)M * a.nb];
  const int64_t W = a.wcum[nseg];
  const int64_t T = W * a.ngroups;
  const int64_t lo = T * blockIdx.x / gridDim.x, hi = T * (blockIdx.x + 1) / gridDim.x;

  const uint32_t slab_u32 = static_cast<uint32_t>(__cvta_generic_to_shared(slab));
  // per-warp ring of staging chunks, after the lines
  const uint32_t stg = slab_u32 + (uint32_t)a.block * KS_LINE + (uint32_t)warp * (KS_RING * KS_CHUNK_BYTES);
  const uint32_t bar0 = static_cast<uint32_t>(__cvta_generic_to_shared(&s_bar[warp][0]));
  if (lane < KS_RING + 2) ks_bar_init(bar0 + 8u * lane, 1);
  const uint32_t hbar0 = bar0 + 8u * KS_RING;
  const uint32_t hbuf0 = slab_u32 + (uint32_t)a.block * KS_LINE + KS_STAGE_BYTES + (uint32_t)warp * (2 * KS_LINE);
  int hfill = 0, huse = 0;            // line buffer of the next request / of the next segment
  uint32_t hpar = 0;                  // bit i: parity of the phase buffer i completes next
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  int fill_buf = 0, use_buf = 0;      // ring positions of the next chunk to request / to consume
  uint32_t use_par = 0;               // parity of the barrier phase the next consumed chunk completes

  uint32_t loff[P];   // byte offset of this lane's m-th load inside a line (and inside a row of the moving factor's
                      // copy): word index (m + a) mod P of pair rp
#pragma unroll
  for (int m = 0; m < P; ++m) {
    int t = m + aslot;
    if (t >= P) t -= P;
    // lanes beyond LANES have no work: they read one of the line's unused words (banks nobody else touches)
    loff[m] = lane_on ? (uint32_t)(rp * P + t) * 8u : (uint32_t)li * 8u;
  }
  const int64_t rkp = (int64_t)a.R * KP;
  const uint32_t lane_entry = (uint32_t)(hw * A + aslot) * 8u;   // this lane's entry inside a step's 2A entries

  int64_t pos = lo;
  while (pos < hi && W > 0) {
    const int g = (int)(pos / W);
    const int64_t goff = pos - (int64_t)g * W;
    const int64_t gend = (hi - (int64_t)g * W) < W ? (hi - (int64_t)g * W) : W;
    const int s_begin = ks_lower_bound(a.wcum, nseg, goff);
    const int s_end = gend >= W ? nseg : ks_lower_bound(a.wcum, nseg, gend);
    pos = (int64_t)g * W + gend;

    const int rp0 = g * RGP;
    const int npair = (a.RP - rp0) < RGP ? (a.RP - rp0) : RGP;
    const bool pair_on = lane_on && rp < npair;
    const int rep0 = 2 * (rp0 + rp), rep1 = rep0 + 1;
    // (the objective pass, MODE 1, covers frozen replicates too: their value is reported with every check)
    bool group_active = a.active == nullptr || MODE == 1;
    if (!group_active)
      for (int r = 2 * rp0; r < 2 * (rp0 + npair) && r < a.R; ++r) group_active |= a.active[r] != 0;
    if (!group_active || s_begin >= s_end) continue;
    const bool on0 = pair_on && rep0 < a.R && (a.active == nullptr || a.active[rep0]);
    const bool on1 = pair_on && rep1 < a.R && (a.active == nullptr || a.active[rep1]);
    double dacc0 = 0.0, dacc1 = 0.0;   // MODE 1: this lane's share of sum x log2(s), replicates rep0 / rep1

    int s = s_begin;
    while (s < s_end) {
      const int b = a.nb > 1 ? a.seg_vrow[s] / M : 0;
      const int blk_end = a.vseg_ptr[(int64_t)(b + 1) * M];
      const int sb_end = s_end < blk_end ? s_end : blk_end;
      const int c0 = b * a.block;
      const int nl = (a.C - c0) < a.block ? (a.C - c0) : a.block;
      const uint32_t sbase = slab_u32 - ((uint32_t)c0 << 7);
      __syncthreads();   // every warp is done with the previous slab
      // lines of this (block, group): 128 bytes each, ngroups lines apart in the copy
      for (int w = threadIdx.x; w < nl * (KS_LINE / 16); w += KS_THREADS) {
        const int line = w >> 3, q = w & 7;
        *reinterpret_cast<uint4*>(slab + line * KS_LINE + q * 16) =
            __ldg(reinterpret_cast<const uint4*>(a.F16 + ((int64_t)(c0 + line) * a.ngroups + g) * 16) + q);
      }

      // The phase's segments [s, sb_end) are dealt to the warps as 16 CONTIGUOUS runs of equal weight, so that a
      // warp's (column, value) pairs -- stored per segment, padded to whole steps, in segment order -- are one
      // contiguous stream: it comes in as bulk copies of whole chunks (one instruction of one lane per KB, completion
      // on an mbarrier), KS_RING chunks ahead of the arithmetic, with no regard for segment boundaries.  (The
      // version before this one staged each segment's entries with 4-byte cp.async copies: ~100 instructions of
      // address arithmetic per chunk, a third of all instructions issued.)
      int bound = 0;
      if (lane <= KS_WARPS) {
        const int64_t w0 = a.wcum[s], w1 = a.wcum[sb_end];
        bound = s + ks_lower_bound(a.wcum + s, sb_end - s, w0 + (w1 - w0) * lane / KS_WARPS);
      }
      const int ws = __shfl_sync(0xffffffffu, bound, warp), we = __shfl_sync(0xffffffffu, bound, warp + 1);
      // stream of this warp, in steps of the pair table: a segment's first step is wcum[seg] - overhead * seg
      int64_t st_pos = 0;
      int st_left = 0;                 // steps not yet requested
      if (ws < we) {
        st_pos = (int64_t)a.wcum[ws] - (int64_t)KS_SEG_OVERHEAD_STEPS * ws;
        st_left = (int)((int64_t)a.wcum[we] - (int64_t)KS_SEG_OVERHEAD_STEPS * we - st_pos);
      }
      auto request_chunk = [&]() {     // the next chunk of the stream into ring slot fill_buf (all lanes call)
        const int n = st_left < CHS ? st_left : CHS;
        if (lane == 0) {
          const uint32_t bar = bar0 + 8u * fill_buf;
          ks_bar_expect(bar, (uint32_t)n * STEP_BYTES);
          ks_bulk_load(stg + (uint32_t)fill_buf * KS_CHUNK_BYTES, a.pairs + st_pos * (2 * A), (uint32_t)n * STEP_BYTES, bar);
        }
        st_pos += n;
        st_left -= n;
        fill_buf = fill_buf + 1 == KS_RING ? 0 : fill_buf + 1;
      };
      int use_left = st_left;          // steps not yet consumed
#pragma unroll 1
      for (int i = 0; i < KS_RING && st_left > 0; ++i) request_chunk();
      __syncthreads();   // the slab is complete

      // A segment's row of the moving factor is one 128-byte line of its copy: it comes in by a bulk copy, a segment
      // ahead, into one of two per-warp line buffers; every lane then picks its rotated words out of shared memory.
      auto request_h = [&](int row) {
        if (lane == 0) {
          const uint32_t bar = hbar0 + 8u * hfill;
          ks_bar_expect(bar, KS_LINE);
          ks_bulk_load(hbuf0 + (uint32_t)hfill * KS_LINE, a.S16 + ((int64_t)row * a.ngroups + g) * 16, KS_LINE, bar);
        }
        hfill ^= 1;
      };
      auto take_h = [&](uint2 (&h)[P]) {
        ks_bar_wait(hbar0 + 8u * huse, (hpar >> huse) & 1u);
        const uint32_t base = hbuf0 + (uint32_t)huse * KS_LINE;
#pragma unroll
        for (int m = 0; m < P; ++m) h[m] = lds64(base + loff[m]);
        hpar ^= 1u << huse;
        huse ^= 1;
      };

      int sg = ws;
      int steps = 0, nsteps = 0, wc = 0;
      uint2 hreg[P];
      if (sg < we) {
        wc = a.wcum[sg + 1];
        steps = wc - a.wcum[sg] - KS_SEG_OVERHEAD_STEPS;
        request_h(a.seg_vrow[sg] - b * M);
      }
      int chunk_left = 0;              // steps left in the chunk being consumed
      bool first_chunk = true;
      uint32_t ep = 0;
      uint2 jx_next = make_uint2(0u, 0u);
      while (sg < we) {
        const bool has_next = sg + 1 < we;
        __syncwarp();                  // every lane has taken its words of the line buffer the next request overwrites
        if (has_next) {                // next segment's table entries and row of the factor, a segment ahead
          request_h(a.seg_vrow[sg + 1] - b * M);
          const int wc2 = a.wcum[sg + 2];
          nsteps = wc2 - wc - KS_SEG_OVERHEAD_STEPS;
          wc = wc2;
        }
        take_h(hreg);
        float n0[KP], n1[KP];   // n0[2m], n0[2m+1]: numerators of k-pair (m + a) mod P, replicate rep0
#pragma unroll
        for (int k = 0; k < KP; ++k) { n0[k] = 0.f; n1[k] = 0.f; }
        float acc0 = 0.f, acc1 = 0.f;

        int seg_left = steps;
        while (seg_left > 0) {
          if (chunk_left == 0) {
            if (!first_chunk) {        // the chunk just consumed: every lane has read it, its slot takes the next request
              __syncwarp();
              if (st_left > 0) request_chunk();
            }
            first_chunk = false;
            ks_bar_wait(bar0 + 8u * use_buf, use_par);
            chunk_left = use_left < CHS ? use_left : CHS;
            use_left -= chunk_left;
            ep = stg + (uint32_t)use_buf * KS_CHUNK_BYTES + lane_entry;
            jx_next = lds64(ep);       // the step's (column, value) one step ahead of its use
            if (++use_buf == KS_RING) { use_buf = 0; use_par ^= 1u; }
          }
          const int n = seg_left < chunk_left ? seg_left : chunk_left;
          seg_left -= n;
          chunk_left -= n;
#pragma unroll 1
          for (int i = 0; i < n; ++i) {
            const uint2 jx = jx_next;                                     // (column, value bits); idle lanes: unused
            ep += STEP_BYTES;
            jx_next = lds64(ep);       // (past the chunk's last step: the next slot or the pad, never used)
            const uint32_t line = sbase + (lane_on ? (jx.x << 7) : ((uint32_t)c0 << 7));   // sbase = slab - c0 * 128
            const float x = lane_on ? __uint_as_float(jx.y) : 0.f;
            uint2 w[P];
#pragma unroll
            for (int m = 0; m < P; ++m) w[m] = lds64(line + loff[m]);
            // s = eps + h.w: the clamp max(s, eps) of the update rule without its instruction (eps = 1e-16 vanishes
            // in any s that is not itself degenerate); one dependent chain per replicate, the two interleave
            float s0 = a.eps, s1 = a.eps;
#pragma unroll
            for (int m = 0; m < P; ++m) {
              s0 = fhfma_lo(hreg[m].x, w[m].x, s0);
              s1 = fhfma_lo(hreg[m].y, w[m].y, s1);
              s0 = fhfma_hi(hreg[m].x, w[m].x, s0);
              s1 = fhfma_hi(hreg[m].y, w[m].y, s1);
            }
            if (MODE == 0) {
              const uint16_t q0 = f32_to_f16_bits(x * fast_rcp(s0)), q1 = f32_to_f16_bits(x * fast_rcp(s1));
#pragma unroll
              for (int m = 0; m < P; ++m) {
                n0[2 * m] = fhfma_q_lo(q0, w[m].x, n0[2 * m]);
                n0[2 * m + 1] = fhfma_q_hi(q0, w[m].x, n0[2 * m + 1]);
                n1[2 * m] = fhfma_q_lo(q1, w[m].y, n1[2 * m]);
                n1[2 * m + 1] = fhfma_q_hi(q1, w[m].y, n1[2 * m + 1]);
              }
            } else {
              acc0 = fmaf(x, __log2f(s0), acc0);   // x = 0 on idle lanes and in the padding of a segment
              acc1 = fmaf(x, __log2f(s1), acc1);
            }
          }
        }

        if (MODE == 0) {
          // un-rotate: the lane with slot a' ends up with k-pair a' -- register m of the lane (a' - m) mod P
          float t0x = n0[0], t0y = n0[1], t1x = n1[0], t1y = n1[1];
#pragma unroll
          for (int m = 1; m < P; ++m) {
            int sa = aslot - m;
            if (sa < 0) sa += P;
            const int src = (lane & 16) + sa * RGP + rp;
            t0x += __shfl_sync(0xffffffffu, n0[2 * m], src);
            t0y += __shfl_sync(0xffffffffu, n0[2 * m + 1], src);
            t1x += __shfl_sync(0xffffffffu, n1[2 * m], src);
            t1y += __shfl_sync(0xffffffffu, n1[2 * m + 1], src);
          }
          // the two half-warps worked on different entries of the same segment; half-warp 0 writes rep0, 1 rep1
          t0x += __shfl_xor_sync(0xffffffffu, t0x, 16);
          t0y += __shfl_xor_sync(0xffffffffu, t0y, 16);
          t1x += __shfl_xor_sync(0xffffffffu, t1x, 16);
          t1y += __shfl_xor_sync(0xffffffffu, t1y, 16);
          const bool on = hw == 0 ? on0 : on1;
          if (on) {
            const int rep = hw == 0 ? rep0 : rep1;
            *reinterpret_cast<float2*>(a.part + (int64_t)sg * rkp + (int64_t)rep * KP + 2 * aslot) =
                hw == 0 ? make_float2(t0x, t0y) : make_float2(t1x, t1y);
          }
        } else {
          dacc0 += (double)acc0;
          dacc1 += (double)acc1;
        }
        ++sg; steps = nsteps;
      }
      s = sb_end;
    }

    if (MODE == 1) {
      __syncthreads();
      // per-warp, per-pair sums through shared memory: 32 lanes -> [2 * RGP] slots, fixed lane order
      for (int q = 0; q < 2 * RGP; ++q) {
        const int want_rp = q >> 1;
        double v = (lane_on && rp == want_rp) ? ((q & 1) ? dacc1 : dacc0) : 0.0;
        v = warp_sum(v);
        if (lane == 0) s_obj[warp][q] = v;
      }
      __syncthreads();
      if (threadIdx.x < 2 * npair) {
        const int rep = 2 * rp0 + threadIdx.x;
        if (rep < a.R) {
          double t = 0.0;
          for (int w2 = 0; w2 < KS_WARPS; ++w2) t += s_obj[w2][threadIdx.x];
          // ln 2 and the scale of the staged values (exact); a CTA meets a replicate once
          a.opart[(int64_t)blockIdx.x * a.R + rep] = t * (0.6931471805599453 / (double)KS_XSCALE);
        }
      }
    }
  }
}

// The pair table: segment s owns steps [wcum[s] - overhead * s, ...) of 2A = Kp entries each; its stored entries come
// first, the rest of its last step is (first column of the block, 0) -- a line that exists, a value that adds nothing.
__global__ void __launch_bounds__(256) kl_pack_pairs_kernel(const int* __restrict__ vseg_ptr, int64_t V, int64_t M,
                                                            int block, const int* __restrict__ seg_vrow,
                                                            const int* __restrict__ seg_beg, const int* __restrict__ vend,
                                                            int seg_len, int step_entries, int c0steps,
                                                            const int* __restrict__ col_idx, const float* __restrict__ val,
                                                            const int* __restrict__ wcum, uint2* __restrict__ pairs) {
  const int nseg = vseg_ptr[V];
  const int lane = threadIdx.x & 31;
  const int nw = gridDim.x * 8;
  for (int s = blockIdx.x * 8 + (threadIdx.x >> 5); s < nseg; s += nw) {
    const int v = seg_vrow[s];
    const int beg = seg_beg[s];
    int end = vend[v];
    if (end > beg + seg_len) end = beg + seg_len;
    const int w0 = wcum[s], w1 = wcum[s + 1];
    const int n = (w1 - w0 - c0steps) * step_entries;
    uint2* dst = pairs + ((int64_t)w0 - (int64_t)c0steps * s) * step_entries;
    const uint32_t padcol = (uint32_t)((v / M) * block);
    for (int e = lane; e < n; e += 32) {
      const int idx = beg + e;
      dst[e] = idx < end ? make_uint2((uint32_t)col_idx[idx], __float_as_uint(val[idx] * KS_XSCALE)) : make_uint2(padcol, 0u);
    }
  }
}

// All rows: numerators = partials of the row's segments, block by block and segment by segment (fixed order);
// multiplicative update of the fp32 master; fp16 copy in the line layout (see kl_to_half_pairs_kernel).
template <int KP>
__global__ void __launch_bounds__(128) kl_finish_slab_kernel(int64_t M, float* __restrict__ state,
                                                             uint32_t* __restrict__ state16,
                                                             const float* __restrict__ fsum,
                                                             const uint8_t* __restrict__ active, int R, int ngroups, float l1,
                                                             float l2, float eps, const int* __restrict__ vseg_ptr, int nb,
                                                             const float* __restrict__ part, int* __restrict__ flag) {
  constexpr int P = KP / 2;
  const int64_t g = (int64_t)blockIdx.x * 128 + threadIdx.x;
  const int64_t row = g / R;
  if (row >= M) return;
  const int r = (int)(g - row * R);
  if (active != nullptr && active[r] == 0) return;
  const int64_t rk = (int64_t)R * KP;
  float num[KP];
#pragma unroll
  for (int k = 0; k < KP; ++k) num[k] = 0.f;
  for (int b = 0; b < nb; ++b) {
    const int64_t v = (int64_t)b * M + row;
    for (int sgi = vseg_ptr[v]; sgi < vseg_ptr[v + 1]; ++sgi) {
      const float2* pp = reinterpret_cast<const float2*>(part + (int64_t)sgi * rk + (int64_t)r * KP);
#pragma unroll
      for (int m = 0; m < P; ++m) {
        const float2 t = pp[m];
        num[2 * m] += t.x;
        num[2 * m + 1] += t.y;
      }
    }
  }
  float2* sp = reinterpret_cast<float2*>(state + row * rk + (int64_t)r * KP);
  const float2* dp = reinterpret_cast<const float2*>(fsum + (int64_t)r * KP);
  constexpr int RGP = 16 / P;
  const int rpair = r >> 1;      // line of (row, group), word (pair in group) * P + m, half r & 1 of the word
  uint32_t* hp = state16 + ((row * ngroups + rpair / RGP) * 16 + (rpair % RGP) * P) * 2 + (r & 1);
  bool bad = false;
#pragma unroll
  for (int m = 0; m < P; ++m) {
    float2 h = sp[m];
    const float2 den = dp[m];
    const float d0 = fmaf(l2, h.x, den.x), d1 = fmaf(l2, h.y, den.y);
    const float u0 = num[2 * m] * (1.f / KS_XSCALE), u1 = num[2 * m + 1] * (1.f / KS_XSCALE);   // exact: a power of two
    const float a0 = l1 > 0.f ? fmaxf(u0 - l1, 0.f) : u0;
    const float a1 = l1 > 0.f ? fmaxf(u1 - l1, 0.f) : u1;
    h.x *= d0 < eps ? 0.f : __fdiv_rn(a0, d0);
    h.y *= d1 < eps ? 0.f : __fdiv_rn(a1, d1);
    sp[m] = h;
    bad |= !(h.x <= 65504.f) || !(h.y <= 65504.f);   // also catches NaN / inf from an overflowed ratio
    const __half2 hh = __floats2half2_rn(h.x, h.y);
    hp[2 * m] = *reinterpret_cast<const uint32_t*>(&hh);
  }
  if (bad) atomicOr(flag, 1);
}

// fp32 master [rows][R*KP] -> fp16 copy in the LINE layout the slab kernel stages and streams: one 128-byte line per
// (row, replicate group), 16 words of 8 bytes; word p * P + m = {k-pair m of replicate 2 * (g * RGP + p), the same of
// the next replicate}; words past RGP * P and replicates past R are zero.
template <int KP>
__global__ void __launch_bounds__(256) kl_to_half_pairs_kernel(const float* __restrict__ src, uint2* __restrict__ dst,
                                                               int64_t rows, int R, int ngroups, int* __restrict__ flag) {
  constexpr int P = KP / 2, RGP = 16 / P;
  const int64_t total = rows * ngroups * 16;
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int word = (int)(i & 15);
    const int64_t t = i >> 4;
    const int g = (int)(t % ngroups);
    const int64_t row = t / ngroups;
    const int pl = word / P, m = word - pl * P;
    const int r0 = 2 * (g * RGP + pl);
    float2 v0 = make_float2(0.f, 0.f), v1 = make_float2(0.f, 0.f);
    if (pl < RGP && r0 < R) v0 = *reinterpret_cast<const float2*>(src + row * R * KP + (int64_t)r0 * KP + 2 * m);
    if (pl < RGP && r0 + 1 < R) v1 = *reinterpret_cast<const float2*>(src + row * R * KP + (int64_t)(r0 + 1) * KP + 2 * m);
    bad |= !(v0.x <= 65504.f) || !(v0.y <= 65504.f) || !(v1.x <= 65504.f) || !(v1.y <= 65504.f);
    const __half2 h0 = __floats2half2_rn(v0.x, v0.y), h1 = __floats2half2_rn(v1.x, v1.y);
    dst[i] = make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
  }
  if (bad) atomicOr(flag, 1);
}

// wcum[s] = steps of the segments before s (a step = 2 * A entries of one segment, plus c0 steps of per-segment
// overhead): the static work split of kl_slab_kernel.  One CTA, chunked scan.
__global__ void __launch_bounds__(1024) kl_work_scan_kernel(const int* __restrict__ vseg_ptr, int64_t V,
                                                            const int* __restrict__ seg_vrow,
                                                            const int* __restrict__ seg_beg, const int* __restrict__ vend,
                                                            int seg_len, int step_entries, int c0, int* __restrict__ wcum) {
  __shared__ int s_w[1024];
  const int nseg = vseg_ptr[V];
  const int t = threadIdx.x;
  const int chunk = (nseg + 1023) / 1024;
  const int lo = t * chunk, hi = lo + chunk < nseg ? lo + chunk : nseg;
  int acc = 0;
  for (int s = lo; s < hi; ++s) {
    int len = vend[seg_vrow[s]] - seg_beg[s];
    if (len > seg_len) len = seg_len;
    acc += (len + step_entries - 1) / step_entries + c0;
  }
  s_w[t] = acc;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    const int v = t >= off ? s_w[t - off] : 0;
    __syncthreads();
    s_w[t] += v;
    __syncthreads();
  }
  acc = s_w[t] - acc;
  for (int s = lo; s < hi; ++s) {
    wcum[s] = acc;
    int len = vend[seg_vrow[s]] - seg_beg[s];
    if (len > seg_len) len = seg_len;
    acc += (len + step_entries - 1) / step_entries + c0;
  }
  if (t == 1023) wcum[nseg] = s_w[1023];
}

// ---- host side --------------------------------------------------------------------------------------------------
// (test knobs: CNMF_B200_KS_LINES caps the columns per block, CNMF_B200_KS_SEG the entries per segment, so that the
// small problems of the test-suite run through several column blocks and cut rows)
int ks_block_cols(int64_t n_cols) {
  static const int cap = [] { const char* e = getenv("CNMF_B200_KS_LINES"); const int v = e ? atoi(e) : 0;
                              return v > 0 && v < KS_MAX_LINES ? v : KS_MAX_LINES; }();
  const int64_t nb = ceil_div(n_cols, cap);
  return (int)ceil_div(n_cols, nb);
}
int ks_seg_len(int Kp) {
  static const int forced = [] { const char* e = getenv("CNMF_B200_KS_SEG"); return e ? atoi(e) : 0; }();
  const int step = 2 * (Kp / 2);                      // entries a warp takes per step
  const int len = forced > 0 ? forced : KS_SEG_LEN;
  return (len + step - 1) / step * step;              // whole steps
}

int launch_kl_work_scan(const KlSegs& sg, int64_t V, int Kp, int* wcum, cudaStream_t st) {
  kl_work_scan_kernel<<<1, 1024, 0, st>>>(sg.vseg_ptr, V, sg.seg_vrow, sg.seg_beg, sg.vend, sg.seg_len, 2 * (Kp / 2),
                                          KS_SEG_OVERHEAD_STEPS, wcum);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int64_t ks_pairs_max(int64_t nnz, int64_t nseg_max, int Kp) { return nnz + nseg_max * (2 * (Kp / 2)); }

int launch_kl_pack_pairs(const KlSegs& sg, int64_t V, int64_t M, int block, int Kp, const int* col_idx, const float* val,
                         const int* wcum, void* pairs, cudaStream_t st) {
  CNMF_NVTX("kl: pack (column, value) pairs");
  kl_pack_pairs_kernel<<<148 * 8, 256, 0, st>>>(sg.vseg_ptr, V, M, block > 0 ? block : 1, sg.seg_vrow, sg.seg_beg, sg.vend,
                                                sg.seg_len, 2 * (Kp / 2), KS_SEG_OVERHEAD_STEPS, col_idx, val, wcum,
                                                reinterpret_cast<uint2*>(pairs));
  CNMF_LAUNCH_CHECK();
  return 0;
}

int launch_kl_to_half_pairs(const float* src, void* dst, int64_t rows, int R, int Kp, int* flag, cudaStream_t st) {
  const int ngroups = ks_num_groups(R, Kp);
  const int64_t total = rows * ngroups * 16;
  int64_t blocks = ceil_div(total, 256 * 4);
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  KS_DISPATCH_KP(Kp, (kl_to_half_pairs_kernel<KP><<<(unsigned)blocks, 256, 0, st>>>(src, reinterpret_cast<uint2*>(dst), rows,
                                                                                    R, ngroups, flag)));
  CNMF_LAUNCH_CHECK();
  return 0;
}

template <int KP, int MODE>
static int launch_slab_inst(const KlSlabArgs& a, int grid, cudaStream_t st) {
  const int smem = a.block * KS_LINE + KS_STAGE_BYTES + KS_HBUF_BYTES;
  static unsigned long long attr_mask = 0;
  if (needs_func_attr(&attr_mask))
    CNMF_CUDA(cudaFuncSetAttribute(kl_slab_kernel<KP, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   KS_MAX_LINES * KS_LINE + KS_STAGE_BYTES + KS_HBUF_BYTES));
  kl_slab_kernel<KP, MODE><<<grid, KS_THREADS, smem, st>>>(a);
  CNMF_LAUNCH_CHECK();
  return 0;
}

static int ks_grid() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    sms[dev] = n;
  }
  return sms[dev];
}

int launch_kl_slab(const KlSlabArgs& a, int Kp, int mode, cudaStream_t st) {
  CNMF_NVTX("kl: slab half-step / objective pass");
  const int grid = a.grid > 0 ? a.grid : ks_grid();
  if (mode == 0) {
    KS_DISPATCH_KP(Kp, return (launch_slab_inst<KP, 0>(a, grid, st)));
  } else {
    KS_DISPATCH_KP(Kp, return (launch_slab_inst<KP, 1>(a, grid, st)));
  }
  return 0;
}

int launch_kl_finish_slab(int64_t M, float* state, void* state16, const float* fsum, const uint8_t* active, int R, int Kp,
                          float l1, float l2, float eps, const KlSegs& sg, int* flag, cudaStream_t st) {
  CNMF_NVTX("kl: finish rows (sum partials, update, fp16 line)");
  const int ngroups = ks_num_groups(R, Kp);
  const int64_t blocks = ceil_div(M * R, 128);
  CNMF_CHECK_ARG(blocks < (1ll << 31), "kl: rows * replicates too large for one launch");
  KS_DISPATCH_KP(Kp, (kl_finish_slab_kernel<KP><<<(unsigned)blocks, 128, 0, st>>>(
                         M, state, reinterpret_cast<uint32_t*>(state16), fsum, active, R, ngroups, l1, l2, eps, sg.vseg_ptr,
                         sg.nb, sg.part, flag)));
  CNMF_LAUNCH_CHECK();
  return 0;
}

int ks_num_groups(int R, int Kp) { return (int)ceil_div((R + 1) / 2, ks_rgp(Kp)); }
int ks_sm_count() { return ks_grid(); }

}  // namespace cnmf
