// Dense half-step on the 5th-gen tensor cores: TMA (SWIZZLE_128B) -> shared memory ->
// tcgen05.mma kind::tf32 (fp32 accumulate in TMEM) -> fused row update in the epilogue.
//
//   D[128 rows of A  x  BN = t*KP columns]  =  A_tile[128 x Kd] * F_tile[BN x Kd]^T
//
// A is X (H half-step) or X^T (W half-step); F is the stacked fixed factor ([R*KP x Kd], K-major);
// the epilogue thread that owns TMEM lane i owns row i of the tile: it reads its KP accumulator
// columns per replicate with tcgen05.ld, applies the MU / HALS row update (common.cuh, same code
// as the SIMT baseline) against the replicate's Gram in shared memory and writes the state
// [R*KP x M] (coalesced across the warp).  P = A F^T is never materialised.
//
// Warp roles (256 threads, 1 CTA / SM, persistent over tiles):
//   warp 0 lane 0 : TMA producer          warp 1 lane 0 : MMA issuer
//   warp 2        : TMEM alloc / dealloc  warps 4..7    : epilogue (TMEM lane quarter = warp % 4)
// Pipelines: smem full/empty ring (TMA <-> MMA), TMEM full/empty double buffer (MMA <-> epilogue).
//
// Replaces x @ W.T / H.T @ x and the update that follows (pattern in the reference at
// /root/reference/src/cnmf/cnmf.py:355-372) for all R replicates of one K in one pass over X.
#include <cuda.h>
#include <cudaTypedefs.h>

#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <type_traits>
#include <vector>

#include "common.cuh"

namespace cnmf {

constexpr int TC_BM = 128;
constexpr int TC_BK = 32;  // floats per k-block: 128 bytes = one swizzle-128B row
constexpr int TC_THREADS = 256;  // 4 control warps + 4 epilogue warps: the minimum; see tc_epi_groups()
// Epilogue warpgroups (128 threads each).  All of them cover the 128 TMEM lanes (warp % 4 = lane quarter);
// warpgroup g takes the replicates rr = g (mod groups) of a tile.  The row update is a per-thread dependency
// chain (HALS sweeps k in order), so the only way to hide its latency is more resident warps per scheduler;
// the register file (64K) bounds how many: ~100 regs/thread at KP<=12 allows 3 groups + the control warps.
#ifndef TC_EG12
#define TC_EG12 2
#endif
#ifndef TC_PF12
#define TC_PF12 1
#endif
__host__ __device__ constexpr int tc_epi_groups(int kp) { return kp <= 8 ? 3 : (kp <= 12 ? TC_EG12 : (kp <= 20 ? 2 : 1)); }
__host__ __device__ constexpr int tc_threads(int kp) { return 128 + 128 * tc_epi_groups(kp); }
constexpr int TC_MAX_BN = 256;
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4;  // 16 KiB
constexpr int TC_SMEM_LIMIT = 227 * 1024;
constexpr int TC_MAX_STAGES = 6;
constexpr int TC_MODE_STORE = 2;  // epilogue modes beyond CNMF_B200_ALGO_MU (0) / _HALS (1)
constexpr int TC_MODE_ACCUM = 3;

struct TcParams {
  int64_t M;       // rows of A (and columns of the state array)
  int64_t Kd;      // contraction length
  float* state;    // [R*KP x lds] fp32 master
  float* state_op; // same layout, tf32-rounded copy that the NEXT half-step's MMA reads (may be null)
  int64_t lds;
  const float* gram;      // [R x KP x KP], or (gram_ntile > 0) the per-tile partials [R x gram_ntile x KP*KP]
  int gram_ntile;
  const uint8_t* active;  // [R] or null
  const uint8_t* rep_k;   // [R] true K of every replicate (KP-3 .. KP), or null: all KP
  float* state_lo;        // same layout as state_op: tf32(state - tf32(state)), the part the operand copy drops (may be null)
  int npass;              // 1, or 2: second pass of the k loop over the operand of tensor map B2
  double* partials;       // [num_m_tiles x R x 2]
  int R;
  int t;  // replicates per tile
  int num_m_tiles, num_m_groups, num_n_tiles, stages;
  float l1, l2, eps;
  int helpers;  // 1: warps 0-3 join the epilogue of the CTA's last tile (MU/HALS, CG = 1)
  int dbg;  // dev probes (CNMF_B200_TC_DEBUG): 1 = main loop without MMAs, 2 = main loop without TMA loads
};

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug traps (kernel aborts with an error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 28)) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// One lane of a converged warp (deterministic for a full mask).  The producer and MMA warps run their loops with
// all 32 lanes and issue through the elected lane: inside an `if (lane == 0)` region ptxas cannot prove
// convergence and wraps every UTCHMMA / UTCBAR / UBLKCP in an ELECT ... BRA.U.ANY loop, which made the issue
// loop (not the tensor pipe) the limiter of the main loop (profiles/NOTES.md, "issue-bound main loop").
__device__ __forceinline__ bool elect_one() {
  uint32_t el;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(el));
  return el != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float& a, float& b, float& c, float& d) {
  uint32_t r0, r1, r2, r3;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(taddr));
  a = __uint_as_float(r0); b = __uint_as_float(r1); c = __uint_as_float(r2); d = __uint_as_float(r3);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float to_tf32(float v) {  // round to nearest, ties away (cvt.rna)
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}
template <int NT>
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory"); }
// the epilogue barrier of a CTA's LAST tile, which warps 0-3 join (see `helpers`): NT epilogue threads + 128
template <int NT>
__device__ __forceinline__ void epi_bar_sync(bool with_helpers) {
  if (with_helpers) asm volatile("bar.sync 2, %0;" ::"n"(NT + 128) : "memory");
  else asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
}

// K-major operand tile, rows at 128-byte pitch, SWIZZLE_128B: SBO = 8 rows * 128 B, LBO unused,
// descriptor version 1 (sm_100), layout type 2.
__device__ __forceinline__ uint64_t make_kmajor_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1024u >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

// ---------------------------------------------------------------------------------------------
// kernel.  CG = 1: one CTA per tile.  CG = 2: a CTA pair (cluster of 2, tcgen05 cta_group::2) works on
// a 256-row x BN tile: each CTA stages its own 128 rows of A and HALF of the B tile, the leader's
// single thread issues M=256 MMAs that read both CTAs' shared memory, each CTA keeps its 128 x BN
// accumulator half in its own tensor memory and runs its own epilogue.  Per SM that cuts the operand
// bytes per k-block from (128 + BN) to (128 + BN/2) rows -- the measured limiter (profiles/NOTES.md).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA load issued by either CTA of a pair; the bytes are accounted on the LEADER's mbarrier
// (peer bit of the shared::cluster address cleared).
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {  // arrives on this barrier in BOTH CTAs
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(static_cast<uint16_t>(3))
               : "memory");
}
__device__ __forceinline__ void tc_mma_tf32_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar, uint32_t cta) {  // arrive on `bar` of CTA `cta`
  asm volatile(
      "{\n\t"
      ".reg .b32 remAddr32;\n\t"
      "mapa.shared::cluster.u32  remAddr32, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64  _, [remAddr32];\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}

// Work iterator shared by the three roles of a CTA: whole tiles, tile = group, group + groups, ...
struct TcWork {
  int num_tiles, num_groups, tile;
  __device__ __forceinline__ TcWork(int group_id, int num_groups_, int num_tiles_)
      : num_tiles(num_tiles_), num_groups(num_groups_), tile(group_id) {}
  __device__ __forceinline__ bool next(int& t_out) {
    if (tile >= num_tiles) return false;
    t_out = tile;
    tile += num_groups;
    return true;
  }
};

template <int KP, int ALGO, int CG, bool SUMS>
__global__ void __launch_bounds__(tc_threads(KP), 1)
tc_half_step_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmB2, const TcParams prm) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment by pointer arithmetic ON the __shared__ array: a round trip through uintptr_t loses the
  // address space and every later access (the Gram reads of the epilogue) becomes a generic LD.E instead of LDS
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = prm.t;
  const int BN = t * KP;
  const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0u;
  const uint32_t b_bytes = static_cast<uint32_t>(BN / CG) * TC_BK * 4;  // this CTA's share of the B tile
  const uint32_t stage_bytes = TC_A_BYTES + b_bytes;
  const int stages = prm.stages;

  float* sgram = reinterpret_cast<float*>(smem + static_cast<size_t>(stages) * stage_bytes);
  float* sred = sgram + t * KP * KP;                       // [4 warps][t][2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sred + 4 * t * 2);  // 8-byte aligned: all counts even
  uint64_t* full = bars;
  uint64_t* empty = bars + stages;
  uint64_t* tfull = empty + stages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  uint8_t* sact = reinterpret_cast<uint8_t*>(tmem_slot + 2);  // [t] active flags of the tile's replicates
  uint8_t* srepk = sact + 64;                                 // [t] true K of the tile's replicates (<= KP)
  float* sginv = reinterpret_cast<float*>(srepk + 64);        // [t][KP] HALS: 1/(G_kk + l2), < 0 where frozen

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    if (prm.npass > 1) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB2)) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull[b], 1);
      mbar_init(&tempty[b], 4 * tc_epi_groups(KP) * CG);  // one arrival per epilogue warp of every CTA of the group
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    if (CG == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"(static_cast<uint32_t>(TC_TMEM_COLS)));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"(static_cast<uint32_t>(TC_TMEM_COLS)));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // tiles are indexed per CTA group: group tile -> (m_group, n_tile); this CTA's rows = m_group*CG + rank
  const int num_tiles = prm.num_m_groups * prm.num_n_tiles;
  const int group_id = blockIdx.x / CG, num_groups = gridDim.x / CG;
  const int num_kb = static_cast<int>((prm.Kd + TC_BK - 1) / TC_BK);

  if (warp == 0) {
    // ===================== TMA producer (every CTA; whole warp, one elected lane issues) =====================
    int stage = 0;
    uint32_t phase = 0;
    TcWork work(group_id, num_groups, num_tiles);
    int tile;
    while (work.next(tile)) {
      const int m_group = tile / prm.num_n_tiles, n_tile = tile - m_group * prm.num_n_tiles;
      const int m_tile = m_group * CG + static_cast<int>(cta_rank);
      // npass = 2: the k loop runs twice, the second time over a second fixed-factor operand (the tf32 residual of the
      // first): the accumulator then holds A (F + F2)^T, the product with the factor to ~2^-22
      for (int kk = 0; kk < prm.npass * num_kb; ++kk) {
        const int kb = kk < num_kb ? kk : kk - num_kb;
        const CUtensorMap* mapB = kk < num_kb ? &tmB : &tmB2;
        mbar_wait(&empty[stage], phase ^ 1u);
        if (elect_one()) {
          uint8_t* sa = smem + static_cast<size_t>(stage) * stage_bytes;
          if (CG == 2 && (prm.dbg & 2)) {
            if (leader) mbar_arrive(&full[stage]);
          } else if (CG == 2) {
            if (leader) mbar_expect_tx(&full[stage], 2u * stage_bytes);  // both CTAs' bytes land on the leader
            tma_load_2d_pair(sa, &tmA, &full[stage], kb * TC_BK, m_tile * TC_BM);
            tma_load_2d_pair(sa + TC_A_BYTES, mapB, &full[stage], kb * TC_BK,
                             n_tile * BN + static_cast<int>(cta_rank) * (BN / 2));
          } else if (prm.dbg & 2) {
            mbar_arrive(&full[stage]);
          } else {
            mbar_expect_tx(&full[stage], stage_bytes);
            tma_load_2d(sa, &tmA, &full[stage], kb * TC_BK, m_tile * TC_BM);
            tma_load_2d(sa + TC_A_BYTES, mapB, &full[stage], kb * TC_BK, n_tile * BN);
          }
        }
        __syncwarp();
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      // ===================== MMA issuer (leader CTA of the group; whole warp, one elected lane issues) ==========
      // instruction descriptor: D=f32 (bits 4-5 = 1), A=B=tf32 (2 at bits 7-9 / 10-12), K-major both,
      // N>>3 at bits 17-22, M>>4 at bits 24-28 (M = 128 per CTA, 256 for a pair)
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(BN >> 3) << 17) |
                             (static_cast<uint32_t>((TC_BM * CG) >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int lt = 0;
      TcWork work(group_id, num_groups, num_tiles);
      int tile;
      for (; work.next(tile); ++lt) {
        const int buf = lt & 1;
        const uint32_t tphase = (lt >> 1) & 1u;
        mbar_wait(&tempty[buf], tphase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(buf * TC_MAX_BN);
        for (int kb = 0; kb < prm.npass * num_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            if (prm.dbg & 1) {
              if (CG == 2) { mbar_arrive_cta(&empty[stage], 0u); mbar_arrive_cta(&empty[stage], 1u); }
              else mbar_arrive(&empty[stage]);
            } else {
              const uint32_t sa = smem_u32(smem + static_cast<size_t>(stage) * stage_bytes);
              const uint64_t a_desc = make_kmajor_sw128_desc(sa);
              const uint64_t b_desc = make_kmajor_sw128_desc(sa + TC_A_BYTES);
#pragma unroll
              for (int k = 0; k < TC_BK / 8; ++k) {
                // advance 8 tf32 = 32 bytes inside the 128-byte swizzle row: +2 in the (addr >> 4) field
                if (CG == 2)
                  tc_mma_tf32_pair(d_tmem, a_desc + static_cast<uint64_t>(2 * k), b_desc + static_cast<uint64_t>(2 * k),
                                   idesc, (kb > 0 || k != 0) ? 1u : 0u);
                else
                  tc_mma_tf32(d_tmem, a_desc + static_cast<uint64_t>(2 * k), b_desc + static_cast<uint64_t>(2 * k),
                              idesc, (kb > 0 || k != 0) ? 1u : 0u);
              }
              // frees the smem slot (in both CTAs of a pair) when these MMAs retire
              if (CG == 2) tc_commit_pair(&empty[stage]); else tc_commit(&empty[stage]);
            }
          }
          __syncwarp();
          if (++stage == stages) { stage = 0; phase ^= 1u; }
        }
        // accumulator complete -> epilogue (of both CTAs)
        if (elect_one()) {
          if (CG == 2) tc_commit_pair(&tfull[buf]); else tc_commit(&tfull[buf]);
        }
        __syncwarp();
      }
    }
  }
  // The update of one tile is instruction-issue bound and takes about half a main loop; it hides behind the next
  // tile's main loop except for the LAST tile of the CTA, when the producer, MMA and allocator warps have nothing
  // left to do: they join that tile's epilogue as one more warpgroup (replicates dealt over EG + 1 groups).
  const bool helpers = (CG == 1) && (ALGO < TC_MODE_STORE) && prm.helpers != 0;
  if (warp >= 4 || helpers) {
    // ===================== epilogue (every CTA, its own 128 rows) =====================
    constexpr int EG = tc_epi_groups(KP), ET = 128 * EG;
    const bool helper = warp < 4;
    const int q = warp & 3;         // TMEM lane quarter == warp % 4
    const int eg = helper ? EG : (warp - 4) >> 2;  // epilogue warpgroup: takes replicates rr = eg (mod groups)
    const int et = helper ? (1 << 30) : static_cast<int>(threadIdx.x) - 128;  // helpers skip the cooperative loops
    int lt = 0;
    TcWork work(group_id, num_groups, num_tiles);
    int tile;
    for (; work.next(tile); ++lt) {
      const bool last = helpers && tile + num_groups >= num_tiles;
      if (helper && !last) continue;
      const int egs = last ? EG + 1 : EG;   // warpgroups sharing this tile
      const int buf = lt & 1;
      const uint32_t tphase = (lt >> 1) & 1u;
      const int m_group = tile / prm.num_n_tiles, n_tile = tile - m_group * prm.num_n_tiles;
      const int m_tile = m_group * CG + static_cast<int>(cta_rank);
      const int r0 = n_tile * t;
      if (ALGO < TC_MODE_STORE) {
        for (int i = et; i < t * KP * KP; i += ET) {
          const int rr = i / (KP * KP);
          float g = 0.f;
          if (r0 + rr < prm.R) {
            if (prm.gram_ntile > 0) {
              // fused final stage of the Gram: fixed-order fp64 sum of the per-tile partials (these warps
              // are otherwise idle while the first tile's main loop runs)
              const float* pp = prm.gram + (static_cast<int64_t>(r0 + rr) * prm.gram_ntile) * (KP * KP) + (i - rr * KP * KP);
              double acc = 0.0;
              for (int tl = 0; tl < prm.gram_ntile; ++tl) acc += static_cast<double>(pp[static_cast<int64_t>(tl) * KP * KP]);
              g = static_cast<float>(acc);
            } else {
              g = prm.gram[static_cast<int64_t>(r0) * KP * KP + i];
            }
          }
          sgram[i] = g;
        }
      }
      // active flags of the tile's replicates: read once here instead of one dependent global load per replicate
      // inside the update loop
      for (int i = et; i < t; i += ET) {
        sact[i] = (r0 + i < prm.R && (prm.active == nullptr || prm.active[r0 + i] != 0)) ? 1 : 0;
        srepk[i] = (prm.rep_k != nullptr && r0 + i < prm.R) ? prm.rep_k[r0 + i] : static_cast<uint8_t>(KP);
      }
      epi_bar_sync<ET>(last);
      if (ALGO == CNMF_B200_ALGO_HALS) {
        // reciprocals of the regularised diagonal, once per tile (these warps idle here while the main loop runs)
        for (int i = et; i < t * KP; i += ET) {
          const int rr = i / KP, k = i - rr * KP;
          const float gkk = sgram[rr * KP * KP + k * KP + k] + prm.l2;
          sginv[i] = (gkk < prm.eps) ? -1.f : __frcp_rn(gkk);
        }
        epi_bar_sync<ET>(last);
      }
      mbar_wait(&tfull[buf], tphase);
      tc_fence_after();
      const int64_t m = static_cast<int64_t>(m_tile) * TC_BM + q * 32 + lane;
      const bool in = m < prm.M;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(buf * TC_MAX_BN);
      // software pipeline over the replicates of the tile: the state of replicate rr+1 is requested
      // before replicate rr is updated, so its global-load latency hides behind the update
      constexpr bool kPrefetch = (KP <= 8 || (KP == 12 && TC_PF12)) && (ALGO != TC_MODE_STORE);
      constexpr bool want_sums = SUMS && (ALGO < TC_MODE_STORE);  // compile-time: p is dead after the update otherwise
      float s[KP], s_nxt[kPrefetch ? KP : 1];
      if (kPrefetch && r0 + eg < prm.R) {
        const float* sp0 = prm.state + static_cast<int64_t>(r0 + eg) * KP * prm.lds + m;
#pragma unroll
        for (int k = 0; k < KP; ++k) s[k] = in ? sp0[static_cast<int64_t>(k) * prm.lds] : 0.f;
      }
      for (int rr = eg; rr < t; rr += egs) {
        const int r = r0 + rr;
        if (r >= prm.R) break;
        float p[KP];
#pragma unroll
        for (int c = 0; c < KP; c += 4)
          tmem_ld4(taddr + static_cast<uint32_t>(rr * KP + c), p[c], p[c + 1], p[c + 2], p[c + 3]);
        const int64_t soff = static_cast<int64_t>(r) * KP * prm.lds + m;
        float* sp = prm.state + soff;
        if (kPrefetch) {
          if (rr + egs < t && r + egs < prm.R) {
            const float* spn = sp + static_cast<int64_t>(KP) * egs * prm.lds;
#pragma unroll
            for (int k = 0; k < KP; ++k) s_nxt[k] = in ? spn[static_cast<int64_t>(k) * prm.lds] : 0.f;
          }
        } else {
#pragma unroll
          for (int k = 0; k < KP; ++k)
            s[k] = (in && ALGO != TC_MODE_STORE) ? sp[static_cast<int64_t>(k) * prm.lds] : 0.f;
        }
        tmem_ld_wait();
        const bool act = sact[rr] != 0;
        float cross = 0.f, ssum = 0.f;
        if (ALGO >= TC_MODE_STORE) {
          // plain GEMM epilogues of the online path: state = P (store) or state += P (accumulate)
          if (in && act) {
#pragma unroll
            for (int k = 0; k < KP; ++k)
              sp[static_cast<int64_t>(k) * prm.lds] = (ALGO == TC_MODE_ACCUM) ? s[k] + p[k] : p[k];
          }
        } else {
          // The replicate's own K (a stack holds K = KP-3 .. KP padded to KP): the update, the state traffic and the
          // sums run over KE components only -- the padded ones are zero and stay zero.  One instantiation per KE;
          // predicated-off instructions would still take issue slots in this issue-bound loop.
          auto body = [&](auto ke_tag) {
            constexpr int KE = decltype(ke_tag)::value;
            if (in) {
              if (act) {
                update_row<KP, ALGO, KE>(p, s, sgram + rr * KP * KP, prm.l1, prm.l2, prm.eps,
                                         ALGO == CNMF_B200_ALGO_HALS ? sginv + rr * KP : nullptr);
#pragma unroll
                for (int k = 0; k < KE; ++k) sp[static_cast<int64_t>(k) * prm.lds] = s[k];
                if (prm.state_op != nullptr) {
                  float* op = prm.state_op + soff;
#pragma unroll
                  for (int k = 0; k < KE; ++k) op[static_cast<int64_t>(k) * prm.lds] = to_tf32(s[k]);
                  if (prm.state_lo != nullptr) {
                    float* lo = prm.state_lo + soff;
#pragma unroll
                    for (int k = 0; k < KE; ++k) lo[static_cast<int64_t>(k) * prm.lds] = to_tf32(s[k] - to_tf32(s[k]));
                  }
                }
              }
              if (want_sums) {
#pragma unroll
                for (int k = 0; k < KE; ++k) {
                  cross = fmaf(s[k], p[k], cross);
                  ssum += s[k];
                }
              }
            }
          };
          switch (KP - static_cast<int>(srepk[rr])) {
            case 1: body(std::integral_constant<int, (KP > 1 ? KP - 1 : KP)>{}); break;
            case 2: body(std::integral_constant<int, (KP > 2 ? KP - 2 : KP)>{}); break;
            case 3: body(std::integral_constant<int, (KP > 3 ? KP - 3 : KP)>{}); break;
            default: body(std::integral_constant<int, KP>{}); break;
          }
          if (want_sums) {
            cross = warp_sum(cross);
            ssum = warp_sum(ssum);
            if (lane == 0) {
              sred[(q * t + rr) * 2 + 0] = cross;
              sred[(q * t + rr) * 2 + 1] = ssum;
            }
          }
        }
        if (kPrefetch) {
#pragma unroll
          for (int k = 0; k < KP; ++k) s[k] = s_nxt[k];
        }
      }
      // all tcgen05.ld of this buffer are complete: hand it back to the MMA warp (of the leader CTA)
      tc_fence_before();
      __syncwarp();
      if (lane == 0 && !helper) {   // (nobody waits for the last tile's buffer)
        if (CG == 2) mbar_arrive_cta(&tempty[buf], 0u); else mbar_arrive(&tempty[buf]);
      }
      epi_bar_sync<ET>(last);
      for (int i = et; i < t * 2 && want_sums && m_tile < prm.num_m_tiles; i += ET) {
        const int rr = i >> 1, r = r0 + rr;
        if (r < prm.R) {
          double v = 0.0;
#pragma unroll
          for (int w = 0; w < 4; ++w) v += static_cast<double>(sred[(w * t + rr) * 2 + (i & 1)]);
          prm.partials[(static_cast<int64_t>(m_tile) * prm.R + r) * 2 + (i & 1)] = v;
        }
      }
      epi_bar_sync<ET>(last);  // sgram / sred / sact are rewritten by the next tile
    }
  }

  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (CG == 2)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                   "r"(static_cast<uint32_t>(TC_TMEM_COLS)));
    else
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                   "r"(static_cast<uint32_t>(TC_TMEM_COLS)));
  }
}

__global__ void round_tf32_kernel(float* __restrict__ dst, const float* __restrict__ src, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = to_tf32(src[i]);
}

__global__ void round_tf32_2d_kernel(float* __restrict__ dst, const float* __restrict__ src, int64_t rows, int64_t cols,
                                     int64_t ld) {
  const int64_t total = rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols, c = i - r * cols;
    dst[r * ld + c] = to_tf32(src[r * ld + c]);
  }
}

int launch_round_tf32_2d(float* dst, const float* src, int64_t rows, int64_t cols, int64_t ld, cudaStream_t st) {
  if (rows * cols == 0) return 0;
  int64_t blocks = ceil_div(rows * cols, 256 * 8);
  if (blocks > 148 * 16) blocks = 148 * 16;
  round_tf32_2d_kernel<<<(unsigned)blocks, 256, 0, st>>>(dst, src, rows, cols, ld);
  CNMF_LAUNCH_CHECK();
  return 0;
}

// dst = tf32(master - rounded): the part of the fp32 master that its tf32 operand copy does not carry.
// A second GEMM pass on it makes the objective's cross term exact to ~2^-22 instead of ~2^-11.
__global__ void residual_tf32_kernel(float* __restrict__ dst, const float* __restrict__ master,
                                     const float* __restrict__ rounded, int64_t rows, int64_t cols, int64_t ld) {
  const int64_t total = rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols, c = i - r * cols;
    dst[r * ld + c] = to_tf32(master[r * ld + c] - rounded[r * ld + c]);
  }
}

int launch_residual_tf32(float* dst, const float* master, const float* rounded, int64_t rows, int64_t cols, int64_t ld,
                         cudaStream_t st) {
  if (rows * cols == 0) return 0;
  int64_t blocks = ceil_div(rows * cols, 256 * 8);
  if (blocks > 148 * 16) blocks = 148 * 16;
  residual_tf32_kernel<<<(unsigned)blocks, 256, 0, st>>>(dst, master, rounded, rows, cols, ld);
  CNMF_LAUNCH_CHECK();
  return 0;
}

int launch_round_tf32(float* dst, const float* src, int64_t n, cudaStream_t st) {
  if (n == 0) return 0;
  int64_t blocks = ceil_div(n, 256 * 8);
  if (blocks > 148 * 16) blocks = 148 * 16;
  round_tf32_kernel<<<(unsigned)blocks, 256, 0, st>>>(dst, src, n);
  CNMF_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
struct TcPlan {
  int t, BN, stages, num_m_tiles, num_m_groups, num_n_tiles, cg;
  size_t smem_bytes;
};

static size_t tc_smem_bytes(int t, int Kp, int stages, int cg) {
  const size_t stage = TC_A_BYTES + static_cast<size_t>(t) * Kp / cg * TC_BK * 4;
  size_t extra = static_cast<size_t>(t) * Kp * Kp * 4 + 4 * static_cast<size_t>(t) * 2 * 4;
  extra += (2 * TC_MAX_STAGES + 4) * 8 + 16 + 128;  // barriers, TMEM slot, per-tile active flags and K (t <= 64)
  extra += static_cast<size_t>(t) * Kp * 4;         // reciprocal diagonals (HALS)
  return stages * stage + extra + 1024;  // +1024: manual alignment of the dynamic base
}

static int tc_num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
      sms = 148;
  }
  return sms;
}

// Pick the CTA group size (1, or a pair sharing the B tile) and replicates-per-tile t (BN = t*Kp, a
// multiple of 16, <= 256) from a cost model calibrated on B200 (profiles/NOTES.md): one SM ingests
// ~43 B/clk from L2 through TMA, so a k-block costs max(operand bytes per SM / 43, tensor-pipe clocks);
// tiles run in waves over the SMs (pairs: over SM pairs); the epilogue of the last tile is exposed.
static bool tc_plan(int64_t M, int64_t Kd, int R, int Kp, int sms, TcPlan* out) {
  const int num_m = static_cast<int>(ceil_div(M, TC_BM));
  const int64_t num_kb = ceil_div(Kd, TC_BK);
  double best = 1e300;
  bool found = false;
  static const int forced_t = [] { const char* e = getenv("CNMF_B200_TC_T"); return e ? atoi(e) : 0; }();
  static const int forced_cg = [] { const char* e = getenv("CNMF_B200_TC_CG"); return e ? atoi(e) : 0; }();
  for (int cg = 1; cg <= 2; ++cg) {
    // CTA pairs (cta_group::2, M = 256 per MMA, each SM stages half of the B tile) are the default where two row tiles
    // exist: measured 13 % faster on both cfg-2 stacks (H half-step of the Kp=12 stack 98.9 -> 86.1 us, the plain GEMM
    // 74.7 -> 65.6 us against 72.4 us for the library's tf32 GEMM of the same shape; profiles/NOTES.md, round 2).
    // CNMF_B200_TC_CG=1 / =2 force one form (dev knob).
    if (cg == 2 && (num_m < 2 || sms < 2 || forced_cg == 1)) continue;
    if (forced_cg == 2 && cg == 1 && num_m >= 2) continue;
    for (int t = 1; t * Kp <= TC_MAX_BN; ++t) {
      const int BN = t * Kp;
      if (BN % 16 != 0) continue;
      if (forced_t > 0 && t != forced_t && (forced_t * Kp) % 16 == 0 && forced_t * Kp <= TC_MAX_BN) continue;
      static const int max_stages = [] { const char* e = getenv("CNMF_B200_TC_STAGES"); return e ? atoi(e) : TC_MAX_STAGES; }();
      int stages = max_stages < TC_MAX_STAGES ? max_stages : TC_MAX_STAGES;
      while (stages >= 2 && tc_smem_bytes(t, Kp, stages, cg) > static_cast<size_t>(TC_SMEM_LIMIT)) --stages;
      if (stages < 2) continue;
      const int num_n = static_cast<int>(ceil_div(R, t));
      const int num_g = static_cast<int>(ceil_div(num_m, cg));
      const int64_t tiles = static_cast<int64_t>(num_g) * num_n;
      const int64_t waves = ceil_div(tiles, sms / cg);
      // clocks per k-block, measured (profiles/NOTES.md): every operand byte crosses the 128 B/clk shared-memory
      // port twice (TMA write + MMA read): 2*(128+BN) clk, over a floor of ~500 clk from the barrier round trip and
      // the 4 MMA issues.  (Pairs: half the B bytes per SM; measured gain 13-14 % per k-block.)
      double kb_clk = 2.0 * (TC_BM + BN) + 40.0;
      if (kb_clk < 500.0 + 0.6 * BN) kb_clk = 500.0 + 0.6 * BN;
      if (cg == 2) kb_clk *= 0.86;
      const double epi = static_cast<double>(t) * (Kp * Kp + 12.0 * Kp) * 1.5 + 1500.0;
      const double cost = waves * (static_cast<double>(num_kb) * kb_clk + 1000.0) + epi + (cg == 2 ? 1500.0 : 0.0);
      if (cost < best) {
        best = cost;
        found = true;
        out->t = t; out->BN = BN; out->stages = stages; out->num_m_tiles = num_m; out->num_m_groups = num_g;
        out->num_n_tiles = num_n; out->cg = cg;
        out->smem_bytes = tc_smem_bytes(t, Kp, stages, cg);
      }
    }
  }
  return found;
}

bool tc_supported(int64_t M, int64_t Kd, int R, int Kp) {
  TcPlan p;
  return M > 0 && Kd > 0 && R > 0 && Kp % 4 == 0 && Kp <= CNMF_B200_MAX_KP && M < (1ll << 31) &&
         Kd < (1ll << 31) && tc_plan(M, Kd, R, Kp, 148, &p);
}

static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

struct MapKey {
  const void* ptr; int64_t rows, cols, ld; int box_rows;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows;
  }
};
struct MapEntry { MapKey key; CUtensorMap map; };

// 2-D fp32 tensor [rows x cols] (row pitch ld), box = [TC_BK cols x box_rows rows], 128B swizzle,
// out-of-bounds elements read as zero.
static int get_tensor_map(const float* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows, CUtensorMap* out) {
  static thread_local std::vector<MapEntry> cache;
  MapKey key{ptr, rows, cols, ld, box_rows};
  for (auto& e : cache)
    if (e.key == key) { *out = e.map; return 0; }
  auto encode = get_encode_fn();
  if (!encode) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return CNMF_B200_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 4};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(TC_BK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult rc = encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), gdim, gstride, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld ld=%lld box=%d", (int)rc, (long long)rows,
              (long long)cols, (long long)ld, box_rows);
    return CNMF_B200_ERR_CUDA;
  }
  if (cache.size() >= 64) cache.erase(cache.begin());
  cache.push_back({key, *out});
  return 0;
}

template <int KP, int ALGO, int CG, bool SUMS>
static int launch_tc_inst3(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mb2, const TcParams& prm, const TcPlan& plan,
                           int grid, cudaStream_t st) {
  static unsigned long long configured = 0;  // per-device bit mask (benign race: setting twice is harmless)
  auto kern = tc_half_step_kernel<KP, ALGO, CG, SUMS>;
  if (needs_func_attr(&configured))
    CNMF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_LIMIT));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(tc_threads(KP));
  cfg.dynamicSmemBytes = plan.smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CNMF_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, mb2, prm));
  count_launch();
  return 0;
}

template <int KP, int ALGO, int CG>
static int launch_tc_inst2(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mb2, const TcParams& prm, const TcPlan& plan,
                           int grid, cudaStream_t st) {
  if (ALGO < TC_MODE_STORE && prm.partials != nullptr) return launch_tc_inst3<KP, ALGO, CG, true>(ma, mb, mb2, prm, plan, grid, st);
  return launch_tc_inst3<KP, ALGO, CG, false>(ma, mb, mb2, prm, plan, grid, st);
}

template <int KP, int ALGO>
static int launch_tc_inst(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mb2, const TcParams& prm, const TcPlan& plan,
                          int grid, cudaStream_t st) {
  if (plan.cg == 2) return launch_tc_inst2<KP, ALGO, 2>(ma, mb, mb2, prm, plan, grid, st);
  return launch_tc_inst2<KP, ALGO, 1>(ma, mb, mb2, prm, plan, grid, st);
}

int launch_tc_half_step_g(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                          float* state, float* state_op, int64_t lds, const float* gram, int gram_ntile,
                          const uint8_t* active, int R, int Kp, int algo, float l1, float l2, float eps,
                          double* partials, cudaStream_t st, const uint8_t* rep_k = nullptr, const float* F2 = nullptr,
                          float* state_lo = nullptr);

int launch_tc_half_step(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                        float* state, float* state_op, int64_t lds, const float* gram, const uint8_t* active, int R,
                        int Kp, int algo, float l1, float l2, float eps, double* partials, cudaStream_t st) {
  return launch_tc_half_step_g(A, lda, M, Kd, F, ldf, state, state_op, lds, gram, 0, active, R, Kp, algo, l1, l2, eps,
                               partials, st);
}

// gram_ntile > 0: `gram` holds the Gram kernel's per-tile partials and the epilogue warps sum them
int launch_tc_half_step_g(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                          float* state, float* state_op, int64_t lds, const float* gram, int gram_ntile,
                          const uint8_t* active, int R, int Kp, int algo, float l1, float l2, float eps,
                          double* partials, cudaStream_t st, const uint8_t* rep_k, const float* F2, float* state_lo) {
  CNMF_NVTX("frob: tcgen05 half-step (GEMM + fused MU/HALS update)");
  TcPlan plan;
  const int sms = tc_num_sms();
  if (!tc_plan(M, Kd, R, Kp, sms, &plan)) {
    set_error("tcgen05 path: no tile plan for M=%lld Kd=%lld R=%d Kp=%d", (long long)M, (long long)Kd, R, Kp);
    return CNMF_B200_ERR_UNSUPPORTED;
  }
  CUtensorMap ma, mb;
  CNMF_TRY(get_tensor_map(A, M, Kd, lda, TC_BM, &ma));
  CNMF_TRY(get_tensor_map(F, static_cast<int64_t>(R) * Kp, Kd, ldf, plan.BN / plan.cg, &mb));
  CUtensorMap mb2 = mb;   // second fixed-factor operand (F2: the tf32 residual of F): the k loop then runs over both
  if (F2 != nullptr) CNMF_TRY(get_tensor_map(F2, static_cast<int64_t>(R) * Kp, Kd, ldf, plan.BN / plan.cg, &mb2));
  TcParams prm;
  prm.M = M; prm.Kd = Kd; prm.state = state; prm.state_op = state_op; prm.lds = lds; prm.gram = gram; prm.gram_ntile = gram_ntile; prm.active = active; prm.rep_k = rep_k; prm.state_lo = state_lo; prm.npass = F2 != nullptr ? 2 : 1;
  prm.partials = partials; prm.R = R; prm.t = plan.t; prm.num_m_tiles = plan.num_m_tiles; prm.num_m_groups = plan.num_m_groups;
  prm.num_n_tiles = plan.num_n_tiles; prm.stages = plan.stages; prm.l1 = l1; prm.l2 = l2; prm.eps = eps;
  static const int dbg = [] { const char* e = getenv("CNMF_B200_TC_DEBUG"); return e ? atoi(e) : 0; }();
  prm.dbg = dbg;
  static const int helpers = [] { const char* e = getenv("CNMF_B200_TC_HELPERS"); return e ? atoi(e) : 1; }();
  prm.helpers = helpers;
  const int64_t tiles = static_cast<int64_t>(plan.num_m_groups) * plan.num_n_tiles;  // per CTA group
  const int64_t slots = sms / plan.cg;
  const int grid = static_cast<int>((tiles < slots ? tiles : slots) * plan.cg);
  if (algo == CNMF_B200_ALGO_MU) {
    CNMF_DISPATCH_KP(Kp, return (launch_tc_inst<KP, CNMF_B200_ALGO_MU>(ma, mb, mb2, prm, plan, grid, st)));
  } else if (algo == CNMF_B200_ALGO_HALS) {
    CNMF_DISPATCH_KP(Kp, return (launch_tc_inst<KP, CNMF_B200_ALGO_HALS>(ma, mb, mb2, prm, plan, grid, st)));
  } else if (algo == TC_MODE_STORE) {
    CNMF_DISPATCH_KP(Kp, return (launch_tc_inst<KP, TC_MODE_STORE>(ma, mb, mb2, prm, plan, grid, st)));
  } else if (algo == TC_MODE_ACCUM) {
    CNMF_DISPATCH_KP(Kp, return (launch_tc_inst<KP, TC_MODE_ACCUM>(ma, mb, mb2, prm, plan, grid, st)));
  }
  set_error("unknown algo %d", algo);
  return CNMF_B200_ERR_INVALID;
}

}  // namespace cnmf

using namespace cnmf;

extern "C" {

int cnmf_b200_round_tf32(float* x, int64_t n, void* stream) {
  CNMF_NVTX("cnmf_b200_round_tf32");
  CNMF_CHECK_ARG(x != nullptr && n >= 0, "round_tf32: bad argument");
  return launch_round_tf32(x, x, n, (cudaStream_t)stream);
}

int cnmf_b200_dense_half_step(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                              float* state, float* state_tf32, int64_t lds, const float* gram,
                              const uint8_t* active, int32_t R, int32_t Kp, int32_t algo, float l1, float l2,
                              float eps, double* partials, void* stream) {
  CNMF_NVTX("cnmf_b200_dense_half_step");
  CNMF_CHECK_ARG(A && F && state && gram, "dense_half_step: null pointer");
  CNMF_CHECK_ARG(M > 0 && Kd > 0 && R > 0, "dense_half_step: empty shape");
  CNMF_CHECK_ARG(lda % 4 == 0 && ldf % 4 == 0 && aligned16(A) && aligned16(F),
                 "dense_half_step: operands need 16-byte aligned rows (ld multiple of 4)");
  CNMF_CHECK_ARG(lda >= Kd && ldf >= Kd && lds >= M, "dense_half_step: leading dimension too small");
  return launch_tc_half_step(A, lda, M, Kd, F, ldf, state, state_tf32, lds, gram, active, R, Kp, algo, l1, l2, eps,
                             partials, (cudaStream_t)stream);
}

/* cnmf_b200_dense_half_step for a stack that holds several K padded to one Kp: rep_k[r] (device, [R]) is the true K of
 * replicate r (Kp-3 .. Kp); the epilogue updates, loads and stores that many components (rows beyond stay zero). */
int cnmf_b200_dense_half_step_k(const float* A, int64_t lda, int64_t M, int64_t Kd, const float* F, int64_t ldf,
                                float* state, float* state_tf32, int64_t lds, const float* gram,
                                const uint8_t* active, const uint8_t* rep_k, int32_t R, int32_t Kp, int32_t algo,
                                float l1, float l2, float eps, double* partials, void* stream) {
  CNMF_NVTX("cnmf_b200_dense_half_step_k");
  CNMF_CHECK_ARG(A && F && state && gram, "dense_half_step_k: null pointer");
  CNMF_CHECK_ARG(M > 0 && Kd > 0 && R > 0, "dense_half_step_k: empty shape");
  CNMF_CHECK_ARG(lda % 4 == 0 && ldf % 4 == 0 && aligned16(A) && aligned16(F),
                 "dense_half_step_k: operands need 16-byte aligned rows (ld multiple of 4)");
  CNMF_CHECK_ARG(lda >= Kd && ldf >= Kd && lds >= M, "dense_half_step_k: leading dimension too small");
  return launch_tc_half_step_g(A, lda, M, Kd, F, ldf, state, state_tf32, lds, gram, 0, active, R, Kp, algo, l1, l2, eps,
                               partials, (cudaStream_t)stream, rep_k);
}

}  // extern "C"
