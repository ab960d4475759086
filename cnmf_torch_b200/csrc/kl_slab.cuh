// Interface of the shared-memory-slab sparse KL kernels (kl_slab.cu) towards the engine entry points in kl_sparse.cu.
#pragma once
#include "common.cuh"
#include "kl_tables.cuh"

namespace cnmf {

constexpr int KS_THREADS = 512;
constexpr int KS_WARPS = KS_THREADS / 32;
constexpr int KS_LINE = 128;               // bytes of shared memory per column of the block
constexpr int KS_RING = 3;                 // staging chunks in flight per warp
constexpr int KS_CHUNK_BYTES = 1024;       // one staging chunk: whole warp steps of (column, value) pairs, <= 1 KB
constexpr int KS_STAGE_BYTES = KS_WARPS * KS_RING * KS_CHUNK_BYTES + 256;   // 48 KB (+ the one-step read-ahead past the last chunk)
constexpr int KS_HBUF_BYTES = KS_WARPS * 2 * KS_LINE;   // two line buffers per warp for the moving factor's rows
constexpr int KS_MAX_LINES = 1360;         // 170 KB slab + 48 KB staging + 4 KB line buffers + static: within 227 KB
constexpr int KS_SEG_LEN = 640;            // entries per segment (64 steps at Kp = 10)
constexpr float KS_XSCALE = 1.f / 256.f;   // the staged values of X are scaled by this (see kl_slab.cu: range of the fp16 ratio)
constexpr int KS_SEG_OVERHEAD_STEPS = 2;   // set-up + un-rotation of a segment, in steps, for the static work split

#define KS_DISPATCH_KP(kp, ...)                                                                       \
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

struct KlSlabArgs {
  const uint2* pairs;      // (column, value bits) of every segment, padded to whole warp steps, in segment order
                           // (kl_pack_pairs_kernel): step t of the table is entries [t * Kp, (t + 1) * Kp)
  int M, C;                // rows; columns (= rows of the fixed factor)
  const uint2* F16;        // fixed factor, fp16 line layout [C][ngroups] x 128 bytes (kl_to_half_pairs_kernel)
  const uint2* S16;        // moving factor (the rows being updated), same layout [M][ngroups] x 128 bytes
  int R, RP, ngroups;      // replicates, replicate pairs, groups of RGP pairs
  const uint8_t* active;   // [R] or NULL
  float eps;
  const int* vseg_ptr;     // segment table (kl_tables.cuh), block-major, one partial slot per segment
  const int* seg_vrow;
  const int* seg_beg;
  const int* vend;
  const int* wcum;         // [nseg + 1] steps before each segment (kl_work_scan_kernel)
  int nb, block, seg_len;
  float* part;             // mode 0: [nseg][R*Kp]
  double* opart;           // mode 1: [grid][R], zeroed by the caller
  int grid;                // 0: one CTA per SM
};

int ks_block_cols(int64_t n_cols);            // columns per block: the smallest number of equal blocks that fit the slab
int ks_seg_len(int Kp);                      // entries per segment: a whole number of warp steps
int ks_num_groups(int R, int Kp);
int ks_sm_count();
int launch_kl_work_scan(const KlSegs& sg, int64_t V, int Kp, int* wcum, cudaStream_t st);
int64_t ks_pairs_max(int64_t nnz, int64_t nseg_max, int Kp);   // entries of the padded pair table, upper bound
int launch_kl_pack_pairs(const KlSegs& sg, int64_t V, int64_t M, int block, int Kp, const int* col_idx, const float* val,
                         const int* wcum, void* pairs, cudaStream_t st);
int launch_kl_to_half_pairs(const float* src, void* dst, int64_t rows, int R, int Kp, int* flag, cudaStream_t st);
int launch_kl_slab(const KlSlabArgs& a, int Kp, int mode, cudaStream_t st);
int launch_kl_finish_slab(int64_t M, float* state, void* state16, const float* fsum, const uint8_t* active, int R, int Kp,
                          float l1, float l2, float eps, const KlSegs& sg, int* flag, cudaStream_t st);

}  // namespace cnmf
