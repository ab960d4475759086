// Segment tables of a CSR matrix, shared by the two sparse-KL kernel families (kl_sparse.cu builds them on the
// device; kl_slab.cu consumes them too).  See the comment above KlSegs in kl_sparse.cu for what a segment is.
#pragma once
#include "common.cuh"

namespace cnmf {

struct KlSegs {
  const int* vseg_ptr;      // [V + 1] first segment of each virtual row, V = nb * M
  const int* row_slot_ptr;  // [M + 1] nb == 1: first partial slot of each cut row (rows of one segment own none)
  const int* seg_vrow;      // [nseg]
  const int* seg_beg;       // [nseg] first entry of the segment
  const int* vend;          // [V] end of each virtual row's entries (nb == 1: row_ptr + 1)
  const int* multi_rows;    // nb == 1: [n_multi] rows that were cut (any order)
  const int* n_multi;       // [1]
  float* part;              // [slots x R*KP]; nb > 1 or all_partial: one slot per segment
  int seg_len;              // entries per segment
  int nb;                   // column blocks
};

struct KlSegLayout {
  int64_t off_vsp, off_rslot, off_svrow, off_sbeg, off_multi, off_nmulti, off_vbeg, off_vend;
  int seg, block, nb;
  int64_t V;
  bool all_partial;         // every row leaves per-segment partials and is finished by a separate pass (nb == 1 too)
};

// bounds of the table sizes (host side, from M, nnz and the segment length only)
static inline int64_t kl_nseg_max(int64_t V, int64_t nnz, int seg) { return V + nnz / seg; }

// explicit segment length / column block (the slab kernels choose them from the shared-memory size)
int64_t kl_seg_layout_explicit(int64_t M, int64_t n_cols, int64_t nnz, int seg, int block, int64_t off, KlSegLayout* o);

// launch == false: only bind `out` to the tables a previous call left in the workspace (a warm problem)
int kl_build_segments(const int* row_ptr, const int* col_idx, int64_t M, char* ws, const KlSegLayout& o, float* part,
                      KlSegs* out, cudaStream_t st, bool launch = true);

}  // namespace cnmf
