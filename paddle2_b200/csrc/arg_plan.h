// Host-side planning for argmax / argmin: the (pre, n, post) view of
// ComputeFullArg (paddle/phi/kernels/gpu/arg_min_max_kernel.cu:85-145) and the choice between
// the row kernels, the split row kernel (flatten / few long rows) and the column kernel
// (post > 1), with the split of the reduced extent across thread blocks.  Pure C++.
#pragma once

#include <stddef.h>

#include <string>

#include "../../include/b200phi.h"
#include "reduce_plan.h"

namespace b200phi {

enum ArgKind {
  kArgRows = 0,      // post == 1, enough rows: one group of threads per row, no scratch
  kArgRowSplit = 1,  // post == 1, few long rows (flatten): the row is cut across CTAs + ticket finalise
  kArgCol = 2        // post > 1: 16-byte loads along post, n split over grid.y + ticket finalise
};

struct ArgPlan {
  int kind;
  int in_dtype, index_dtype;
  long long pre, n, post;
  // kArgRowSplit
  long long seg;
  int aligned;  // every segment starts 16-byte aligned and is a whole number of vectors
  // kArgCol
  int vec, TX;
  long long tiles, rows_per_split;
  // launch
  unsigned grid_x, grid_y;
  int splits;
  // workspace: [values: splits * out_numel * val_size][indices: ... * 8][tickets], each 256-aligned
  size_t val_size;
  size_t ws_val_bytes, ws_idx_bytes, ws_ticket_bytes, ws_total_bytes;
};

// Validates the descriptor (messages as the reference's) and fills *p.  `x_aligned16`: whether
// d->x sits on a 16-byte boundary (the workspace query passes both and takes the larger answer).
int plan_arg(const b200phi_argreduce_desc* d, const Tunables& t, bool x_aligned16, ArgPlan* p,
             std::string* err);

}  // namespace b200phi
