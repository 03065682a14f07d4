// Host-side planning for the n-ary broadcast / elementwise engine.  Pure C++.
//
// Re-states what the reference spreads over
//   funcs::GetBroadcastDimsArrays     common_shape.h:35-92      (axis alignment, dim check)
//   BroadcastDimsSimplifier           dims_simplifier.h:25-224  (rank extension + merging)
//   BroadcastTypeClassifier           broadcast_function.h:34-113
// Differences by design: dims stay in outer->inner order, strides are 64-bit
// element strides (0 on a broadcast dim), and the plan keeps the innermost merged
// dim apart so kernels never do a div/mod per element.
#pragma once

#include <stdint.h>

#include <string>

#include "../../include/b200phi.h"

namespace b200phi {

struct BcastPlan {
  int arity;
  int ndim;  // merged dims, outer -> inner, >= 1
  long long extent[B200PHI_MAX_RANK];
  long long stride[B200PHI_MAX_ARITY][B200PHI_MAX_RANK];
  long long out_numel;
  // derived: rows x inner view
  long long inner, rows;
  int use_rows_kernel;  // 1: per-row base offsets + vector inner loop; 0: per-element unravel
};

int plan_broadcast(const b200phi_elementwise_desc* d, BcastPlan* p, std::string* err);

}  // namespace b200phi
