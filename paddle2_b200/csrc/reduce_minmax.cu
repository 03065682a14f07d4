// MAX / MIN / AMAX / AMIN instantiations
// (kps/reduce_kernel.cu:61-84,100-133, legacy/kps/reduce_max_kernel.cu:22-33).
#include "reduce_launch.cuh"

namespace b200phi {
int B200PHI_LAUNCH_NAME(launch_minmax)(const ReduceArgs& a) {
  const int op = a.plan->op;
  if (op == B200PHI_MAX || op == B200PHI_AMAX) return dispatch_reduce_natural<MaxOp>(a);
  return dispatch_reduce_natural<MinOp>(a);
}
}  // namespace b200phi
