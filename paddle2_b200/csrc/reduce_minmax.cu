// MAX / MIN / AMAX / AMIN instantiations
// (kps/reduce_kernel.cu:61-84,100-133, legacy/kps/reduce_max_kernel.cu:22-33).
#include "reduce_launch.cuh"

namespace b200phi {
int B200PHI_LAUNCH_NAME(launch_minmax)(const ReduceArgs& a) {
  const int op = a.plan->op;
  // float8: the reference registers `max` alone for them (kps/reduce_kernel.cu:345-355); the
  // planner has already refused every other op
  if (a.plan->in_dtype == B200PHI_FLOAT8_E4M3FN)
    return launch_reduce_typed<__nv_fp8_e4m3, float, MaxOp<float>>(a);
  if (a.plan->in_dtype == B200PHI_FLOAT8_E5M2)
    return launch_reduce_typed<__nv_fp8_e5m2, float, MaxOp<float>>(a);
  if (op == B200PHI_MAX || op == B200PHI_AMAX) return dispatch_reduce_natural<MaxOp>(a);
  return dispatch_reduce_natural<MinOp>(a);
}
}  // namespace b200phi
