// SUM / MEAN instantiations (kps/reduce_kernel.cu:110-120,185-271 in the reference).
#include "reduce_launch.cuh"

namespace b200phi {
int B200PHI_LAUNCH_NAME(launch_sum)(const ReduceArgs& a) {
  switch (a.plan->acc) {
    case 0:
      return dispatch_reduce_in<float, SumOp>(a);
    case 1:
      return dispatch_reduce_in<double, SumOp>(a);
    case 2:
      return dispatch_reduce_in<long long, SumOp>(a);
    default:
      return B200PHI_UNSUPPORTED;
  }
}
}  // namespace b200phi
