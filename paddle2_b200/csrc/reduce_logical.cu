// ANY / ALL instantiations (kps/reduce_kernel.cu:48-58,87-97): the cast-to-bool
// pre-pass of the reference is the `!= 0` in the loader.
#include "reduce_launch.cuh"

namespace b200phi {
int B200PHI_LAUNCH_NAME(launch_logical)(const ReduceArgs& a) {
  if (a.plan->op == B200PHI_ANY) return dispatch_reduce_in<int, AnyOp>(a);
  return dispatch_reduce_in<int, AllOp>(a);
}
}  // namespace b200phi
