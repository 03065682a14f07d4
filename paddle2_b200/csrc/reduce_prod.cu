// PROD instantiations (kps/reduce_kernel.cu:35-45 in the reference).
#include "reduce_launch.cuh"

namespace b200phi {
int B200PHI_LAUNCH_NAME(launch_prod)(const ReduceArgs& a) { return dispatch_reduce_natural<ProdOp>(a); }
}  // namespace b200phi
