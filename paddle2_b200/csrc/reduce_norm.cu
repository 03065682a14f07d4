// Norm-type reductions: the base operators over a transformed element
// (gpu/p_norm_kernel.cu:76-95, gpu/squared_l2_norm_kernel.cu:24-34 in the reference).
// Floating-point inputs only; fp16/bf16 accumulate in fp32 (MPType).
#include "reduce_launch.cuh"

namespace b200phi {

template <template <typename> class OpT>
static int dispatch_norm(const ReduceArgs& a) {
  switch (a.plan->in_dtype) {
    case B200PHI_FLOAT32:
      return launch_reduce_typed<float, float, OpT<float>>(a);
    case B200PHI_FLOAT64:
      return launch_reduce_typed<double, double, OpT<double>>(a);
    case B200PHI_FLOAT16:
      return launch_reduce_typed<__half, float, OpT<float>>(a);
    case B200PHI_BFLOAT16:
      return launch_reduce_typed<__nv_bfloat16, float, OpT<float>>(a);
    default:
      return B200PHI_UNSUPPORTED;
  }
}

int B200PHI_LAUNCH_NAME(launch_norm)(const ReduceArgs& a) {
  switch (a.plan->op) {
    case B200PHI_SUM_ABS:
      return dispatch_norm<SumAbsOp>(a);
    case B200PHI_SUM_SQUARE:
      return dispatch_norm<SumSquareOp>(a);
    case B200PHI_L2_NORM:
      return dispatch_norm<L2NormOp>(a);
    case B200PHI_MAX_ABS:
      return dispatch_norm<MaxAbsOp>(a);
    case B200PHI_MIN_ABS:
      return dispatch_norm<MinAbsOp>(a);
    case B200PHI_COUNT_NONZERO:
      return dispatch_norm<CountNonzeroOp>(a);
    default:
      return B200PHI_UNSUPPORTED;
  }
}

}  // namespace b200phi
