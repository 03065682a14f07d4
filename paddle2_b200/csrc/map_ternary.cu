// b200phi_map, backward functors: three operands (x, out|y, dout) through the broadcast
// engine in one launch where the reference runs Equal then Multiply
// (gpu/reduce_kernel.cu:74-119) or one ternary functor (funcs/elementwise_functor.h:173-179,
// 478-531), plus the two-operand helpers EQUAL_T and SUB_EXP.
#include "map_kernels.cuh"

namespace b200phi {

template <int kOp>
static int grad_dtype(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->in_dtype) {
    case B200PHI_INT32:
      return launch_map_op<kOp, int32_t>(d, bp, s);
    case B200PHI_INT64:
      return launch_map_op<kOp, long long>(d, bp, s);
    case B200PHI_FLOAT32:
      return launch_map_op<kOp, float>(d, bp, s);
    case B200PHI_FLOAT64:
      return launch_map_op<kOp, double>(d, bp, s);
    case B200PHI_FLOAT16:
      return launch_map_op<kOp, __half>(d, bp, s);
    case B200PHI_BFLOAT16:
      return launch_map_op<kOp, __nv_bfloat16>(d, bp, s);
    default:
      return set_error(B200PHI_UNSUPPORTED, "map: dtype not built for this functor");
  }
}

template <int kOp>
static int float_dtype(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->in_dtype) {
    case B200PHI_FLOAT32:
      return launch_map_op<kOp, float>(d, bp, s);
    case B200PHI_FLOAT64:
      return launch_map_op<kOp, double>(d, bp, s);
    case B200PHI_FLOAT16:
      return launch_map_op<kOp, __half>(d, bp, s);
    case B200PHI_BFLOAT16:
      return launch_map_op<kOp, __nv_bfloat16>(d, bp, s);
    default:
      return set_error(B200PHI_UNSUPPORTED, "map: this functor needs a floating-point dtype");
  }
}

int map_ternary(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->op) {
    case B200PHI_MAP_EQUAL_MUL:
      return grad_dtype<B200PHI_MAP_EQUAL_MUL>(d, bp, s);
    case B200PHI_MAP_MAX_GRAD_X:
      return grad_dtype<B200PHI_MAP_MAX_GRAD_X>(d, bp, s);
    case B200PHI_MAP_MAX_GRAD_Y:
      return grad_dtype<B200PHI_MAP_MAX_GRAD_Y>(d, bp, s);
    case B200PHI_MAP_MIN_GRAD_X:
      return grad_dtype<B200PHI_MAP_MIN_GRAD_X>(d, bp, s);
    case B200PHI_MAP_MIN_GRAD_Y:
      return grad_dtype<B200PHI_MAP_MIN_GRAD_Y>(d, bp, s);
    case B200PHI_MAP_DIV_GRAD_Y:
      return grad_dtype<B200PHI_MAP_DIV_GRAD_Y>(d, bp, s);
    case B200PHI_MAP_EQUAL_T:
      return grad_dtype<B200PHI_MAP_EQUAL_T>(d, bp, s);
    default:
      return float_dtype<B200PHI_MAP_SUB_EXP>(d, bp, s);
  }
}

}  // namespace b200phi
