// b200phi_map, unary activation family (map_unary_b): the Cuda*Functor forms of
// funcs/activation_functor.h that gpu/activation_kernel.cu runs through ElementwiseKernel;
// fp16/bf16 compute in fp32 (MPType).
#include "map_kernels.cuh"

namespace b200phi {

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

// abs and relu also take the integer dtypes (gpu/abs_kernel.cu, activation_kernel.cu relu)
template <int kOp>
static int number_dtype(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->in_dtype) {
    case B200PHI_INT32:
      return launch_map_op<kOp, int32_t>(d, bp, s);
    case B200PHI_INT64:
      return launch_map_op<kOp, long long>(d, bp, s);
    default:
      return float_dtype<kOp>(d, bp, s);
  }
}

int map_unary_b(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->op) {
    case B200PHI_MAP_SOFTPLUS:
      return float_dtype<B200PHI_MAP_SOFTPLUS>(d, bp, s);
    case B200PHI_MAP_FLOOR:
      return float_dtype<B200PHI_MAP_FLOOR>(d, bp, s);
    case B200PHI_MAP_CEIL:
      return float_dtype<B200PHI_MAP_CEIL>(d, bp, s);
    case B200PHI_MAP_EXPM1:
      return float_dtype<B200PHI_MAP_EXPM1>(d, bp, s);
    case B200PHI_MAP_LOG1P:
      return float_dtype<B200PHI_MAP_LOG1P>(d, bp, s);
    case B200PHI_MAP_LOG2:
      return float_dtype<B200PHI_MAP_LOG2>(d, bp, s);
    case B200PHI_MAP_LOG10:
      return float_dtype<B200PHI_MAP_LOG10>(d, bp, s);
    case B200PHI_MAP_RELU6:
      return float_dtype<B200PHI_MAP_RELU6>(d, bp, s);
    case B200PHI_MAP_HARD_SIGMOID:
      return float_dtype<B200PHI_MAP_HARD_SIGMOID>(d, bp, s);
    case B200PHI_MAP_SWISH:
      return float_dtype<B200PHI_MAP_SWISH>(d, bp, s);
    case B200PHI_MAP_MISH:
      return float_dtype<B200PHI_MAP_MISH>(d, bp, s);
    case B200PHI_MAP_SOFTSIGN:
      return float_dtype<B200PHI_MAP_SOFTSIGN>(d, bp, s);
    case B200PHI_MAP_LOG_SIGMOID:
      return float_dtype<B200PHI_MAP_LOG_SIGMOID>(d, bp, s);
    case B200PHI_MAP_POW_ABS:
      return float_dtype<B200PHI_MAP_POW_ABS>(d, bp, s);
    case B200PHI_MAP_ABS:
      return number_dtype<B200PHI_MAP_ABS>(d, bp, s);
    case B200PHI_MAP_RELU:
      return number_dtype<B200PHI_MAP_RELU>(d, bp, s);
    default:
      return set_error(B200PHI_INVALID_ARGUMENT, "map: unknown functor");
  }
}

}  // namespace b200phi
