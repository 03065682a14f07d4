// b200phi_map, unary activation family (map_unary_a): the Cuda*Functor forms of
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

int map_unary_a(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->op) {
    case B200PHI_MAP_EXP:
      return float_dtype<B200PHI_MAP_EXP>(d, bp, s);
    case B200PHI_MAP_LOG:
      return float_dtype<B200PHI_MAP_LOG>(d, bp, s);
    case B200PHI_MAP_SQRT:
      return float_dtype<B200PHI_MAP_SQRT>(d, bp, s);
    case B200PHI_MAP_RSQRT:
      return float_dtype<B200PHI_MAP_RSQRT>(d, bp, s);
    case B200PHI_MAP_SQUARE:
      return float_dtype<B200PHI_MAP_SQUARE>(d, bp, s);
    case B200PHI_MAP_RECIPROCAL:
      return float_dtype<B200PHI_MAP_RECIPROCAL>(d, bp, s);
    case B200PHI_MAP_SIN:
      return float_dtype<B200PHI_MAP_SIN>(d, bp, s);
    case B200PHI_MAP_COS:
      return float_dtype<B200PHI_MAP_COS>(d, bp, s);
    case B200PHI_MAP_TANH:
      return float_dtype<B200PHI_MAP_TANH>(d, bp, s);
    case B200PHI_MAP_SIGMOID:
      return float_dtype<B200PHI_MAP_SIGMOID>(d, bp, s);
    case B200PHI_MAP_SILU:
      return float_dtype<B200PHI_MAP_SILU>(d, bp, s);
    case B200PHI_MAP_LEAKY_RELU:
      return float_dtype<B200PHI_MAP_LEAKY_RELU>(d, bp, s);
    case B200PHI_MAP_ELU:
      return float_dtype<B200PHI_MAP_ELU>(d, bp, s);
    default:
      return set_error(B200PHI_INVALID_ARGUMENT, "map: unknown functor");
  }
}

}  // namespace b200phi
