// b200phi_map, the arithmetic binary functors that kps/elementwise_kernel.cu:183-293 and
// legacy/kps/elementwise_kernel.cu:139-165 register next to add .. minimum: remainder,
// floor_divide, elementwise_pow, copysign, fmax, fmin, heaviside.  Same broadcast engine, out
// dtype = in dtype; dtype lists as the reference registers them.
#include "map_kernels.cuh"

namespace b200phi {

// kMask bits: 1 floats (fp32, fp64, fp16, bf16), 2 int32 / int64, 4 the narrow integers
// (uint8, int8, int16), 8 bool
template <int kOp, int kMask>
static int binary_dtype(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->in_dtype) {
    case B200PHI_FLOAT32:
      if constexpr (kMask & 1) return launch_map_op<kOp, float>(d, bp, s);
      break;
    case B200PHI_FLOAT64:
      if constexpr (kMask & 1) return launch_map_op<kOp, double>(d, bp, s);
      break;
    case B200PHI_FLOAT16:
      if constexpr (kMask & 1) return launch_map_op<kOp, __half>(d, bp, s);
      break;
    case B200PHI_BFLOAT16:
      if constexpr (kMask & 1) return launch_map_op<kOp, __nv_bfloat16>(d, bp, s);
      break;
    case B200PHI_INT32:
      if constexpr (kMask & 2) return launch_map_op<kOp, int32_t>(d, bp, s);
      break;
    case B200PHI_INT64:
      if constexpr (kMask & 2) return launch_map_op<kOp, long long>(d, bp, s);
      break;
    case B200PHI_UINT8:
      if constexpr (kMask & 4) return launch_map_op<kOp, uint8_t>(d, bp, s);
      break;
    case B200PHI_INT8:
      if constexpr (kMask & 4) return launch_map_op<kOp, int8_t>(d, bp, s);
      break;
    case B200PHI_INT16:
      if constexpr (kMask & 4) return launch_map_op<kOp, int16_t>(d, bp, s);
      break;
    case B200PHI_BOOL:
      if constexpr (kMask & 8) return launch_map_op<kOp, Bool8>(d, bp, s);
      break;
    default:
      break;
  }
  return set_error(B200PHI_UNSUPPORTED, "map: dtype not built for this functor");
}

int map_binary(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->op) {
    case B200PHI_MAP_REMAINDER:  // float, double, int, int64, float16, bfloat16
      return binary_dtype<B200PHI_MAP_REMAINDER, 3>(d, bp, s);
    case B200PHI_MAP_FLOOR_DIVIDE:  // + uint8, int8, int16
      return binary_dtype<B200PHI_MAP_FLOOR_DIVIDE, 7>(d, bp, s);
    case B200PHI_MAP_POW:
      return binary_dtype<B200PHI_MAP_POW, 3>(d, bp, s);
    case B200PHI_MAP_COPYSIGN:  // every dtype, bool included
      return binary_dtype<B200PHI_MAP_COPYSIGN, 15>(d, bp, s);
    case B200PHI_MAP_FMAX:
      return binary_dtype<B200PHI_MAP_FMAX, 3>(d, bp, s);
    case B200PHI_MAP_FMIN:
      return binary_dtype<B200PHI_MAP_FMIN, 3>(d, bp, s);
    default:
      return binary_dtype<B200PHI_MAP_HEAVISIDE, 3>(d, bp, s);
  }
}

}  // namespace b200phi
