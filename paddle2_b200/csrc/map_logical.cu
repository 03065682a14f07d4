// b200phi_map, logical and bitwise families: kps/logical_kernel.cu + funcs/logical_functor.h
// (any dtype by truthiness, out BOOL); kps/bitwise_kernel.cu + funcs/bitwise_functors.h
// (bool and integer dtypes, out = in dtype).
#include "map_kernels.cuh"

namespace b200phi {

template <int kOp>
static int logical_dtype(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->in_dtype) {
    case B200PHI_BOOL:
      return launch_map_op<kOp, Bool8>(d, bp, s);
    case B200PHI_UINT8:
      return launch_map_op<kOp, uint8_t>(d, bp, s);
    case B200PHI_INT8:
      return launch_map_op<kOp, int8_t>(d, bp, s);
    case B200PHI_INT16:
      return launch_map_op<kOp, int16_t>(d, bp, s);
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
      return set_error(B200PHI_UNSUPPORTED, "map: dtype not built for logical ops");
  }
}

int map_logical(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->op) {
    case B200PHI_MAP_LOGICAL_AND:
      return logical_dtype<B200PHI_MAP_LOGICAL_AND>(d, bp, s);
    case B200PHI_MAP_LOGICAL_OR:
      return logical_dtype<B200PHI_MAP_LOGICAL_OR>(d, bp, s);
    case B200PHI_MAP_LOGICAL_XOR:
      return logical_dtype<B200PHI_MAP_LOGICAL_XOR>(d, bp, s);
    default:
      return logical_dtype<B200PHI_MAP_LOGICAL_NOT>(d, bp, s);
  }
}

template <int kOp>
static int bitwise_dtype(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->in_dtype) {
    case B200PHI_BOOL:
      return launch_map_op<kOp, Bool8>(d, bp, s);
    case B200PHI_UINT8:
      return launch_map_op<kOp, uint8_t>(d, bp, s);
    case B200PHI_INT8:
      return launch_map_op<kOp, int8_t>(d, bp, s);
    case B200PHI_INT16:
      return launch_map_op<kOp, int16_t>(d, bp, s);
    case B200PHI_INT32:
      return launch_map_op<kOp, int32_t>(d, bp, s);
    case B200PHI_INT64:
      return launch_map_op<kOp, long long>(d, bp, s);
    default:
      return set_error(B200PHI_UNSUPPORTED, "map: bitwise ops take bool and integer dtypes");
  }
}

template <int kOp>
static int shift_dtype(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->in_dtype) {
    case B200PHI_UINT8:
      return launch_map_op<kOp, uint8_t>(d, bp, s);
    case B200PHI_INT8:
      return launch_map_op<kOp, int8_t>(d, bp, s);
    case B200PHI_INT16:
      return launch_map_op<kOp, int16_t>(d, bp, s);
    case B200PHI_INT32:
      return launch_map_op<kOp, int32_t>(d, bp, s);
    case B200PHI_INT64:
      return launch_map_op<kOp, long long>(d, bp, s);
    default:
      return set_error(B200PHI_UNSUPPORTED, "map: shifts take uint8 / int8 / int16 / int32 / int64");
  }
}

int map_bitwise(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->op) {
    case B200PHI_MAP_LEFT_SHIFT_ARITHMETIC:
      return shift_dtype<B200PHI_MAP_LEFT_SHIFT_ARITHMETIC>(d, bp, s);
    case B200PHI_MAP_LEFT_SHIFT_LOGIC:
      return shift_dtype<B200PHI_MAP_LEFT_SHIFT_LOGIC>(d, bp, s);
    case B200PHI_MAP_RIGHT_SHIFT_ARITHMETIC:
      return shift_dtype<B200PHI_MAP_RIGHT_SHIFT_ARITHMETIC>(d, bp, s);
    case B200PHI_MAP_RIGHT_SHIFT_LOGIC:
      return shift_dtype<B200PHI_MAP_RIGHT_SHIFT_LOGIC>(d, bp, s);
    case B200PHI_MAP_BITWISE_AND:
      return bitwise_dtype<B200PHI_MAP_BITWISE_AND>(d, bp, s);
    case B200PHI_MAP_BITWISE_OR:
      return bitwise_dtype<B200PHI_MAP_BITWISE_OR>(d, bp, s);
    case B200PHI_MAP_BITWISE_XOR:
      return bitwise_dtype<B200PHI_MAP_BITWISE_XOR>(d, bp, s);
    default:
      return bitwise_dtype<B200PHI_MAP_BITWISE_NOT>(d, bp, s);
  }
}

}  // namespace b200phi
