// b200phi_map, compare family: kps/compare_kernel.cu + funcs/compare_functors.h, out BOOL.
#include "map_kernels.cuh"

namespace b200phi {

template <int kOp>
static int compare_dtype(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
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
      return set_error(B200PHI_UNSUPPORTED, "map: dtype not built for compare");
  }
}

int map_compare(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  switch (d->op) {
    case B200PHI_MAP_EQUAL:
      return compare_dtype<B200PHI_MAP_EQUAL>(d, bp, s);
    case B200PHI_MAP_NOT_EQUAL:
      return compare_dtype<B200PHI_MAP_NOT_EQUAL>(d, bp, s);
    case B200PHI_MAP_LESS_THAN:
      return compare_dtype<B200PHI_MAP_LESS_THAN>(d, bp, s);
    case B200PHI_MAP_LESS_EQUAL:
      return compare_dtype<B200PHI_MAP_LESS_EQUAL>(d, bp, s);
    case B200PHI_MAP_GREATER_THAN:
      return compare_dtype<B200PHI_MAP_GREATER_THAN>(d, bp, s);
    default:
      return compare_dtype<B200PHI_MAP_GREATER_EQUAL>(d, bp, s);
  }
}

}  // namespace b200phi
