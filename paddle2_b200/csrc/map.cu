// b200phi_map: validation + family dispatch.  The functors are in map_functors.cuh, the
// kernels in map_kernels.cuh, one translation unit per family (map_*.cu).
#include "map_kernels.cuh"

namespace b200phi {

static bool is_float_dtype(int dt) {
  return dt == B200PHI_FLOAT32 || dt == B200PHI_FLOAT64 || dt == B200PHI_FLOAT16 ||
         dt == B200PHI_BFLOAT16;
}

static int map_expected_arity(int op) {
  switch (op) {
    case B200PHI_MAP_LOGICAL_NOT:
    case B200PHI_MAP_BITWISE_NOT:
      return 1;
    case B200PHI_MAP_EQUAL_MUL:
    case B200PHI_MAP_MAX_GRAD_X:
    case B200PHI_MAP_MAX_GRAD_Y:
    case B200PHI_MAP_MIN_GRAD_X:
    case B200PHI_MAP_MIN_GRAD_Y:
    case B200PHI_MAP_DIV_GRAD_Y:
      return 3;
    default:
      return op >= B200PHI_MAP_EXP ? 1 : 2;
  }
}

int run_map(const b200phi_map_desc* d, cudaStream_t stream) {
  if (!d) return set_error(B200PHI_INVALID_ARGUMENT, "null descriptor");
  const int op = d->op;
  const bool known = (op >= B200PHI_MAP_EQUAL && op <= B200PHI_MAP_GREATER_EQUAL) ||
                     (op >= B200PHI_MAP_LOGICAL_AND && op <= B200PHI_MAP_LOGICAL_NOT) ||
                     (op >= B200PHI_MAP_BITWISE_AND && op <= B200PHI_MAP_RIGHT_SHIFT_LOGIC) ||
                     (op >= B200PHI_MAP_EQUAL_MUL && op <= B200PHI_MAP_DIV_GRAD_Y) ||
                     op == B200PHI_MAP_EQUAL_T || op == B200PHI_MAP_SUB_EXP || op == B200PHI_MAP_HEAVISIDE ||
                     (op >= B200PHI_MAP_REMAINDER && op <= B200PHI_MAP_FMIN) ||
                     (op >= B200PHI_MAP_EXP && op <= B200PHI_MAP_RELU);
  if (!known) return set_error(B200PHI_INVALID_ARGUMENT, "map: unknown functor");
  if (d->arity != map_expected_arity(op))
    return set_error(B200PHI_INVALID_ARGUMENT,
                     "map: the functor takes " + std::to_string(map_expected_arity(op)) +
                         " operand(s), got " + std::to_string(d->arity));
  const int implied = map_out_is_bool(op) ? B200PHI_BOOL : d->in_dtype;
  if (d->out_dtype != B200PHI_UNDEFINED && d->out_dtype != implied)
    return set_error(B200PHI_INVALID_ARGUMENT, "map: out_dtype does not match what the functor writes");

  // same placement / broadcast rules as b200phi_elementwise
  b200phi_elementwise_desc e = {};
  e.dtype = d->in_dtype;
  e.arity = d->arity;
  e.axis = d->axis;
  for (int i = 0; i < d->arity && i < B200PHI_MAX_ARITY; ++i) e.in[i] = d->in[i];
  e.out_rank = d->out_rank;
  for (int k = 0; k < B200PHI_MAX_RANK; ++k) e.out_dims[k] = d->out_dims[k];
  BcastPlan bp;
  std::string err;
  int rc = plan_broadcast(&e, &bp, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  if (bp.out_numel == 0) return B200PHI_OK;
  if (!d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null output pointer");
  for (int i = 0; i < d->arity; ++i)
    if (!d->in[i].data) return set_error(B200PHI_INVALID_ARGUMENT, "null input pointer");

  if (op <= B200PHI_MAP_GREATER_EQUAL) return map_compare(d, bp, stream);
  if (op <= B200PHI_MAP_LOGICAL_NOT) return map_logical(d, bp, stream);
  if (op <= B200PHI_MAP_RIGHT_SHIFT_LOGIC) return map_bitwise(d, bp, stream);
  if (op == B200PHI_MAP_HEAVISIDE || (op >= B200PHI_MAP_REMAINDER && op <= B200PHI_MAP_FMIN))
    return map_binary(d, bp, stream);
  if (op <= B200PHI_MAP_SUB_EXP) {
    if (op == B200PHI_MAP_SUB_EXP && !is_float_dtype(d->in_dtype))
      return set_error(B200PHI_UNSUPPORTED, "map: this functor needs a floating-point dtype");
    return map_ternary(d, bp, stream);
  }
  return map_unary(d, bp, stream);
}

int map_unary_a(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);
int map_unary_b(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);
int map_unary(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s) {
  return d->op <= B200PHI_MAP_ELU ? map_unary_a(d, bp, s) : map_unary_b(d, bp, s);
}

}  // namespace b200phi

extern "C" int b200phi_map(const b200phi_map_desc* d, void* stream) {
  return b200phi::run_map(d, static_cast<cudaStream_t>(stream));
}
