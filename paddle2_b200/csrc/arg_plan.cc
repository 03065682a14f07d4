// See arg_plan.h.  Host-only; compiled by g++.
#include "arg_plan.h"

#include <algorithm>

namespace b200phi {

static long long a_ceil_div(long long a, long long b) { return (a + b - 1) / b; }
static long long a_round_up(long long a, long long b) { return a_ceil_div(a, b) * b; }
static int a_pow2_ceil(long long v) {
  int p = 1;
  while (p < v && p < (1 << 30)) p <<= 1;
  return p;
}
static int a_fail(std::string* err, int code, const char* msg) {
  if (err) *err = msg;
  return code;
}

// the type the comparison runs in: float for 16-bit floats, int32 for the narrow integers
static size_t arg_val_size(int dtype) {
  switch (dtype) {
    case B200PHI_FLOAT64:
    case B200PHI_INT64:
      return 8;
    case B200PHI_FLOAT32:
    case B200PHI_BFLOAT16:
    case B200PHI_FLOAT16:
    case B200PHI_INT32:
    case B200PHI_INT16:
    case B200PHI_UINT8:
      return 4;
    default:
      return 0;
  }
}

int plan_arg(const b200phi_argreduce_desc* d, const Tunables& t, bool x_aligned16, ArgPlan* p,
             std::string* err) {
  if (!d) return a_fail(err, B200PHI_INVALID_ARGUMENT, "null descriptor");
  *p = ArgPlan();
  if (d->op != B200PHI_ARGMAX && d->op != B200PHI_ARGMIN)
    return a_fail(err, B200PHI_INVALID_ARGUMENT, "unknown arg op");
  if (d->rank < 0 || d->rank > B200PHI_MAX_RANK)
    return a_fail(err, B200PHI_INVALID_ARGUMENT, "rank must be in [0, 9]");
  p->index_dtype = d->index_dtype == B200PHI_UNDEFINED ? B200PHI_INT64 : d->index_dtype;
  if (p->index_dtype != B200PHI_INT32 && p->index_dtype != B200PHI_INT64)
    return a_fail(err, B200PHI_INVALID_ARGUMENT, "argmax/argmin index dtype must be int32 or int64");
  long long numel = 1;
  for (int i = 0; i < d->rank; ++i) {
    if (d->dims[i] < 0) return a_fail(err, B200PHI_INVALID_ARGUMENT, "negative extent");
    numel *= d->dims[i];
  }
  if (numel <= 0)
    return a_fail(err, B200PHI_INVALID_ARGUMENT, "argmin/argmax input numel must > 0, bug got 0");
  if (d->flatten || d->rank == 0) {
    p->pre = 1;
    p->n = numel;
    p->post = 1;
  } else {
    long long axis = d->axis;
    if (axis < -d->rank || axis >= d->rank)
      return a_fail(err, B200PHI_INVALID_ARGUMENT, "argmax/argmin axis out of range");
    if (axis < 0) axis += d->rank;
    p->pre = p->post = 1;
    for (int i = 0; i < axis; ++i) p->pre *= d->dims[i];
    p->n = d->dims[axis];
    for (int i = static_cast<int>(axis) + 1; i < d->rank; ++i) p->post *= d->dims[i];
  }
  if (p->index_dtype == B200PHI_INT32 && p->n > 0x7fffffffLL)
    return a_fail(err, B200PHI_INVALID_ARGUMENT, "int32 indices need the reduced extent <= INT_MAX");
  p->in_dtype = d->in_dtype;
  p->val_size = arg_val_size(d->in_dtype);
  const size_t esize = dtype_size(d->in_dtype);
  if (p->val_size == 0 || esize == 0)
    return a_fail(err, B200PHI_UNSUPPORTED, "argmax/argmin: dtype not built");

  const int VE = static_cast<int>(16 / esize);
  const long long want_ctas = static_cast<long long>(t.sm_count) * t.ctas_per_sm;
  const long long out_numel = p->pre * p->post;
  p->kind = kArgRows;
  p->splits = 1;
  p->grid_x = p->grid_y = 1;

  if (p->post == 1) {
    const long long row_bytes = p->n * static_cast<long long>(esize);
    if (p->pre < static_cast<long long>(t.sm_count) * 2) {
      long long splits = std::min(a_ceil_div(want_ctas, p->pre), row_bytes / 32768);
      splits = std::min<long long>(splits, 65535);
      if (splits > 1) {
        const long long seg = a_round_up(a_ceil_div(p->n, splits), static_cast<long long>(VE) * 32);
        splits = a_ceil_div(p->n, seg);
        if (splits > 1) {
          p->kind = kArgRowSplit;
          p->seg = seg;
          p->splits = static_cast<int>(splits);
          p->aligned = (x_aligned16 && p->n % VE == 0) ? 1 : 0;
          p->grid_x = static_cast<unsigned>(splits);
          p->grid_y = static_cast<unsigned>(p->pre);
        }
      }
    }
  } else {
    p->kind = kArgCol;
    const bool vec_ok = VE > 1 && p->post % VE == 0 && x_aligned16;
    p->vec = vec_ok ? VE : 1;
    const long long post_v = p->post / p->vec;
    p->TX = std::min(256, a_pow2_ceil(post_v));
    const int TY = 256 / p->TX;
    p->tiles = a_ceil_div(post_v, p->TX);
    const long long bxy = p->pre * p->tiles;
    if (bxy > 0x7fffffffLL) return a_fail(err, B200PHI_UNSUPPORTED, "argreduce: output too large");
    const long long max_splits = std::max<long long>(
        1, std::min(p->n / (static_cast<long long>(TY) * 8),
                    p->n * p->TX * p->vec * static_cast<long long>(esize) / (256 << 10)));
    long long splits = pick_splits(bxy, want_ctas, max_splits,  // whole waves (reduce_plan.cc)
                                  static_cast<double>(p->n) * esize / (p->val_size + 8));
    p->rows_per_split = a_round_up(a_ceil_div(p->n, splits), TY);
    if (p->rows_per_split > 0x7fffffffLL)
      return a_fail(err, B200PHI_UNSUPPORTED, "argreduce: reduced extent too long for the column plan");
    splits = a_ceil_div(p->n, p->rows_per_split);
    p->splits = static_cast<int>(splits);
    p->grid_x = static_cast<unsigned>(bxy);
    p->grid_y = static_cast<unsigned>(splits);
  }
  if (p->splits > 1) {
    const size_t parts = static_cast<size_t>(p->splits) * static_cast<size_t>(out_numel);
    p->ws_val_bytes = static_cast<size_t>(a_round_up(static_cast<long long>(parts * p->val_size), 256));
    p->ws_idx_bytes = static_cast<size_t>(a_round_up(static_cast<long long>(parts * 8), 256));
    p->ws_ticket_bytes = static_cast<size_t>(p->kind == kArgRowSplit ? p->pre : p->grid_x) * 4;
    p->ws_total_bytes = p->ws_val_bytes + p->ws_idx_bytes + p->ws_ticket_bytes;
  }
  return B200PHI_OK;
}

}  // namespace b200phi
