// See bcast_plan.h.  Host-only.
#include "bcast_plan.h"

#include <stdio.h>

namespace b200phi {

static int fail(std::string* err, int code, const std::string& msg) {
  if (err) *err = msg;
  return code;
}

template <typename I>
static std::string dims_str(const I* d, int n) {
  std::string s = "[";
  for (int i = 0; i < n; ++i) {
    s += std::to_string(d[i]);
    if (i + 1 < n) s += ", ";
  }
  return s + "]";
}

int plan_broadcast(const b200phi_elementwise_desc* d, BcastPlan* p, std::string* err) {
  if (!d) return fail(err, B200PHI_INVALID_ARGUMENT, "null descriptor");
  *p = BcastPlan();
  if (d->arity < 1 || d->arity > B200PHI_MAX_ARITY)
    return fail(err, B200PHI_INVALID_ARGUMENT, "arity must be 1, 2 or 3");
  if (d->out_rank < 0 || d->out_rank > B200PHI_MAX_RANK)
    return fail(err, B200PHI_INVALID_ARGUMENT, "out rank must be in [0, 9]");
  const int R = d->out_rank;
  p->arity = d->arity;

  // ---- place every input inside the output rank ----------------------------------
  long long in_dims[B200PHI_MAX_ARITY][B200PHI_MAX_RANK];
  for (int i = 0; i < d->arity; ++i) {
    const b200phi_tensor_view& t = d->in[i];
    if (t.rank < 0 || t.rank > R) {
      return fail(err, B200PHI_INVALID_ARGUMENT,
                  "input rank exceeds the output rank (InferMeta fixes out_rank = max rank)");
    }
    int offset = R - t.rank;  // numpy right alignment
    if (i < 2 && d->arity >= 2 && d->axis >= 0) {
      // legacy `axis`: the lower-rank operand of (in0, in1) starts at dim `axis`
      const int r0 = d->in[0].rank, r1 = d->in[1].rank;
      const bool lower = (i == 0) ? (r0 < r1) : (r1 < r0);
      if (r0 == r1 && d->axis != 0)
        return fail(err, B200PHI_INVALID_ARGUMENT,
                    "axis should be -1 or 0 while the dimension of tensor X is equal to the "
                    "dimension of tensor Y");
      if (lower) {
        offset = d->axis;
        if (offset + t.rank > R)
          return fail(err, B200PHI_INVALID_ARGUMENT, "Axis should be less than or equal to max rank");
      }
    }
    for (int k = 0; k < R; ++k) in_dims[i][k] = 1;
    for (int k = 0; k < t.rank; ++k) in_dims[i][offset + k] = t.dims[k];
    for (int k = 0; k < R; ++k) {
      const long long a = in_dims[i][k], o = d->out_dims[k];
      if (a < 0 || o < 0) return fail(err, B200PHI_INVALID_ARGUMENT, "negative extent");
      if (a != o && a != 1) {
        return fail(err, B200PHI_INVALID_ARGUMENT,
                    "Broadcast dimension mismatch. Operands could not be broadcast together with "
                    "the shape of input " + std::to_string(i) + " = " + dims_str(in_dims[i], R) +
                    " and the shape of Out = " + dims_str(d->out_dims, R) + ".");
      }
    }
  }

  // ---- strides (0 on broadcast dims), drop extent-1 dims, merge ---------------------
  long long ext[B200PHI_MAX_RANK];
  long long st[B200PHI_MAX_ARITY][B200PHI_MAX_RANK];
  long long run[B200PHI_MAX_ARITY];
  for (int i = 0; i < d->arity; ++i) run[i] = 1;
  // raw strides, inner -> outer
  long long raw[B200PHI_MAX_ARITY][B200PHI_MAX_RANK];
  for (int k = R - 1; k >= 0; --k) {
    for (int i = 0; i < d->arity; ++i) {
      raw[i][k] = (in_dims[i][k] == 1 && d->out_dims[k] != 1) ? 0 : run[i];
      run[i] *= in_dims[i][k];
    }
  }
  int m = 0;
  long long numel = 1;
  for (int k = 0; k < R; ++k) {
    const long long e = d->out_dims[k];
    numel *= e;
    if (e == 1) continue;
    bool mergeable = m > 0;
    if (mergeable) {
      for (int i = 0; i < d->arity; ++i)
        if (st[i][m - 1] != raw[i][k] * e) mergeable = false;
    }
    if (mergeable) {
      ext[m - 1] *= e;
      for (int i = 0; i < d->arity; ++i) st[i][m - 1] = raw[i][k];
    } else {
      ext[m] = e;
      for (int i = 0; i < d->arity; ++i) st[i][m] = raw[i][k];
      ++m;
    }
  }
  if (m == 0) {
    ext[0] = 1;
    for (int i = 0; i < d->arity; ++i) st[i][0] = 0;
    m = 1;
  }
  p->ndim = m;
  p->out_numel = numel;
  for (int k = 0; k < m; ++k) {
    p->extent[k] = ext[k];
    for (int i = 0; i < d->arity; ++i) p->stride[i][k] = st[i][k];
  }
  p->inner = ext[m - 1];
  p->rows = numel / (p->inner ? p->inner : 1);
  // the row kernel wants a useful amount of contiguous work per row
  p->use_rows_kernel = (m == 1 || p->inner >= 32) ? 1 : 0;
  return B200PHI_OK;
}

}  // namespace b200phi
