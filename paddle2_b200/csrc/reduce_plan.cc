// See reduce_plan.h.  Host-only; compiled by g++ (unit tests) and nvcc (library).
#include "reduce_plan.h"

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

namespace b200phi {

size_t dtype_size(int dtype) {
  switch (dtype) {
    case B200PHI_BOOL:
    case B200PHI_UINT8:
    case B200PHI_INT8:
      return 1;
    case B200PHI_INT16:
    case B200PHI_FLOAT16:
    case B200PHI_BFLOAT16:
      return 2;
    case B200PHI_INT32:
    case B200PHI_FLOAT32:
      return 4;
    case B200PHI_INT64:
    case B200PHI_FLOAT64:
    case B200PHI_COMPLEX64:
      return 8;
    case B200PHI_COMPLEX128:
      return 16;
    case B200PHI_FLOAT8_E4M3FN:
    case B200PHI_FLOAT8_E5M2:
      return 1;
    default:
      return 0;
  }
}

bool dtype_known(int dtype) { return dtype_size(dtype) != 0; }

bool dtype_is_float(int dtype) {
  return dtype == B200PHI_FLOAT32 || dtype == B200PHI_FLOAT64 || dtype == B200PHI_FLOAT16 ||
         dtype == B200PHI_BFLOAT16;
}
static bool dtype_is_complex(int dtype) { return dtype == B200PHI_COMPLEX64 || dtype == B200PHI_COMPLEX128; }
static bool dtype_is_fp8(int dtype) { return dtype == B200PHI_FLOAT8_E4M3FN || dtype == B200PHI_FLOAT8_E5M2; }
static bool dtype_is_int(int dtype) {
  return dtype == B200PHI_UINT8 || dtype == B200PHI_INT8 || dtype == B200PHI_INT16 || dtype == B200PHI_INT32 ||
         dtype == B200PHI_INT64;
}

static long long env_ll(const char* name, long long dflt) {
  const char* s = getenv(name);
  if (!s || !*s) return dflt;
  return atoll(s);
}

// The environment is read ONCE per process (first use); a knob changed later is not seen.
static Tunables read_env_tunables() {
  Tunables t;
  t.sm_count = 148;
  t.row_bytes_per_thread = env_ll("B200PHI_ROW_BYTES_PER_THREAD", 64);
  t.ctas_per_sm = static_cast<int>(env_ll("B200PHI_CTAS_PER_SM", 4));  // measured: profiles/r01_microbench.md
  t.force_kind = static_cast<int>(env_ll("B200PHI_FORCE_KIND", -1));
  t.allow_narrow_out = 0;
  t.no_two_stage = static_cast<int>(env_ll("B200PHI_NO_TWO_STAGE", 0));
  t.no_row_batch = static_cast<int>(env_ll("B200PHI_NO_ROW_BATCH", 0));
  t.row_batch_waves = static_cast<int>(env_ll("B200PHI_ROW_BATCH_WAVES", 8));
  t.no_tiled = static_cast<int>(env_ll("B200PHI_NO_TILED", 0));
  t.no_row_peel = static_cast<int>(env_ll("B200PHI_NO_ROW_PEEL", 0));
  t.no_row_staged = static_cast<int>(env_ll("B200PHI_NO_ROW_STAGED", 0));
  t.p2p_timeout_ms = env_ll("B200PHI_P2P_TIMEOUT_MS", 4000);
  return t;
}

long long staged_max_row_bytes() {
  static const long long v = env_ll("B200PHI_STAGED_MAX_BYTES", 0);
  return v;
}

long long staged_min_row_bytes() {
  static const long long v = env_ll("B200PHI_STAGED_MIN_BYTES", 0);
  return v;
}

int staged_aligned_mode() {
  static const int v = static_cast<int>(env_ll("B200PHI_STAGED_ALIGNED", 1));
  return v;
}

Tunables default_tunables(int sm_count) {
  static const Tunables env = read_env_tunables();
  Tunables t = env;
  t.sm_count = sm_count > 0 ? sm_count : 148;
  return t;
}

static long long ceil_div(long long a, long long b) { return (a + b - 1) / b; }
static long long round_up(long long a, long long b) { return ceil_div(a, b) * b; }
static int pow2_ceil(long long v) {
  int p = 1;
  while (p < v && p < (1 << 30)) p <<= 1;
  return p;
}

// How many ways to cut the reduced extent when `bxy` thread blocks already come from the
// outputs.  The blocks run in waves of `slots` and every block does 1/s of a column's work, so
// compute time ~ ceil(bxy * s / slots) / s; a split also writes and re-reads one partial per
// output per split, `ratio` = (input bytes per output) / (partial bytes per output) says how
// much that weighs.  The smallest s within 0.5 % of the best cost: 392 tiles on 592 slots run as
// 3 splits in two full waves (0.67), not as 1 or 2 (one and a third waves: 1.0); 16384 tiles of
// a 64-row reduce are never split (28 waves either way, and the partials would be a quarter of
// the input).
long long pick_splits(long long bxy, long long slots, long long max_splits, double ratio) {
  if (bxy < 1) bxy = 1;
  if (ratio < 1.0) ratio = 1.0;
  long long hi = std::min<long long>(max_splits, ceil_div(slots, bxy) * 4 + 4);
  hi = std::max<long long>(1, std::min<long long>(hi, 65535));
  auto cost = [&](long long s) {
    const double waves = static_cast<double>(ceil_div(bxy * s, slots)) / static_cast<double>(s);
    // (5 % flat: the ticket, the finalise and the memset are not free either)
    return waves * (1.0 + (s > 1 ? 0.05 + 2.0 * static_cast<double>(s) / ratio : 0.0));
  };
  double best = 1e300;
  for (long long s = 1; s <= hi; ++s) best = std::min(best, cost(s));
  for (long long s = 1; s <= hi; ++s)
    if (cost(s) <= best * 1.005) return s;
  return 1;
}

static int fail(std::string* err, int code, const std::string& msg) {
  if (err) *err = msg;
  return code;
}

void two_stage_descs(const ReducePlan& p, const b200phi_reduce_desc* d, void* tmp,
                     b200phi_reduce_desc* d1, b200phi_reduce_desc* d2) {
  // stage 1: the merged tensor, reduced over the peeled dim alone
  *d1 = b200phi_reduce_desc();
  d1->op = p.ts_op1;
  d1->in_dtype = p.in_dtype;
  d1->out_dtype = p.ts_tmp_dtype;
  d1->rank = p.ts_rank;
  for (int i = 0; i < p.ts_rank; ++i) d1->dims[i] = p.ts_dims[i];
  d1->naxes = 1;
  d1->axes[0] = p.ts_peel;
  d1->x = d->x;
  d1->out = tmp;
  // stage 2: what is left of it, reduced over the remaining reduced dims
  *d2 = b200phi_reduce_desc();
  d2->op = p.ts_op2;
  d2->in_dtype = p.ts_tmp_dtype;
  d2->out_dtype = p.out_dtype;
  d2->rank = p.ts_rank - 1;
  d2->naxes = 0;
  for (int i = 0, k = 0; i < p.ts_rank; ++i) {
    if (i == p.ts_peel) continue;
    d2->dims[k] = p.ts_dims[i];
    if (p.ts_isred[i]) d2->axes[d2->naxes++] = k;
    ++k;
  }
  d2->x = tmp;
  d2->out = d->out;
}

int plan_reduce(const b200phi_reduce_desc* d, const Tunables& t, ReducePlan* p, std::string* err) {
  char buf[256];
  if (!d) return fail(err, B200PHI_INVALID_ARGUMENT, "null descriptor");
  *p = ReducePlan();
  if (d->op < B200PHI_SUM || d->op > B200PHI_COUNT_NONZERO)
    return fail(err, B200PHI_INVALID_ARGUMENT, "unknown reduce op");
  if (!dtype_known(d->in_dtype)) return fail(err, B200PHI_UNSUPPORTED, "unsupported input dtype");
  if (d->rank < 0 || d->rank > B200PHI_MAX_RANK)
    return fail(err, B200PHI_INVALID_ARGUMENT, "rank must be in [0, 9]");
  if (d->naxes < 0 || d->naxes > B200PHI_MAX_RANK)
    return fail(err, B200PHI_INVALID_ARGUMENT, "naxes must be in [0, 9]");

  const bool logical = d->op == B200PHI_ANY || d->op == B200PHI_ALL;
  int out_dtype = d->out_dtype;
  if (out_dtype == B200PHI_UNDEFINED) out_dtype = logical ? B200PHI_BOOL : d->in_dtype;
  if (!dtype_known(out_dtype)) return fail(err, B200PHI_UNSUPPORTED, "unsupported output dtype");

  // ---- complex64 / complex128: SUM and MEAN are component-wise, so the tensor is planned (and
  //      run, by the same kernels) as its float lanes: dims + [2] with the trailing 2 kept
  if (dtype_is_complex(d->in_dtype) || dtype_is_complex(out_dtype)) {
    if (d->in_dtype != out_dtype)
      return fail(err, B200PHI_UNSUPPORTED, "complex reductions keep the dtype");
    if (d->op != B200PHI_SUM && d->op != B200PHI_MEAN)
      return fail(err, B200PHI_UNSUPPORTED, "complex dtypes: sum and mean only");
    if (d->rank >= B200PHI_MAX_RANK)
      return fail(err, B200PHI_UNSUPPORTED, "complex dtypes: rank 9 is not built");
    b200phi_reduce_desc lanes = *d;
    lanes.in_dtype = lanes.out_dtype = d->in_dtype == B200PHI_COMPLEX64 ? B200PHI_FLOAT32 : B200PHI_FLOAT64;
    lanes.rank = d->rank + 1;
    lanes.dims[d->rank] = 2;
    lanes.reduce_all = 0;
    if (d->reduce_all != 0 || d->naxes == 0 || d->rank == 0) {
      lanes.naxes = d->rank;
      for (int i = 0; i < d->rank; ++i) lanes.axes[i] = i;
      if (d->rank == 0) {  // a 0-D complex tensor: one (re, im) pair, nothing to combine
        lanes.rank = 2;
        lanes.dims[0] = 1;
        lanes.dims[1] = 2;
        lanes.naxes = 1;
        lanes.axes[0] = 0;
      }
    } else {
      for (int i = 0; i < d->naxes; ++i) {
        int a = d->axes[i];
        if (a < -d->rank || a >= d->rank) {
          snprintf(buf, sizeof(buf),
                   "The reduce dim index %d should be in the range [ -dimension(X), dimension(X) ) "
                   "which dimension = %d. But received dim index = %d.",
                   i, d->rank, a);
          return fail(err, B200PHI_INVALID_ARGUMENT, buf);
        }
        lanes.axes[i] = a < 0 ? a + d->rank : a;
      }
    }
    return plan_reduce(&lanes, t, p, err);
  }

  if (logical && out_dtype != B200PHI_BOOL)
    return fail(err, B200PHI_INVALID_ARGUMENT, "any/all write BOOL");
  if (!logical && out_dtype == B200PHI_BOOL && d->in_dtype != B200PHI_BOOL)
    return fail(err, B200PHI_UNSUPPORTED, "BOOL output needs any/all or a BOOL input");
  if (dtype_is_fp8(d->in_dtype) || dtype_is_fp8(out_dtype)) {
    if (d->in_dtype != out_dtype || (d->op != B200PHI_MAX && d->op != B200PHI_AMAX))
      return fail(err, B200PHI_UNSUPPORTED, "float8 dtypes: max / amax only, same output dtype");
  }

  const bool sum_like = d->op == B200PHI_SUM || d->op == B200PHI_MEAN || d->op == B200PHI_SUM_ABS ||
                        d->op == B200PHI_SUM_SQUARE || d->op == B200PHI_L2_NORM ||
                        d->op == B200PHI_COUNT_NONZERO;
  if (d->op >= B200PHI_SUM_ABS && !dtype_is_float(d->in_dtype))
    return fail(err, B200PHI_UNSUPPORTED, "the norm reductions take floating-point dtypes");
  if (d->op >= B200PHI_SUM_ABS && out_dtype != d->in_dtype && !t.allow_narrow_out)
    return fail(err, B200PHI_UNSUPPORTED, "the norm reductions keep the input dtype");

  // ---- out_dtype != x.dtype.  The reference casts x to the output dtype FIRST and reduces that
  //      (phi::Cast pre-pass, gpu/reduce.h:48-69).  Fused into the load wherever that is the same
  //      arithmetic: sums (the loader converts each element to the accumulation type of the
  //      output dtype), and max / min between floating-point types or into a wider integer
  //      (rounding is monotonic, so max(round(x)) == round(max(x))).  Where it is NOT the same --
  //      a sum into bf16 / fp16 from another dtype (each element is rounded to 16 bits before it is
  //      added), a product, a narrowing integer conversion -- the plan is the reference's own: a cast
  //      pass into scratch of the output dtype, then the reduce (kCastFirst).
  bool cast_first = false;
  if (out_dtype != d->in_dtype && !logical && !t.allow_narrow_out) {
    const bool minmax = d->op == B200PHI_MAX || d->op == B200PHI_MIN || d->op == B200PHI_AMAX ||
                        d->op == B200PHI_AMIN;
    const bool out16 = out_dtype == B200PHI_FLOAT16 || out_dtype == B200PHI_BFLOAT16;
    if (sum_like) {
      cast_first = out16;
    } else if (minmax) {
      const bool both_float = dtype_is_float(d->in_dtype) && dtype_is_float(out_dtype);
      const bool widening_int = (dtype_is_int(d->in_dtype) || d->in_dtype == B200PHI_BOOL) &&
                                dtype_is_int(out_dtype) && dtype_size(out_dtype) > dtype_size(d->in_dtype);
      cast_first = !(both_float || widening_int);
    } else {
      cast_first = true;  // prod
    }
  }

  p->op = d->op;
  p->in_dtype = d->in_dtype;
  p->out_dtype = out_dtype;
  // accumulation type: sums follow the OUTPUT dtype (the loader converts, which is the reference's
  // Cast pre-pass); max / min / prod compute in the INPUT dtype's natural type and convert at the store
  const int acc_of = sum_like ? out_dtype : d->in_dtype;
  if (logical)
    p->acc = 3;
  else if (acc_of == B200PHI_FLOAT64)
    p->acc = 1;
  else if (dtype_is_float(acc_of) || dtype_is_fp8(acc_of))
    p->acc = 0;
  else
    p->acc = 2;
  const size_t acc_size = (p->acc == 0 || p->acc == 3) ? 4 : 8;

  // ---- numel, emptiness (reduce_function.h:1052-1055) -------------------------
  long long numel = 1;
  for (int i = 0; i < d->rank; ++i) {
    if (d->dims[i] < 0) return fail(err, B200PHI_INVALID_ARGUMENT, "negative extent");
    numel *= d->dims[i];
  }
  if (numel == 0)
    return fail(err, B200PHI_INVALID_ARGUMENT, "Tensor need be reduced must not empty.");
  p->in_numel = numel;

  // ---- which axes are reduced (reduce_function.h:100-123, kernel_utils.h:405-416)
  bool red[B200PHI_MAX_RANK] = {false};
  bool all = d->reduce_all != 0 || d->naxes == 0 || d->rank == 0;
  if (!all) {
    for (int i = 0; i < d->naxes; ++i) {
      int a = d->axes[i];
      if (a < -d->rank || a >= d->rank) {
        snprintf(buf, sizeof(buf),
                 "The reduce dim index %d should be in the range [ -dimension(X), dimension(X) ) "
                 "which dimension = %d. But received dim index = %d.",
                 i, d->rank, a);
        return fail(err, B200PHI_INVALID_ARGUMENT, buf);
      }
      if (a < 0) a += d->rank;
      red[a] = true;
    }
  } else {
    for (int i = 0; i < d->rank; ++i) red[i] = true;
  }

  // ---- drop extent-1 dims, merge neighbours with the same role -------------------
  long long ext[B200PHI_MAX_RANK];
  bool isred[B200PHI_MAX_RANK];
  int m = 0;
  for (int i = 0; i < d->rank; ++i) {
    if (d->dims[i] == 1) continue;
    if (m > 0 && isred[m - 1] == red[i]) {
      ext[m - 1] *= d->dims[i];
    } else {
      ext[m] = d->dims[i];
      isred[m] = red[i];
      ++m;
    }
  }
  long long stride[B200PHI_MAX_RANK];
  long long s = 1;
  for (int i = m - 1; i >= 0; --i) {
    stride[i] = s;
    s *= ext[i];
  }
  long long red_numel = 1, out_numel = 1;
  p->nkeep = p->nred = 0;
  for (int i = 0; i < m; ++i) {
    if (isred[i]) {
      p->red_extent[p->nred] = ext[i];
      p->red_stride[p->nred++] = stride[i];
      red_numel *= ext[i];
    } else {
      p->keep_extent[p->nkeep] = ext[i];
      p->keep_stride[p->nkeep++] = stride[i];
      out_numel *= ext[i];
    }
  }
  p->red_numel = red_numel;
  p->out_numel = out_numel;
  if (cast_first && red_numel > 1) {  // (a lone cast, reduce_num == 1, converts at the store)
    // stage 1: the cast, spelled as a reduce over an appended extent-1 dim (reduce_num == 1);
    // stage 2: the reduce in the output dtype
    if (d->rank >= B200PHI_MAX_RANK)
      return fail(err, B200PHI_UNSUPPORTED, "a cast pre-pass on a rank-9 tensor is not built");
    p->kind = kCastFirst;
    p->out_numel = 0;
    p->ts_tmp_dtype = out_dtype;
    p->ts_tmp_bytes = static_cast<size_t>(round_up(numel * static_cast<long long>(dtype_size(out_dtype)), 256));
    b200phi_reduce_desc d2 = *d;
    d2.in_dtype = out_dtype;
    d2.out_dtype = out_dtype;
    d2.x = reinterpret_cast<const void*>(static_cast<uintptr_t>(256));
    ReducePlan p2;
    int rc2 = plan_reduce(&d2, t, &p2, err);
    if (rc2 != B200PHI_OK) return rc2;
    p->out_numel = p2.out_numel;
    p->red_numel = p2.red_numel;
    p->ts_ws2_bytes = static_cast<size_t>(round_up(static_cast<long long>(p2.ws_total_bytes), 256));
    p->ws_partial_bytes = p->ts_tmp_bytes;
    p->ws_total_bytes = p->ts_tmp_bytes + p->ts_ws2_bytes;
    p->grid_x = p2.grid_x;
    p->grid_y = p2.grid_y;
    return B200PHI_OK;
  }

  const size_t esize = dtype_size(d->in_dtype);
  const int VE = static_cast<int>(16 / esize);
  const long long want_ctas = static_cast<long long>(t.sm_count) * t.ctas_per_sm;

  // ---- kind ---------------------------------------------------------------------
  int kind = kGeneric;
  if (red_numel > 1) {
    if (m == 1) {  // [cols] all reduced
      kind = kRowBlock;
      p->rows = 1;
      p->cols = ext[0];
    } else if (m == 2 && !isred[0] && isred[1]) {
      kind = kRowBlock;
      p->rows = ext[0];
      p->cols = ext[1];
    } else if (m == 2 && isred[0] && !isred[1]) {
      kind = kCol;
      p->outer = 1;
      p->R = ext[0];
      p->inner = ext[1];
    } else if (m == 3 && !isred[0] && isred[1] && !isred[2]) {
      kind = kCol;
      p->outer = ext[0];
      p->R = ext[1];
      p->inner = ext[2];
    }
  }
  if (kind == kCol && p->R < 16) kind = kGeneric;

  // ---- two stages: peel ONE reduced dim with a streaming kernel, then reduce what is left ------
  // [A, K, B] with A and B reduced (per-channel statistics of [N,C,H,W]) peels B with the row
  // kernel and finishes with the column kernel over [A, K].  Any other interleaving
  // ([R,K,R,K], [K,R,K,R], ...) peels its longest reduced dim the same way: with everything
  // inside that dim treated as kept, stage 1 is a plain column (or row) reduce at streaming speed
  // over the whole input, and stage 2 sees a tensor smaller by that dim's extent.  The dim must be
  // long enough (>= 64 bytes of elements) for the intermediate's write + read to stay a small
  // fraction of the input; otherwise the single-pass tiled kernels below take the shape.
  int peel = -1;
  if (kind == kGeneric && red_numel > 1 && m >= 3 && !t.no_two_stage && t.force_kind < 0) {
    if (m == 3 && isred[0] && !isred[1] && isred[2] && ext[2] >= 16) {
      peel = 2;
    } else if (esize <= 2 && isred[m - 1] && ext[m - 1] * static_cast<long long>(esize) < 512 &&
               !t.no_tiled && red_numel < 0x7fffffffLL && out_numel < 0x7fffffffLL) {
      // 16-bit tensors whose contiguous run is short and reduced: the single-pass kernel reads
      // them faster than two stages through an fp32 intermediate twice their element size
      // (bf16 sum([256,64,112,112], (1,3)): 6.0 TB/s against 4.9)
    } else {
      for (int i = 0; i < m; ++i) {
        if (!isred[i]) continue;
        const bool innermost = i == m - 1;
        const bool friendly = ext[i] * static_cast<long long>(esize) >= 64 &&
                              (innermost || (ext[i] >= 16 && stride[i] * static_cast<long long>(esize) >= 64));
        if (friendly && (peel < 0 || ext[i] > ext[peel])) peel = i;
      }
    }
  }
  if (peel >= 0) {
    int op1 = -1, op2 = -1, tmp_dtype = B200PHI_UNDEFINED;
    const int acc_dtype = p->acc == 0 ? B200PHI_FLOAT32 : (p->acc == 1 ? B200PHI_FLOAT64 : B200PHI_INT64);
    const bool wide_in = esize >= 4;
    switch (d->op) {
      case B200PHI_SUM:
      case B200PHI_SUM_ABS:
      case B200PHI_SUM_SQUARE:
      case B200PHI_COUNT_NONZERO:
        op1 = d->op, op2 = B200PHI_SUM, tmp_dtype = acc_dtype;
        break;
      case B200PHI_MEAN:  // equal-sized groups: the mean of the row means is the mean
        if (p->acc <= 1) op1 = op2 = B200PHI_MEAN, tmp_dtype = acc_dtype;
        break;
      case B200PHI_MAX:
      case B200PHI_AMAX:
      case B200PHI_MIN:
      case B200PHI_AMIN:
        op1 = op2 = d->op, tmp_dtype = d->in_dtype;
        break;
      case B200PHI_MAX_ABS:
        op1 = d->op, op2 = B200PHI_MAX, tmp_dtype = d->in_dtype;
        break;
      case B200PHI_MIN_ABS:
        op1 = d->op, op2 = B200PHI_MIN, tmp_dtype = d->in_dtype;
        break;
      case B200PHI_PROD:  // a 16-bit intermediate would round the partial products
        if (wide_in) op1 = op2 = d->op, tmp_dtype = d->in_dtype;
        break;
      case B200PHI_ANY:
      case B200PHI_ALL:
        op1 = op2 = d->op, tmp_dtype = B200PHI_BOOL;
        break;
      default:  // L2_NORM: the sqrt belongs after the second stage only
        break;
    }
    if (op1 >= 0) {
      p->kind = kTwoStage;
      p->ts_rank = m;
      p->ts_peel = peel;
      for (int i = 0; i < m; ++i) {
        p->ts_dims[i] = ext[i];
        p->ts_isred[i] = isred[i] ? 1 : 0;
      }
      p->ts_op1 = op1;
      p->ts_op2 = op2;
      p->ts_tmp_dtype = tmp_dtype;
      p->ts_tmp_bytes = static_cast<size_t>(
          round_up((numel / ext[peel]) * static_cast<long long>(dtype_size(tmp_dtype)), 256));
      b200phi_reduce_desc d1, d2;
      two_stage_descs(*p, d, reinterpret_cast<void*>(static_cast<uintptr_t>(256)), &d1, &d2);
      Tunables t2 = t;
      t2.allow_narrow_out = 1;
      ReducePlan p1, p2;
      int rc = plan_reduce(&d1, t, &p1, err);
      if (rc != B200PHI_OK) return rc;
      rc = plan_reduce(&d2, t2, &p2, err);
      if (rc != B200PHI_OK) return rc;
      p->ts_ws1_bytes = static_cast<size_t>(round_up(static_cast<long long>(p1.ws_total_bytes), 256));
      p->ts_ws2_bytes = static_cast<size_t>(round_up(static_cast<long long>(p2.ws_total_bytes), 256));
      p->ws_partial_bytes = p->ts_tmp_bytes;
      p->ws_ticket_bytes = 0;
      p->ws_total_bytes = p->ts_tmp_bytes + p->ts_ws1_bytes + p->ts_ws2_bytes;
      p->grid_x = p1.grid_x;
      p->grid_y = p1.grid_y;
      return B200PHI_OK;
    }
  }

  // ---- every other keep/reduce interleaving: the tiled kernels (32-bit index spaces) ----------
  if (kind == kGeneric && red_numel > 1 && m >= 2 && !t.no_tiled && t.force_kind < 0 &&
      red_numel < 0x7fffffffLL && out_numel < 0x7fffffffLL) {
    const bool inner_kept = !isred[m - 1];
    const long long inner = ext[m - 1];
    const bool vec_ok = VE > 1 && inner % VE == 0 && reinterpret_cast<uintptr_t>(d->x) % 16 == 0;
    p->vec = vec_ok ? VE : 1;
    p->td_nkeep = p->td_nred = 0;
    for (int i = 0; i < m; ++i) {
      long long e = ext[i], st = stride[i];
      const bool last = i == m - 1;
      if (!isred[i]) {
        if (last) e /= p->vec, st = p->vec;  // the kept innermost dim is walked in vectors
        p->td_keep_extent[p->td_nkeep] = static_cast<unsigned>(e);
        p->td_keep_stride[p->td_nkeep++] = st;
      } else if (!last) {  // the reduced innermost dim is the contiguous run, not a divided dim
        p->td_red_extent[p->td_nred] = static_cast<unsigned>(e);
        p->td_red_stride[p->td_nred++] = st;
      }
    }
    if (inner_kept) {
      kind = kTiledKeep;
      p->nq = static_cast<unsigned>(out_numel / p->vec);
      p->TX = std::min(256, pow2_ceil(p->nq));
      const int TY = 256 / p->TX;
      const long long gx = ceil_div(p->nq, p->TX);
      // a split must keep at least 8 rows per thread and 256 KiB per block: below that the block's
      // start-up and the finalise cost more than the better wave fit gives back
      const long long max_splits = std::max<long long>(
          1, std::min(red_numel / (static_cast<long long>(TY) * 8),
                      red_numel * p->TX * p->vec * static_cast<long long>(esize) / (256 << 10)));
      long long splits = pick_splits(gx, want_ctas, max_splits,
                                      static_cast<double>(red_numel) * esize / acc_size);
      p->rows_per_split = round_up(ceil_div(red_numel, splits), TY);
      splits = ceil_div(red_numel, p->rows_per_split);
      p->splits = static_cast<int>(splits);
      p->grid_x = static_cast<unsigned>(gx);
      p->grid_y = static_cast<unsigned>(splits);
    } else {
      const long long Lv = inner / p->vec;
      const long long W = (red_numel / inner) * Lv;
      if (W < 0x7fffffffLL) {
        kind = kTiledRed;
        p->Lv = static_cast<unsigned>(Lv);
        p->W = static_cast<unsigned>(W);
        // a warp per output unless an output is long (>= 128 KiB) or there are too few outputs to
        // fill the machine with warps; a CTA per 28 KiB output spent its time in barriers
        p->G = (W >= 8192 || out_numel < 8 * want_ctas) ? 256 : 32;
        const long long gx = ceil_div(out_numel, 256 / p->G);
        long long splits = 1;
        if (p->G == 256) {
          splits = std::max<long long>(1, ceil_div(want_ctas, gx));
          splits = std::min(splits, std::max<long long>(1, W / (256 * 16)));
          splits = std::min<long long>(splits, 65535);
        }
        p->chunk = static_cast<unsigned>(round_up(ceil_div(W, splits), 256));
        splits = ceil_div(W, p->chunk);
        p->splits = static_cast<int>(splits);
        p->grid_x = static_cast<unsigned>(gx);
        p->grid_y = static_cast<unsigned>(splits);
      }
    }
    if (kind != kGeneric) {
      p->kind = kind;
      if (p->splits > 1) {
        p->ws_partial_bytes = static_cast<size_t>(p->splits) * out_numel * acc_size;
        p->ws_ticket_bytes = static_cast<size_t>(p->grid_x) * 4;
      }
      p->ws_partial_bytes = static_cast<size_t>(round_up(static_cast<long long>(p->ws_partial_bytes), 256));
      p->ws_total_bytes = p->ws_partial_bytes + p->ws_ticket_bytes;
      return B200PHI_OK;
    }
  }

  if (kind == kRowBlock) {
    const long long row_bytes = p->cols * static_cast<long long>(esize);
    // threads per row: each thread should own about 4 16-byte vectors of the row
    int G = pow2_ceil(ceil_div(row_bytes, std::max<long long>(16, t.row_bytes_per_thread)));
    G = std::min(256, std::max(32, G));
    if (t.force_kind == kRowWarp) G = 32;
    if (t.force_kind == kRowBlock) G = 256;
    p->G = G;
    p->aligned = (reinterpret_cast<uintptr_t>(d->x) % 16 == 0 && row_bytes % 16 == 0) ? 1 : 0;
    if (p->rows < static_cast<long long>(t.sm_count) * 2) {
      long long splits = std::min(ceil_div(want_ctas, p->rows), row_bytes / 32768);
      splits = std::min<long long>(splits, 65535);
      if (splits > 1) {
        long long seg = round_up(ceil_div(p->cols, splits), static_cast<long long>(VE) * 32);
        splits = ceil_div(p->cols, seg);
        if (splits > 1) {
          kind = kRowSplit;
          p->seg = seg;
          p->splits = static_cast<int>(splits);
        }
      }
    }
    if (kind != kRowSplit) kind = (G == 256) ? kRowBlock : kRowWarp;
    // rows under 16 KiB: every load of R rows in one round (row_batch_kernel)
    const long long nvec_row = row_bytes / 16;
    if (kind != kRowSplit && p->aligned && nvec_row < 1024 && (p->cols > 128 || nvec_row > 16) && !t.no_row_batch &&
        t.force_kind < 0) {
      kind = kRowBatch;
      if (nvec_row <= 32) {  // sub-warp groups: a 3- or 4-step shuffle tree instead of 5 per row
        p->G = 8, p->nv = 4, p->rpg = 2;
      } else if (nvec_row <= 64) {
        p->G = 16, p->nv = 4, p->rpg = 2;
      } else if (nvec_row <= 128) {
        p->G = 16, p->nv = 8, p->rpg = 1;
      } else if (nvec_row <= 256) {
        p->G = 32, p->nv = 8, p->rpg = 1;
      } else if (nvec_row <= 512) {
        p->G = 64, p->nv = 8, p->rpg = 1;
      } else {
        p->G = 128, p->nv = 8, p->rpg = 1;
      }
    }
    // aligned rows of at most 16 vectors
    if ((kind == kRowWarp || kind == kRowBlock) && p->aligned && nvec_row >= 1 && nvec_row <= 16 &&
        p->cols <= 128 && p->rows >= 64 && t.force_kind < 0 && !t.no_row_batch) {
      kind = kRowTiny;
      p->G = pow2_ceil(nvec_row);
      p->rpg = 8;
    }
    // short rows that are not whole aligned vectors
    // (rows of 33..128 elements take it whether aligned or not: the vector kernels start above that)
    if ((kind == kRowWarp || kind == kRowBlock) && p->cols > 32 && p->cols < 1024 &&
        (p->cols <= 128 || !p->aligned) && p->rows >= 64 && t.force_kind < 0 && !t.no_row_batch) {
      kind = kRowScalar;
      if (p->cols <= 64) {
        p->G = 8, p->nv = 8;
      } else if (p->cols <= 128) {
        p->G = 8, p->nv = 16;
      } else if (p->cols < 256) {
        p->G = 16, p->nv = 16;
      } else if (p->cols < 512) {
        p->G = 32, p->nv = 16;
      } else {
        p->G = 32, p->nv = 32;
      }
    }
    // 1- and 2-byte rows that are not whole aligned vectors, 4 .. 256 vectors long: head element,
    // vectors and tail element per lane (row_peel_kernel) instead of 2-byte loads
    if ((kind == kRowWarp || kind == kRowBlock || kind == kRowScalar) && esize <= 2 && !p->aligned &&
        p->cols / VE >= 4 && p->cols / VE <= 512 && p->rows >= 64 && t.force_kind < 0 && !t.no_row_batch &&
        !t.no_row_peel) {
      const long long nv_row = p->cols / VE;
      const int gmin = VE > 8 ? 16 : 8;  // a lane per possible head / tail element
      kind = kRowPeel;
      if (nv_row <= 16 && gmin == 8) {
        p->G = 8, p->nv = 2;
      } else if (nv_row <= 32 && gmin == 8) {
        p->G = 8, p->nv = 4;
      } else if (nv_row <= 64) {
        p->G = gmin == 8 ? 8 : 16, p->nv = gmin == 8 ? 8 : 4;
      } else if (nv_row <= 128) {
        p->G = 16, p->nv = 8;
      } else if (nv_row <= 256) {
        p->G = 32, p->nv = 8;
      } else {
        p->G = 32, p->nv = 16;  // 4..8 KiB rows: a warp per row, 16 vectors per lane
      }
      p->rpg = 1;
    }
    // short rows: fewer than 32 lanes per row (about 8 elements per lane)
    if (kind == kRowWarp && p->cols <= 128 && p->rows >= 64 && t.force_kind < 0) {
      int g = pow2_ceil(ceil_div(p->cols, 8));
      p->G = std::min(16, std::max(1, g));
      kind = kRowSmall;
    }
  }
  // short rows that are not whole aligned vectors, whatever the element type: a thread per row out
  // of shared memory (the sub-warp kernels above are bound by instruction issue there)
  if ((kind == kRowSmall || kind == kRowScalar || kind == kRowPeel || kind == kRowWarp || kind == kRowTiny ||
       kind == kRowBatch) &&
      t.force_kind < 0 &&
      !t.no_row_staged && !t.no_row_batch && staged_rows_ok(p->rows, p->cols * static_cast<long long>(esize), esize, kStagedReduce,
                     reinterpret_cast<uintptr_t>(d->x) % 16 == 0)) {
    kind = kRowStaged;
  }
  p->kind = kind;

  // ---- geometry -----------------------------------------------------------------
  p->splits = (kind == kRowSplit) ? p->splits : 1;
  switch (kind) {
    case kRowWarp:
    case kRowBlock:
    case kRowSmall: {
      const long long ctas = ceil_div(p->rows, 256 / p->G);
      p->grid_x = static_cast<unsigned>(std::min<long long>(ctas, 0x7fffffffLL));
      p->grid_y = 1;
      break;
    }
    case kRowStaged:  // (the launcher sizes the persistent grid from the ring's shared memory)
      p->grid_x = 1;
      p->grid_y = 1;
      break;
    case kRowTiny: {
      const long long per_cta = 8LL * (32 / p->G) * p->rpg;
      long long ctas = ceil_div(p->rows, per_cta);
      const long long cap = static_cast<long long>(t.sm_count) * 8 * std::max(1, t.row_batch_waves);
      if (ctas > cap) ctas = cap;
      p->grid_x = static_cast<unsigned>(ctas);
      p->grid_y = 1;
      break;
    }
    case kRowPeel:
    case kRowScalar: {
      const long long per_cta = 8LL * (32 / p->G) * (kind == kRowPeel ? p->rpg : 1);
      long long ctas = ceil_div(p->rows, per_cta);
      const long long cap = static_cast<long long>(t.sm_count) * 8 * std::max(1, t.row_batch_waves);
      if (ctas > cap) ctas = cap;
      p->grid_x = static_cast<unsigned>(ctas);
      p->grid_y = 1;
      break;
    }
    case kRowBatch: {
      const long long per_cta = static_cast<long long>(256 / p->G) * p->rpg;
      long long ctas = ceil_div(p->rows, per_cta);
      const long long cap = static_cast<long long>(t.sm_count) * 8 * std::max(1, t.row_batch_waves);
      if (ctas > cap) ctas = cap;  // grid-stride beyond that
      p->grid_x = static_cast<unsigned>(ctas);
      p->grid_y = 1;
      break;
    }
    case kRowSplit:
      p->grid_x = static_cast<unsigned>(p->splits);
      p->grid_y = static_cast<unsigned>(p->rows);
      p->ws_partial_bytes = static_cast<size_t>(p->rows) * p->splits * acc_size;
      p->ws_ticket_bytes = static_cast<size_t>(p->rows) * 4;
      break;
    case kCol: {
      const bool vec_ok =
          VE > 1 && (p->inner % VE == 0) && (reinterpret_cast<uintptr_t>(d->x) % 16 == 0);
      p->vec = vec_ok ? VE : 1;
      const long long inner_v = p->inner / p->vec;
      p->TX = std::min(256, pow2_ceil(inner_v));
      const int TY = 256 / p->TX;
      p->tiles = ceil_div(inner_v, p->TX);
      const long long bxy = p->outer * p->tiles;
      if (bxy > 0x7fffffffLL) return fail(err, B200PHI_UNSUPPORTED, "output too large for col plan");
      // (a split keeps at least 8 rows per thread and 256 KiB per block)
      const long long max_splits = std::max<long long>(
          1, std::min(p->R / (static_cast<long long>(TY) * 8),
                      p->R * p->TX * p->vec * static_cast<long long>(esize) / (256 << 10)));
      // fill the resident CTA slots in whole waves (pick_splits): 128 tiles on 592 slots run as 4
      // splits in one wave, not 5 splits in two ([8192,131072] axis 0: 5.71 TB/s with 5)
      long long splits = pick_splits(bxy, want_ctas, max_splits,
                                      static_cast<double>(p->R) * esize / acc_size);
      p->rows_per_split = round_up(ceil_div(p->R, splits), TY);
      splits = ceil_div(p->R, p->rows_per_split);
      p->splits = static_cast<int>(splits);
      p->grid_x = static_cast<unsigned>(bxy);
      p->grid_y = static_cast<unsigned>(splits);
      if (splits > 1) {
        p->ws_partial_bytes = static_cast<size_t>(splits) * out_numel * acc_size;
        p->ws_ticket_bytes = static_cast<size_t>(bxy) * 4;
      }
      break;
    }
    default: {  // kGeneric
      const bool inner_reduced = m > 0 && isred[m - 1];
      int TY = std::min(256, pow2_ceil(red_numel));
      if (!inner_reduced) TY = std::min(TY, 8);
      p->TX = 256 / TY;
      const long long gx = ceil_div(out_numel, p->TX);
      if (gx > 0x7fffffffLL) return fail(err, B200PHI_UNSUPPORTED, "output too large");
      p->grid_x = static_cast<unsigned>(gx);
      p->grid_y = 1;
      break;
    }
  }
  p->ws_partial_bytes = static_cast<size_t>(round_up(static_cast<long long>(p->ws_partial_bytes), 256));
  p->ws_total_bytes = p->ws_partial_bytes + p->ws_ticket_bytes;
  return B200PHI_OK;
}

}  // namespace b200phi
