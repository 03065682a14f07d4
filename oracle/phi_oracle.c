/*
 * phi_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, CPU restatement ("port") of the reference's phi CPU arithmetic for
 * the reduce / argmax / broadcast hot path.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this file's
 * library; the product library (paddle2_b200/libb200phi.so) never does.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here
 *   (a) against oracle/_ref/libphi_ref.so -- the reference's own functors + the
 *       vendored Eigen compiled from /root/reference (oracle/ref_src), and
 *   (b) against tests/golden/ fixtures generated from that library
 *       (tests/golden/make_golden.py), which also hold the literal answers of
 *       the reference's docstring examples.
 *
 * What is restated, with the reference lines it follows:
 *   cast-then-reduce in the output dtype      paddle/phi/kernels/cpu/reduce.h:25-67
 *   any/all: cast to bool first               cpu/reduce.h:69-99
 *   reduce_all / full-dim rule                kernel_utils.h:405-416, cpu/reduce.h:33-46
 *   rank > 6: reduced dims moved last, 2-D    funcs/reduce_function.h:1316-1349
 *   summation order of Eigen's reducers       third_party/eigen3/unsupported/Eigen/CXX11/src/Tensor/
 *       TensorReduction.h:177-260 (InnerMostDimReducer: 1024-packet leaves, tree above,
 *       two packet accumulators), TensorFunctors.h:95-170 (Sum/Mean), :300-320 (Prod)
 *       with AVX packets of 8 floats / 4 doubles (the _ref build uses -mavx2)
 *   argmax/argmin: in-order scan, strict compare, seed (0, lowest/highest)
 *       TensorFunctors.h:369-405, cpu/arg_min_max_kernel.cc:35-66,96-140
 *   broadcast index walk                      funcs/elementwise_base.h:273-307, common_shape.h:35-92
 *   functors                                  funcs/elementwise_functor.h:39+, funcs/functors.h:36-64,
 *                                             funcs/activation_functor.h:1473-1480 (relu)
 *   paddle.var composite                      python/paddle/tensor/stat.py:184-197
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define EXPORT __attribute__((visibility("default")))

/* phi::DataType values (paddle/phi/common/data_type.h:42-85) */
enum { DT_BOOL = 1, DT_UINT8 = 2, DT_INT8 = 3, DT_INT16 = 5, DT_INT32 = 7, DT_INT64 = 9,
       DT_F32 = 10, DT_F64 = 11 };
enum { OP_SUM = 0, OP_MEAN, OP_MAX, OP_MIN, OP_PROD, OP_ANY, OP_ALL, OP_AMAX, OP_AMIN };

static size_t dt_size(int dt) {
  switch (dt) {
    case DT_BOOL: case DT_UINT8: case DT_INT8: return 1;
    case DT_INT16: return 2;
    case DT_INT32: case DT_F32: return 4;
    case DT_INT64: case DT_F64: return 8;
    default: return 0;
  }
}

/* ---- element cast (phi::Cast) ------------------------------------------------ */
static double load_as_double(const void* p, int dt, int64_t i) {
  switch (dt) {
    case DT_BOOL: return ((const uint8_t*)p)[i] != 0;
    case DT_UINT8: return ((const uint8_t*)p)[i];
    case DT_INT8: return ((const int8_t*)p)[i];
    case DT_INT16: return ((const int16_t*)p)[i];
    case DT_INT32: return ((const int32_t*)p)[i];
    case DT_INT64: return (double)((const int64_t*)p)[i];
    case DT_F32: return ((const float*)p)[i];
    default: return ((const double*)p)[i];
  }
}
static int64_t load_as_i64(const void* p, int dt, int64_t i) {
  switch (dt) {
    case DT_BOOL: return ((const uint8_t*)p)[i] != 0;
    case DT_UINT8: return ((const uint8_t*)p)[i];
    case DT_INT8: return ((const int8_t*)p)[i];
    case DT_INT16: return ((const int16_t*)p)[i];
    case DT_INT32: return ((const int32_t*)p)[i];
    case DT_INT64: return ((const int64_t*)p)[i];
    case DT_F32: return (int64_t)((const float*)p)[i];
    default: return (int64_t)((const double*)p)[i];
  }
}
static int is_float_dt(int dt) { return dt == DT_F32 || dt == DT_F64; }

static void cast_buffer(const void* src, int sdt, void* dst, int ddt, int64_t n) {
  if (sdt == ddt) { memcpy(dst, src, (size_t)n * dt_size(sdt)); return; }
  for (int64_t i = 0; i < n; ++i) {
    if (ddt == DT_F32) ((float*)dst)[i] = is_float_dt(sdt) ? (float)load_as_double(src, sdt, i)
                                                           : (float)load_as_i64(src, sdt, i);
    else if (ddt == DT_F64) ((double*)dst)[i] = is_float_dt(sdt) ? load_as_double(src, sdt, i)
                                                                 : (double)load_as_i64(src, sdt, i);
    else if (ddt == DT_BOOL) ((uint8_t*)dst)[i] = is_float_dt(sdt) ? (load_as_double(src, sdt, i) != 0.0)
                                                                   : (load_as_i64(src, sdt, i) != 0);
    else {
      int64_t v = load_as_i64(src, sdt, i);
      switch (ddt) {
        case DT_UINT8: ((uint8_t*)dst)[i] = (uint8_t)v; break;
        case DT_INT8: ((int8_t*)dst)[i] = (int8_t)v; break;
        case DT_INT16: ((int16_t*)dst)[i] = (int16_t)v; break;
        case DT_INT32: ((int32_t*)dst)[i] = (int32_t)v; break;
        default: ((int64_t*)dst)[i] = v; break;
      }
    }
  }
}

/* ---- Eigen's inner-most-dim reducers, restated ------------------------------------ */
static int64_t divup(int64_t a, int64_t b) { return (a + b - 1) / b; }
#define KLEAF 1024

/* predux of an AVX packet: (lo half + hi half) lane-wise, then (0+2) + (1+3) */
#define DEF_FLOAT_REDUCERS(T, NAME, P)                                                        \
  static T predux_add_##NAME(const T* a) {                                                    \
    T h[P / 2];                                                                               \
    for (int i = 0; i < P / 2; ++i) h[i] = a[i] + a[i + P / 2];                               \
    if (P == 8) return (h[0] + h[2]) + (h[1] + h[3]);                                         \
    return h[0] + h[1];                                                                       \
  }                                                                                           \
  static T predux_mul_##NAME(const T* a) {                                                    \
    T h[P / 2];                                                                               \
    for (int i = 0; i < P / 2; ++i) h[i] = a[i] * a[i + P / 2];                               \
    if (P == 8) return (h[0] * h[2]) * (h[1] * h[3]);                                         \
    return h[0] * h[1];                                                                       \
  }                                                                                           \
  /* SumReducer: InnerMostDimReducer<Self,Op,true,true> (TensorReduction.h:213-260) */        \
  static T sum_tree_##NAME(const T* x, int64_t first, int64_t n) {                            \
    if (n > (int64_t)P * KLEAF) {                                                             \
      const int64_t split = P * divup(first + divup(n, 2), P);                                \
      int64_t nl = split - first; if (nl > n) nl = n;                                         \
      T acc = (T)0;                                                                           \
      acc = acc + sum_tree_##NAME(x, first, nl);                                              \
      if (nl < n) acc = acc + sum_tree_##NAME(x, split, n - nl);                              \
      return acc;                                                                             \
    }                                                                                         \
    const int64_t unroll = (n / (2 * P)) * 2 * P, vec = (n / P) * P;                          \
    T p1[P], p2[P];                                                                           \
    for (int i = 0; i < P; ++i) p1[i] = p2[i] = (T)0;                                         \
    for (int64_t j = 0; j < unroll; j += 2 * P)                                               \
      for (int i = 0; i < P; ++i) { p1[i] += x[first + j + i]; p2[i] += x[first + j + P + i]; } \
    for (int64_t j = unroll; j < vec; j += P)                                                 \
      for (int i = 0; i < P; ++i) p1[i] += x[first + j + i];                                  \
    for (int i = 0; i < P; ++i) p1[i] += p2[i];                                               \
    T s = (T)0;                                                                               \
    for (int64_t j = vec; j < n; ++j) s += x[first + j];                                      \
    return s + predux_add_##NAME(p1);                                                         \
  }                                                                                           \
  /* Mean/Prod: InnerMostDimReducer<Self,Op,true,false> (TensorReduction.h:177-195) */        \
  static T sum_flat_##NAME(const T* x, int64_t first, int64_t n) {                            \
    const int64_t vec = (n / P) * P;                                                          \
    T p1[P];                                                                                  \
    for (int i = 0; i < P; ++i) p1[i] = (T)0;                                                 \
    for (int64_t j = 0; j < vec; j += P)                                                      \
      for (int i = 0; i < P; ++i) p1[i] += x[first + j + i];                                  \
    T s = (T)0;                                                                               \
    for (int64_t j = vec; j < n; ++j) s += x[first + j];                                      \
    return s + predux_add_##NAME(p1);                                                         \
  }                                                                                           \
  static T prod_flat_##NAME(const T* x, int64_t first, int64_t n) {                           \
    const int64_t vec = (n / P) * P;                                                          \
    T p1[P];                                                                                  \
    for (int i = 0; i < P; ++i) p1[i] = (T)1;                                                 \
    for (int64_t j = 0; j < vec; j += P)                                                      \
      for (int i = 0; i < P; ++i) p1[i] *= x[first + j + i];                                  \
    T s = (T)1;                                                                               \
    for (int64_t j = vec; j < n; ++j) s *= x[first + j];                                      \
    return s * predux_mul_##NAME(p1);                                                         \
  }

DEF_FLOAT_REDUCERS(float, f32, 8)
DEF_FLOAT_REDUCERS(double, f64, 4)

/* ---- generic reduce driver ----------------------------------------------------------- */
typedef struct {
  int rank;
  int64_t dims[9];
  int red[9];            /* 1 if the axis is reduced */
  int full;              /* every axis is reduced (reduce_all) */
  int64_t out_numel, red_numel;
} rplan;

static int make_plan(rplan* p, int rank, const int64_t* dims, int naxes, const int32_t* axes,
                     int reduce_all) {
  if (rank < 0 || rank > 9) return 1;
  p->rank = rank;
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) { p->dims[i] = dims[i]; p->red[i] = 0; numel *= dims[i]; }
  if (numel <= 0) return 1;  /* "Tensor need be reduced must not empty." */
  if (reduce_all || naxes == 0 || rank == 0) {
    for (int i = 0; i < rank; ++i) p->red[i] = 1;
  } else {
    for (int i = 0; i < naxes; ++i) {
      int a = axes[i];
      if (a < -rank || a >= rank) return 1;
      if (a < 0) a += rank;
      p->red[a] = 1;
    }
  }
  p->out_numel = p->red_numel = 1;
  for (int i = 0; i < rank; ++i) {
    if (p->red[i]) p->red_numel *= dims[i]; else p->out_numel *= dims[i];
  }
  p->full = 1;
  for (int i = 0; i < rank; ++i) if (!p->red[i]) p->full = 0;
  return 0;
}

/* offset of output element `o` (row-major over kept axes) and iteration over the
 * reduced axes in row-major order; `gather` collects the reduced values in that order */
static void kept_base_and_strides(const rplan* p, int64_t o, int64_t* base, int64_t* rstride,
                                  int64_t* rext, int* nred) {
  int64_t stride[9];
  int64_t s = 1;
  for (int i = p->rank - 1; i >= 0; --i) { stride[i] = s; s *= p->dims[i]; }
  int64_t b = 0;
  for (int i = p->rank - 1; i >= 0; --i) {
    if (!p->red[i]) { b += (o % p->dims[i]) * stride[i]; o /= p->dims[i]; }
  }
  *base = b;
  int k = 0;
  for (int i = 0; i < p->rank; ++i)
    if (p->red[i]) { rstride[k] = stride[i]; rext[k] = p->dims[i]; ++k; }
  *nred = k;
}

#define DEF_REDUCE_TYPED(T, NAME, ISFLOAT, LOWEST, HIGHEST)                                   \
  static void reduce_##NAME(int op, const rplan* p, const T* x, T* out) {                     \
    if (p->out_numel == 1 && p->full) {                                                       \
      /* reduce_all: flattened 1-D tensor -> FullReducer -> InnerMostDimReducer               \
       * (reduce_function.h:1367-1375; TensorReduction.h:177-260) */                          \
      const int64_t n = p->red_numel;                                                         \
      T r;                                                                                    \
      if (op == OP_SUM) r = ISFLOAT ? sum_tree_dispatch_##NAME(x, 0, n) : seq_##NAME(OP_SUM, x, n); \
      else if (op == OP_MEAN) r = ISFLOAT ? (T)(sum_flat_dispatch_##NAME(x, 0, n) / (T)n)     \
                                          : (T)(seq_##NAME(OP_SUM, x, n) / (T)n);             \
      else if (op == OP_PROD) r = ISFLOAT ? prod_flat_dispatch_##NAME(x, 0, n)                \
                                          : seq_##NAME(OP_PROD, x, n);                        \
      else r = seq_##NAME(op, x, n);                                                          \
      out[0] = r;                                                                             \
      return;                                                                                 \
    }                                                                                         \
    /* Partial reductions: the reduce dims are a run-time Eigen::array, so neither            \
     * ReducingInnerMostDims nor PreservingInnerMostDims is statically known and every        \
     * output coefficient goes through GenericDimReducer (TensorReduction.h:142-160,          \
     * :788-797): plain in-order accumulation where the LAST reduced axis is the outer        \
     * loop and the FIRST reduced axis the inner one.  rank > 6 first moves the reduced       \
     * axes to the end (HandleLargeDim) and reduces a 2-D [unreduced, reduced] matrix         \
     * over axis 1: a single reduced axis, walked in row-major order of the moved axes. */    \
    const int transposed = p->rank > 6;                                                       \
    for (int64_t o = 0; o < p->out_numel; ++o) {                                              \
      int64_t base, rs[9], re[9]; int nr;                                                     \
      kept_base_and_strides(p, o, &base, rs, re, &nr);                                        \
      int64_t idx[9] = {0};                                                                   \
      T acc = (op == OP_PROD) ? (T)1 : (op == OP_MAX || op == OP_AMAX) ? (T)(LOWEST)          \
              : (op == OP_MIN || op == OP_AMIN) ? (T)(HIGHEST) : (T)0;                        \
      for (int64_t j = 0; j < p->red_numel; ++j) {                                            \
        int64_t off = 0;                                                                      \
        for (int k = 0; k < nr; ++k) off += idx[k] * rs[k];                                   \
        const T v = x[base + off];                                                            \
        if (op == OP_SUM || op == OP_MEAN) acc = (T)(acc + v);                                \
        else if (op == OP_PROD) acc = (T)(acc * v);                                           \
        else if (op == OP_MAX || op == OP_AMAX) acc = (v > acc) ? v : acc;                    \
        else acc = (v < acc) ? v : acc;                                                       \
        if (transposed) {                                                                     \
          for (int k = nr - 1; k >= 0; --k) { if (++idx[k] < re[k]) break; idx[k] = 0; }      \
        } else {                                                                              \
          for (int k = 0; k < nr; ++k) { if (++idx[k] < re[k]) break; idx[k] = 0; }           \
        }                                                                                     \
      }                                                                                       \
      out[o] = (op == OP_MEAN) ? (T)(acc / (T)p->red_numel) : acc;                            \
    }                                                                                         \
  }

#define DEF_SEQ(T, NAME, LOWEST, HIGHEST)                                                     \
  static T seq_##NAME(int op, const T* x, int64_t n) {                                        \
    T acc = (op == OP_PROD) ? (T)1 : (op == OP_MAX || op == OP_AMAX) ? (T)(LOWEST)            \
            : (op == OP_MIN || op == OP_AMIN) ? (T)(HIGHEST) : (T)0;                          \
    for (int64_t j = 0; j < n; ++j) {                                                         \
      const T v = x[j];                                                                       \
      if (op == OP_SUM) acc = (T)(acc + v);                                                   \
      else if (op == OP_PROD) acc = (T)(acc * v);                                             \
      else if (op == OP_MAX || op == OP_AMAX) acc = (v > acc) ? v : acc;                      \
      else acc = (v < acc) ? v : acc;                                                         \
    }                                                                                         \
    return acc;                                                                               \
  }

/* integer types never take the packet paths: give them no-op dispatchers */
#define DEF_INT_DISPATCH(T, NAME)                                                             \
  static T sum_tree_dispatch_##NAME(const T* x, int64_t f, int64_t n) { (void)x; (void)f; (void)n; return 0; } \
  static T sum_flat_dispatch_##NAME(const T* x, int64_t f, int64_t n) { (void)x; (void)f; (void)n; return 0; } \
  static T prod_flat_dispatch_##NAME(const T* x, int64_t f, int64_t n) { (void)x; (void)f; (void)n; return 0; }

static float sum_tree_dispatch_f32(const float* x, int64_t f, int64_t n) { return sum_tree_f32(x, f, n); }
static float sum_flat_dispatch_f32(const float* x, int64_t f, int64_t n) { return sum_flat_f32(x, f, n); }
static float prod_flat_dispatch_f32(const float* x, int64_t f, int64_t n) { return prod_flat_f32(x, f, n); }
static double sum_tree_dispatch_f64(const double* x, int64_t f, int64_t n) { return sum_tree_f64(x, f, n); }
static double sum_flat_dispatch_f64(const double* x, int64_t f, int64_t n) { return sum_flat_f64(x, f, n); }
static double prod_flat_dispatch_f64(const double* x, int64_t f, int64_t n) { return prod_flat_f64(x, f, n); }
DEF_INT_DISPATCH(int32_t, i32)
DEF_INT_DISPATCH(int64_t, i64)
DEF_INT_DISPATCH(int16_t, i16)
DEF_INT_DISPATCH(int8_t, i8)
DEF_INT_DISPATCH(uint8_t, u8)

DEF_SEQ(float, f32, -3.402823466e+38f, 3.402823466e+38f)
DEF_SEQ(double, f64, -1.7976931348623157e+308, 1.7976931348623157e+308)
DEF_SEQ(int32_t, i32, INT32_MIN, INT32_MAX)
DEF_SEQ(int64_t, i64, INT64_MIN, INT64_MAX)
DEF_SEQ(int16_t, i16, INT16_MIN, INT16_MAX)
DEF_SEQ(int8_t, i8, INT8_MIN, INT8_MAX)
DEF_SEQ(uint8_t, u8, 0, UINT8_MAX)

DEF_REDUCE_TYPED(float, f32, 1, -3.402823466e+38f, 3.402823466e+38f)
DEF_REDUCE_TYPED(double, f64, 1, -1.7976931348623157e+308, 1.7976931348623157e+308)
DEF_REDUCE_TYPED(int32_t, i32, 0, INT32_MIN, INT32_MAX)
DEF_REDUCE_TYPED(int64_t, i64, 0, INT64_MIN, INT64_MAX)
DEF_REDUCE_TYPED(int16_t, i16, 0, INT16_MIN, INT16_MAX)
DEF_REDUCE_TYPED(int8_t, i8, 0, INT8_MIN, INT8_MAX)
DEF_REDUCE_TYPED(uint8_t, u8, 0, 0, UINT8_MAX)

/* any/all on bool bytes (cpu/reduce.h:69-99) */
static void reduce_bool(int op, const rplan* p, const uint8_t* x, uint8_t* out) {
  for (int64_t o = 0; o < p->out_numel; ++o) {
    int64_t base, rs[9], re[9]; int nr;
    kept_base_and_strides(p, o, &base, rs, re, &nr);
    int64_t idx[9] = {0};
    int acc = (op == OP_ALL) ? 1 : 0;
    for (int64_t j = 0; j < p->red_numel; ++j) {
      int64_t off = 0;
      for (int k = 0; k < nr; ++k) off += idx[k] * rs[k];
      const int v = x[base + off] != 0;
      acc = (op == OP_ALL) ? (acc && v) : (acc || v);
      for (int k = nr - 1; k >= 0; --k) { if (++idx[k] < re[k]) break; idx[k] = 0; }
    }
    out[o] = (uint8_t)acc;
  }
}

/* Reduce exactly as phi::Reduce<CPUContext,...> / BoolReduceKernel would.
 * in_dtype/out_dtype use phi::DataType values; bf16/fp16 inputs must be widened to
 * fp32 by the caller (the CPU kernels for mean/max/argmax are not registered for
 * them and CPU bf16 sum accumulates in bf16: SURVEY.md probe P7).
 * Returns 0, or 1 for an invalid call (empty tensor, axis out of range), 2 unsupported. */
EXPORT int phi_oracle_reduce(int op, int in_dtype, int out_dtype, int rank, const int64_t* dims,
                             int naxes, const int32_t* axes, int reduce_all, const void* x,
                             void* out) {
  rplan p;
  if (make_plan(&p, rank, dims, naxes, axes, reduce_all)) return 1;
  if (!dt_size(in_dtype)) return 2;
  const int logical = (op == OP_ANY || op == OP_ALL);
  if (out_dtype == 0) out_dtype = logical ? DT_BOOL : in_dtype;
  if (!dt_size(out_dtype)) return 2;
  if (logical && out_dtype != DT_BOOL) return 1;
  int64_t numel = p.out_numel * p.red_numel;
  void* tmp = NULL;
  const void* src = x;
  if (in_dtype != out_dtype) {  /* phi::Cast<T>(x, out_dtype) first */
    tmp = malloc((size_t)numel * dt_size(out_dtype));
    if (!tmp) return 2;
    cast_buffer(x, in_dtype, tmp, out_dtype, numel);
    src = tmp;
  }
  int rc = 0;
  if (logical) reduce_bool(op, &p, (const uint8_t*)src, (uint8_t*)out);
  else switch (out_dtype) {
    case DT_F32: reduce_f32(op, &p, (const float*)src, (float*)out); break;
    case DT_F64: reduce_f64(op, &p, (const double*)src, (double*)out); break;
    case DT_INT32: reduce_i32(op, &p, (const int32_t*)src, (int32_t*)out); break;
    case DT_INT64: reduce_i64(op, &p, (const int64_t*)src, (int64_t*)out); break;
    case DT_INT16: reduce_i16(op, &p, (const int16_t*)src, (int16_t*)out); break;
    case DT_INT8: reduce_i8(op, &p, (const int8_t*)src, (int8_t*)out); break;
    case DT_UINT8: reduce_u8(op, &p, (const uint8_t*)src, (uint8_t*)out); break;
    case DT_BOOL: {  /* max/min/prod/sum on bool storage: treat as u8 then clamp to 0/1 */
      reduce_u8(op, &p, (const uint8_t*)src, (uint8_t*)out);
      for (int64_t i = 0; i < p.out_numel; ++i) ((uint8_t*)out)[i] = ((uint8_t*)out)[i] != 0;
      break;
    }
    default: rc = 2;
  }
  free(tmp);
  return rc;
}

/* ---- argmax / argmin ----------------------------------------------------------------- */
#define DEF_ARG(T, NAME, LOWEST, HIGHEST)                                                     \
  static void arg_##NAME(int is_max, const T* x, int64_t pre, int64_t n, int64_t post,        \
                         int64_t* out) {                                                      \
    for (int64_t a = 0; a < pre; ++a)                                                         \
      for (int64_t c = 0; c < post; ++c) {                                                    \
        int64_t best = 0;                                                                     \
        T bv = is_max ? (T)(LOWEST) : (T)(HIGHEST);                                           \
        for (int64_t k = 0; k < n; ++k) {                                                     \
          const T v = x[(a * n + k) * post + c];                                              \
          if (is_max ? (v > bv) : (v < bv)) { bv = v; best = k; }                             \
        }                                                                                     \
        out[a * post + c] = best;                                                             \
      }                                                                                       \
  }
DEF_ARG(float, f32, -3.402823466e+38f, 3.402823466e+38f)
DEF_ARG(double, f64, -1.7976931348623157e+308, 1.7976931348623157e+308)
DEF_ARG(int32_t, i32, INT32_MIN, INT32_MAX)
DEF_ARG(int64_t, i64, INT64_MIN, INT64_MAX)
DEF_ARG(int16_t, i16, INT16_MIN, INT16_MAX)
DEF_ARG(uint8_t, u8, 0, UINT8_MAX)

/* out is int64 [pre*post]; the caller narrows to int32 when the op asked for it
 * (Eigen's .cast<Tout>()). op: 0 argmax, 1 argmin. */
EXPORT int phi_oracle_argreduce(int op, int in_dtype, int rank, const int64_t* dims, int64_t axis,
                                int flatten, const void* x, int64_t* out) {
  if (rank < 0 || rank > 9) return 1;
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) numel *= dims[i];
  if (numel <= 0) return 1;
  int64_t pre = 1, n = numel, post = 1;
  if (!flatten && rank > 0) {
    if (axis < -rank || axis >= rank) return 1;
    if (axis < 0) axis += rank;
    pre = post = 1;
    for (int i = 0; i < axis; ++i) pre *= dims[i];
    n = dims[axis];
    for (int i = (int)axis + 1; i < rank; ++i) post *= dims[i];
  }
  const int is_max = (op == 0);
  switch (in_dtype) {
    case DT_F32: arg_f32(is_max, (const float*)x, pre, n, post, out); break;
    case DT_F64: arg_f64(is_max, (const double*)x, pre, n, post, out); break;
    case DT_INT32: arg_i32(is_max, (const int32_t*)x, pre, n, post, out); break;
    case DT_INT64: arg_i64(is_max, (const int64_t*)x, pre, n, post, out); break;
    case DT_INT16: arg_i16(is_max, (const int16_t*)x, pre, n, post, out); break;
    case DT_UINT8: arg_u8(is_max, (const uint8_t*)x, pre, n, post, out); break;
    default: return 2;
  }
  return 0;
}

/* ---- broadcast binary / fused add+activation ------------------------------------------- */
static double unary_d(int act, double v, float scale) {
  switch (act) {
    case 1: return v > 0 ? v : 0;
    case 2: return v * (double)scale;
    case 3: return 0.5 * v * (1.0 + erf(v * M_SQRT1_2));
    case 4: return 1.0 / (1.0 + exp(-v));
    case 5: return tanh(v);
    default: return v;
  }
}
static float unary_f(int act, float v, float scale) {
  switch (act) {
    case 1: return v > 0 ? v : 0;
    case 2: return v * scale;
    case 3: return 0.5f * v * (1.0f + erff(v * (float)M_SQRT1_2));
    case 4: return 1.0f / (1.0f + expf(-v));
    case 5: return tanhf(v);
    default: return v;
  }
}
#define BIN(op, a, b) ((op) == 0 ? (a) + (b) : (op) == 1 ? (a) - (b) : (op) == 2 ? (a) * (b) \
                       : (op) == 3 ? (a) / (b) : (op) == 4 ? ((a) > (b) ? (a) : (b)) : ((a) < (b) ? (a) : (b)))

/* in_dims[i*9 + k]: input i padded to out_rank (1 where it broadcasts) -- the arrays
 * GetBroadcastDimsArrays builds (common_shape.h:35-92). dtype: F32, F64, INT32, INT64. */
EXPORT int phi_oracle_elementwise(int binary, int act, int compound, float scale, int dtype,
                                  int arity, int out_rank, const int64_t* out_dims,
                                  const int64_t* in_dims, const void* const* ins, void* out,
                                  void* inter) {
  if (arity < 1 || arity > 3 || out_rank < 0 || out_rank > 9) return 1;
  int64_t numel = 1;
  for (int i = 0; i < out_rank; ++i) numel *= out_dims[i];
  for (int i = 0; i < arity; ++i)
    for (int k = 0; k < out_rank; ++k) {
      const int64_t a = in_dims[i * 9 + k];
      if (a != out_dims[k] && a != 1) return 1;  /* "Broadcast dimension mismatch." */
    }
  int64_t idx[9] = {0};
  for (int64_t n = 0; n < numel; ++n) {
    int64_t off[3] = {0, 0, 0};
    for (int i = 0; i < arity; ++i)
      for (int k = 0; k < out_rank; ++k) {
        const int64_t e = in_dims[i * 9 + k];
        off[i] = off[i] * e + (e == 1 ? 0 : idx[k]);
      }
#define EW_BODY(T, UN)                                                                        \
    {                                                                                         \
      T v0 = ((const T*)ins[0])[off[0]];                                                      \
      T v1 = arity > 1 ? ((const T*)ins[1])[off[1]] : (T)0;                                   \
      T v2 = arity > 2 ? ((const T*)ins[2])[off[2]] : (T)0;                                   \
      T t, r;                                                                                 \
      if (arity == 1) { t = v0; r = UN(act, t, scale); }                                      \
      else if (compound) { T u = UN(act, v1, scale); t = BIN(binary, v0, u); r = t; }         \
      else { t = BIN(binary, v0, v1); if (arity == 3) t = BIN(binary, t, v2); r = UN(act, t, scale); } \
      ((T*)out)[n] = r;                                                                       \
      if (inter) ((T*)inter)[n] = t;                                                          \
    }
#define UN_INT(act, v, scale) ((act) == 1 ? ((v) > 0 ? (v) : 0) : (v))
    switch (dtype) {
      case DT_F32: EW_BODY(float, unary_f) break;
      case DT_F64: EW_BODY(double, unary_d) break;
      case DT_INT32: EW_BODY(int32_t, UN_INT) break;
      case DT_INT64: EW_BODY(int64_t, UN_INT) break;
      default: return 2;
    }
    for (int k = out_rank - 1; k >= 0; --k) { if (++idx[k] < out_dims[k]) break; idx[k] = 0; }
  }
  return 0;
}

/* ---- row mean / variance: paddle.var's composite in fp32 (stat.py:184-197) ---------------- */
/* u = mean(x); out = sum((x-u)^2) / (n-1 if unbiased and n>1 else n). fp32 in, fp32 out;
 * both reductions are partial ones, i.e. plain in-order accumulation (see reduce above). */
EXPORT int phi_oracle_rowstats_f32(int64_t rows, int64_t cols, int unbiased, const float* x,
                                   float* mean, float* var) {
  if (rows <= 0 || cols <= 0) return 1;
  float* d = (float*)malloc((size_t)cols * sizeof(float));
  if (!d) return 2;
  for (int64_t r = 0; r < rows; ++r) {
    const float* row = x + r * cols;
    float su = 0.f;
    for (int64_t j = 0; j < cols; ++j) su += row[j];
    const float u = su / (float)cols;
    for (int64_t j = 0; j < cols; ++j) { const float t = row[j] - u; d[j] = t * t; }
    float s = 0.f;
    for (int64_t j = 0; j < cols; ++j) s += d[j];
    const float n = (float)((unbiased && cols > 1) ? cols - 1 : cols);
    if (mean) mean[r] = u;
    if (var) var[r] = s / n;
  }
  free(d);
  return 0;
}

/* ---- bf16 helpers (test inputs are widened before the fp32 oracle runs) ---------------- */
EXPORT void phi_oracle_bf16_to_f32(const uint16_t* src, float* dst, int64_t n) {
  for (int64_t i = 0; i < n; ++i) { uint32_t u = (uint32_t)src[i] << 16; memcpy(&dst[i], &u, 4); }
}

/* ---- multi-threaded driver for the CPU baseline (bench.py only) ---------------------------
 * phi's CPU kernels are single-threaded (Eigen::DefaultDevice, cpu_context.cc:40).  The
 * "all host cores" baseline splits independent output slices across OpenMP threads and
 * runs the same single-thread routine on each: column blocks for an axis-0 sum. */
EXPORT int phi_oracle_sum_axis0_f32_mt(int64_t rows, int64_t cols, const float* x, float* out,
                                       int threads) {
  if (rows <= 0 || cols <= 0) return 1;
  if (threads < 1) threads = 1;
  #pragma omp parallel for num_threads(threads) schedule(static)
  for (int t = 0; t < threads; ++t) {
    const int64_t c0 = cols * t / threads, c1 = cols * (t + 1) / threads;
    for (int64_t c = c0; c < c1; ++c) out[c] = 0.f;
    /* InnerMostDimPreserver order: for each output, in-order over the reduced index; walk rows
     * outermost so the memory stream is contiguous (same per-element order of additions) */
    for (int64_t r = 0; r < rows; ++r) {
      const float* row = x + r * cols;
      for (int64_t c = c0; c < c1; ++c) out[c] += row[c];
    }
  }
  return 0;
}

/* ---- functor plug-ins on the broadcast engine (b200phi_map) ------------------------------- */
/* Op ids follow b200phi_map_op (include/b200phi.h).  Restated from
 *   funcs/compare_functors.h:20-60   funcs/logical_functor.h:20-42   funcs/bitwise_functors.h:20-48
 *   funcs/elementwise_functor.h:173-179 (DivGradY), :478-531 (Max/MinGrad X/Y)
 *   gpu/reduce_kernel.cu:74-119 (min/max grad = Equal then Multiply)
 *   funcs/activation_functor.h (one formula per functor, quoted at each case)
 *   gpu/p_norm_kernel.cu:43-52 (UnsignedPowFunctor) */
static int equal_f32(float a, float b) {
  if (isinf(a) || isinf(b)) return a == b;
  if (isnan(a) || isnan(b)) return 0;
  return fabs((double)(a - b)) < 1e-8;
}
static int equal_f64(double a, double b) {
  if (isinf((float)a) || isinf((float)b)) return a == b;
  if (isnan((float)a) || isnan((float)b)) return 0;
  return fabs(a - b) < 1e-8;
}
#define MAP_ACT(T, SUF, NAME)                                                                     \
  static T NAME(int op, T x, float p0, float p1) {                                                \
    const T one = 1, zero = 0;                                                                    \
    switch (op) {                                                                                 \
      case 64: return exp##SUF(x);                                                                \
      case 65: return log##SUF(x);                                                                \
      case 66: return sqrt##SUF(x);                                                               \
      case 67: return one / sqrt##SUF(x);                                                         \
      case 68: return x * x;                                                                      \
      case 69: return one / x;                                                                    \
      case 70: return sin##SUF(x);                                                                \
      case 71: return cos##SUF(x);                                                                \
      case 72: return tanh##SUF(x);                                                               \
      case 73: return one / (one + exp##SUF(-x));                         /* sigmoid */            \
      case 74: return x / (one + exp##SUF(-x));                           /* silu */               \
      case 75: return x > zero ? x : (T)p0 * x;                           /* leaky_relu */         \
      case 76: return x > zero ? x : (T)p0 * (exp##SUF(x) - one);         /* elu */                \
      case 77: { T xb = x * (T)p0;                                        /* softplus */           \
                 return xb > (T)p1 ? x : log##SUF(one + exp##SUF(xb)) / (T)p0; }                  \
      case 78: return floor##SUF(x);                                                              \
      case 79: return ceil##SUF(x);                                                               \
      case 80: return expm1##SUF(x);                                                              \
      case 81: return log##SUF(one + x);                                  /* log1p: log(1 + x) */  \
      case 82: return log2##SUF(x);                                                               \
      case 83: return log10##SUF(x);                                                              \
      case 84: return x <= zero ? zero : (x < (T)p0 ? x : (T)p0);         /* relu6 */              \
      case 85: { T t = x * (T)p0 + (T)p1; T hi = t > zero ? t : zero;     /* hard_sigmoid */       \
                 return hi < one ? hi : one; }                                                    \
      case 86: return x / (one + exp##SUF(-(T)p0 * x));                   /* swish */              \
      case 87: { T sp = x > (T)p0 ? x : log##SUF(one + exp##SUF(x));      /* mish */               \
                 return x * tanh##SUF(sp); }                                                      \
      case 88: return x / (one + (x > -x ? x : -x));                      /* softsign */           \
      case 89: { T t = x > zero ? zero : -x;                              /* log_sigmoid */        \
                 return -t - log##SUF(exp##SUF(-t) + exp##SUF(-x - t)); }                         \
      case 90: return x < zero ? -x : x;                                                          \
      case 91: return pow##SUF(x < zero ? -x : x, (T)p0);                 /* |x|^p */              \
      case 92: return x > zero ? x : zero;                                                        \
      default: return x;                                                                          \
    }                                                                                             \
  }
MAP_ACT(float, f, map_act_f32)
MAP_ACT(double, , map_act_f64)

static int map_op_arity(int op) {
  if (op == 19 || op == 35 || op >= 64) return 1;
  return (op >= 48 && op <= 53) ? 3 : 2;
}

/* in_dims[i*9 + k]: operand i padded to out_rank.  in_dtype: BOOL, UINT8, INT8, INT16, INT32,
 * INT64, F32, F64.  out: BOOL bytes for op < 32, else in_dtype. */
EXPORT int phi_oracle_map(int op, int in_dtype, float p0, float p1, int arity, int out_rank,
                          const int64_t* out_dims, const int64_t* in_dims, const void* const* ins,
                          void* out) {
  if (arity != map_op_arity(op) || out_rank < 0 || out_rank > 9) return 1;
  const int flt = is_float_dt(in_dtype);
  if (op >= 32 && op <= 39 && flt) return 2;
  if (op >= 36 && op <= 39 && in_dtype == DT_BOOL) return 2;
  if ((op == 57 || (op >= 64 && op != 90 && op != 92)) && !flt) return 2;
  int64_t numel = 1;
  for (int i = 0; i < out_rank; ++i) numel *= out_dims[i];
  for (int i = 0; i < arity; ++i)
    for (int k = 0; k < out_rank; ++k) {
      const int64_t a = in_dims[i * 9 + k];
      if (a != out_dims[k] && a != 1) return 1;
    }
  int64_t idx[9] = {0};
  for (int64_t n = 0; n < numel; ++n) {
    int64_t off[3] = {0, 0, 0};
    for (int i = 0; i < arity; ++i)
      for (int k = 0; k < out_rank; ++k) {
        const int64_t e = in_dims[i * 9 + k];
        off[i] = off[i] * e + (e == 1 ? 0 : idx[k]);
      }
    if (flt) {
      /* operands in their own precision */
      double a = load_as_double(ins[0], in_dtype, off[0]);
      double b = arity > 1 ? load_as_double(ins[1], in_dtype, off[1]) : 0;
      double c = arity > 2 ? load_as_double(ins[2], in_dtype, off[2]) : 0;
      const int f32 = in_dtype == DT_F32;
      const int eq = f32 ? equal_f32((float)a, (float)b) : equal_f64(a, b);
      int bres = -1;
      double r = 0;
      switch (op) {
        case 0: bres = eq; break;
        case 1: bres = !eq; break;
        case 2: bres = a < b; break;
        case 3: bres = a <= b; break;
        case 4: bres = a > b; break;
        case 5: bres = a >= b; break;
        case 16: bres = (a != 0) && (b != 0); break;
        case 17: bres = (a != 0) || (b != 0); break;
        case 18: bres = (a != 0) != (b != 0); break;
        case 19: bres = !(a != 0); break;
        case 48: r = f32 ? (double)((float)c * (float)eq) : c * (double)eq; break;
        case 49: r = f32 ? (double)((float)c * (float)(a > b)) : c * (double)(a > b); break;
        case 50: r = f32 ? (double)((float)c * (float)(a <= b)) : c * (double)(a <= b); break;
        case 51: r = f32 ? (double)((float)c * (float)(a < b)) : c * (double)(a < b); break;
        case 52: r = f32 ? (double)((float)c * (float)(a >= b)) : c * (double)(a >= b); break;
        case 53:
          if (f32) { float fa = (float)a, fb = (float)b, fc = (float)c; r = -fa * ((fb / fc) / fc); }
          else r = -a * ((b / c) / c);
          break;
        case 56: r = eq; break;
        case 57: r = f32 ? (double)expf((float)a - (float)b) : exp(a - b); break;
        /* the arithmetic binary functors of funcs/elementwise_functor.h:190-290,540-734,844-963,
         * each in the operand's own precision */
        case 54: r = (a == 0) ? b : (double)(a > 0); break;
        case 58:
          if (f32) { float fb = (float)b, res = fmodf((float)a, fb); if (res != 0 && ((res < 0) != (fb < 0))) res += fb; r = res; }
          else { double res = fmod(a, b); if (res != 0 && ((res < 0) != (b < 0))) res += b; r = res; }
          break;
        case 59:
          if (f32) {
            float fa = (float)a, fb = (float)b;
            if (fb == 0) { r = fa / fb; break; }
            float mod = fmodf(fa, fb), dv = (fa - mod) / fb, fd;
            if ((mod != 0) && ((fb < 0) != (mod < 0))) dv -= 1.0f;
            if (dv != 0) { fd = floorf(dv); if (dv - fd > 0.5f) fd += 1.0f; } else fd = copysignf(0.0f, fa / fb);
            r = fd;
          } else {
            if (b == 0) { r = a / b; break; }
            double mod = fmod(a, b), dv = (a - mod) / b, fd;
            if ((mod != 0) && ((b < 0) != (mod < 0))) dv -= 1.0;
            if (dv != 0) { fd = floor(dv); if (dv - fd > 0.5) fd += 1.0; } else fd = copysign(0.0, a / b);
            r = fd;
          }
          break;
        case 60: r = f32 ? (double)powf((float)a, (float)b) : pow(a, b); break;
        case 61: r = f32 ? (double)copysignf((float)a, (float)b) : copysign(a, b); break;
        case 62: r = f32 ? (double)fmaxf((float)a, (float)b) : fmax(a, b); break;
        case 63: r = f32 ? (double)fminf((float)a, (float)b) : fmin(a, b); break;
        default:
          r = f32 ? (double)map_act_f32(op, (float)a, p0, p1) : map_act_f64(op, a, p0, p1);
          break;
      }
      if (bres >= 0) ((uint8_t*)out)[n] = (uint8_t)bres;
      else if (f32) ((float*)out)[n] = (float)r;
      else ((double*)out)[n] = r;
    } else {
      int64_t a = load_as_i64(ins[0], in_dtype, off[0]);
      int64_t b = arity > 1 ? load_as_i64(ins[1], in_dtype, off[1]) : 0;
      int64_t c = arity > 2 ? load_as_i64(ins[2], in_dtype, off[2]) : 0;
      int bres = -1;
      int64_t r = 0;
      switch (op) {
        case 0: bres = a == b; break;
        case 1: bres = a != b; break;
        case 2: bres = a < b; break;
        case 3: bres = a <= b; break;
        case 4: bres = a > b; break;
        case 5: bres = a >= b; break;
        case 16: bres = (a != 0) && (b != 0); break;
        case 17: bres = (a != 0) || (b != 0); break;
        case 18: bres = (a != 0) != (b != 0); break;
        case 19: bres = !(a != 0); break;
        case 32: r = a & b; break;
        case 33: r = a | b; break;
        case 34: r = a ^ b; break;
        case 35: r = in_dtype == DT_BOOL ? !a : ~a; break;
        /* shifts, funcs/bitwise_functors.h:49-206: count outside [0, bits) gives 0, or the sign
         * fill for the arithmetic right shift of a signed type; the logical right shift works on
         * the unsigned image of the storage type; the store below truncates to the storage type */
        case 36: case 37: {
          const int bits = (int)dt_size(in_dtype) * 8;
          r = (b < 0 || b >= bits) ? 0 : (int64_t)((uint64_t)a << b);
          break;
        }
        case 38: {
          const int bits = (int)dt_size(in_dtype) * 8;
          if (b < 0 || b >= bits) r = (in_dtype != DT_UINT8 && a < 0) ? -1 : 0;
          else r = a >> b;
          break;
        }
        case 39: {
          const int bits = (int)dt_size(in_dtype) * 8;
          const uint64_t image = bits == 64 ? (uint64_t)a : ((uint64_t)a & ((1ull << bits) - 1));
          r = (b < 0 || b >= bits) ? 0 : (int64_t)(image >> b);
          break;
        }
        case 48: r = c * (a == b); break;
        case 49: r = c * (a > b); break;
        case 50: r = c * (a <= b); break;
        case 51: r = c * (a < b); break;
        case 52: r = c * (a >= b); break;
        case 53: r = c == 0 ? 0 : -a * ((b / c) / c); break;
        case 56: r = a == b; break;
        case 54: r = (a == 0) ? b : (a > 0); break;
        case 58:  /* the remainder takes the divisor's sign; b == 0 is an assert in the reference */
          if (b == 0) r = 0;
          else { r = a % b; if (r != 0 && ((b ^ r) < 0)) r += b; }
          break;
        case 59:
          if (b == 0) r = 0;
          else if ((a < 0) != (b < 0)) { const int64_t q = a / b, rem = a % b; r = rem ? q - 1 : q; }
          else r = a / b;
          break;
        case 60: r = (int64_t)pow((double)a, (double)b); break;       /* host branch of compute_pow */
        case 61: r = (int64_t)copysign((double)a, (double)b); break;
        case 62:  /* FMaxFunctor<int> goes through float, <int64_t> through double */
          r = in_dtype == DT_INT32 ? (int64_t)lrintf(fmaxf((float)a, (float)b)) : (int64_t)llrint(fmax((double)a, (double)b));
          break;
        case 63:
          r = in_dtype == DT_INT32 ? (int64_t)lrintf(fminf((float)a, (float)b)) : (int64_t)llrint(fmin((double)a, (double)b));
          break;
        case 90: r = a < 0 ? -a : a; break;
        case 92: r = a > 0 ? a : 0; break;
        default: return 2;
      }
      if (bres >= 0) ((uint8_t*)out)[n] = (uint8_t)bres;
      else switch (in_dtype) {
        case DT_BOOL: ((uint8_t*)out)[n] = r != 0; break;
        case DT_UINT8: ((uint8_t*)out)[n] = (uint8_t)r; break;
        case DT_INT8: ((int8_t*)out)[n] = (int8_t)r; break;
        case DT_INT16: ((int16_t*)out)[n] = (int16_t)r; break;
        case DT_INT32: ((int32_t*)out)[n] = (int32_t)r; break;
        default: ((int64_t*)out)[n] = r; break;
      }
    }
    for (int k = out_rank - 1; k >= 0; --k) { if (++idx[k] < out_dims[k]) break; idx[k] = 0; }
  }
  return 0;
}
