/*
 * b200phi.h -- C ABI of the B200-native phi reduce / broadcast hot path.
 *
 * Every entry point is plain C: pointers, sizes and small POD descriptors.
 * No phi, torch or C++ types cross this boundary.  All device work is enqueued
 * asynchronously on the caller's stream (a cudaStream_t passed as void*); no
 * entry point synchronises the stream, allocates device memory or keeps a
 * pointer after it returns.  Scratch memory, where a kernel needs some, is
 * supplied by the caller (size from the matching *_workspace query), exactly
 * like the reference obtains temporaries from dev_ctx.Alloc
 * (paddle/phi/kernels/funcs/reduce_function.h:1118-1140).
 *
 * Return value: 0 on success, a b200phi_status otherwise; the message is
 * available per host thread through b200phi_last_error().  The reference
 * raises EnforceNotMet C++ exceptions for the same conditions
 * (reduce_function.h:1052-1055, dims_simplifier.h:116-126); the phi adapter
 * (adapter/phi_b200_kernels.cu) turns a non-zero status back into PADDLE_THROW.
 *
 * Tensors are contiguous row-major with int64 extents and rank <= 9
 * (paddle/common/ddim.h:71, paddle/phi/core/dense_tensor.h:43-199).
 */
#ifndef B200PHI_H_
#define B200PHI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200PHI_API __attribute__((visibility("default")))
#else
#define B200PHI_API
#endif

#define B200PHI_MAX_RANK 9
#define B200PHI_MAX_ARITY 3

/* Numeric values equal phi::DataType (paddle/phi/common/data_type.h:42-85) so
 * the adapter can pass static_cast<int>(tensor.dtype()) straight through. */
typedef enum {
  B200PHI_UNDEFINED = 0,
  B200PHI_BOOL = 1,
  B200PHI_UINT8 = 2,
  B200PHI_INT8 = 3,
  B200PHI_INT16 = 5,
  B200PHI_INT32 = 7,
  B200PHI_INT64 = 9,
  B200PHI_FLOAT32 = 10,
  B200PHI_FLOAT64 = 11,
  /* SUM / MEAN reductions and ADD / SUBTRACT only (kps/reduce_kernel.cu:382-397,
   * kps/elementwise_kernel.cu:304-381): component-wise on the (re, im) pairs */
  B200PHI_COMPLEX64 = 12,
  B200PHI_COMPLEX128 = 13,
  B200PHI_FLOAT16 = 15,
  B200PHI_BFLOAT16 = 16,
  /* MAX / AMAX reductions only (the reference registers `max` for them, kps/reduce_kernel.cu:345-355) */
  B200PHI_FLOAT8_E4M3FN = 17,
  B200PHI_FLOAT8_E5M2 = 18
} b200phi_dtype;

typedef enum {
  B200PHI_OK = 0,
  B200PHI_INVALID_ARGUMENT = 1, /* EnforceNotMet(InvalidArgument) in phi */
  B200PHI_UNSUPPORTED = 2,      /* dtype / op combination not built */
  B200PHI_CUDA_ERROR = 3,
  B200PHI_WORKSPACE_TOO_SMALL = 4,
  B200PHI_NCCL_ERROR = 5,
  B200PHI_PEER_TIMEOUT = 6 /* b200phi_sharded_reduce_p2p: a peer never published its partial */
} b200phi_status;

/* ------------------------------------------------------------------------ */
/* Reductions                                                               */
/* ------------------------------------------------------------------------ */

/* Replaces the op binding done per kernel in
 * paddle/phi/kernels/kps/reduce_kernel.cu:35-133,185-271 and
 * paddle/phi/kernels/legacy/kps/reduce_max_kernel.cu:22-33. */
typedef enum {
  B200PHI_SUM = 0,  /* kps::AddFunctor + IdentityFunctor            */
  B200PHI_MEAN = 1, /* sum in MPType, then / reduce_num (IsMean)    */
  B200PHI_MAX = 2,  /* kps::MaxFunctor                              */
  B200PHI_MIN = 3,  /* kps::MinFunctor                              */
  B200PHI_PROD = 4, /* kps::MulFunctor                              */
  B200PHI_ANY = 5,  /* kps::LogicalOrFunctor on (x != 0), out BOOL  */
  B200PHI_ALL = 6,  /* kps::LogicalAndFunctor on (x != 0), out BOOL */
  B200PHI_AMAX = 7, /* same arithmetic as MAX (differs only in grad) */
  B200PHI_AMIN = 8,
  /* Reductions over a transformed element: the TransformOp slot of funcs::ReduceKernel as
   * p_norm and squared_l2_norm fill it (gpu/p_norm_kernel.cu:26-52,76-95,
   * gpu/squared_l2_norm_kernel.cu:24-34).  Floating-point dtypes, out dtype = in dtype. */
  B200PHI_SUM_ABS = 9,        /* sum |x|            p_norm porder 1 */
  B200PHI_SUM_SQUARE = 10,    /* sum x*x            squared_l2_norm */
  B200PHI_L2_NORM = 11,       /* sqrt(sum x*x)      p_norm porder 2 */
  B200PHI_MAX_ABS = 12,       /* max |x|            p_norm porder +inf */
  B200PHI_MIN_ABS = 13,       /* min |x|            p_norm porder -inf */
  B200PHI_COUNT_NONZERO = 14  /* sum (x != 0)       p_norm porder 0 (NonzeroFunctor) */
} b200phi_reduce_op;

/* One reduction call.  Stands in for the argument list of
 *   phi::SumKernel(ctx, x, IntArray dims, DataType out_dtype, bool keep_dim, out)
 *   (paddle/phi/kernels/reduce_sum_kernel.h:22-37) and its siblings
 *   MeanKernel/MaxKernel/MinKernel/ProdKernel/AnyKernel/AllKernel/AMax/AMin,
 * after InferMeta has fixed out's dims/dtype.  keep_dim is not needed: it only
 * changes out's shape, never its linear layout. */
typedef struct {
  int32_t op;         /* b200phi_reduce_op */
  int32_t in_dtype;   /* b200phi_dtype of x */
  int32_t out_dtype;  /* UNDEFINED => in_dtype (ANY/ALL: BOOL) */
  int32_t rank;       /* 0..9; 0 => 0-D tensor with one element */
  int64_t dims[B200PHI_MAX_RANK];
  int32_t naxes;      /* 0 => reduce over every axis */
  int32_t axes[B200PHI_MAX_RANK]; /* each in [-rank, rank) */
  int32_t reduce_all; /* forces a full reduction, as the phi attribute does */
  const void* x;      /* device pointer, element-aligned */
  void* out;          /* device pointer, element-aligned */
  void* workspace;    /* device scratch, >= b200phi_reduce_workspace() bytes, 256-B aligned; may be NULL if 0 */
  size_t workspace_bytes;
} b200phi_reduce_desc;

/* Bytes of device scratch b200phi_reduce needs for this descriptor on the
 * current device (0 for most shapes; non-zero when the reduced extent is split
 * across thread blocks and finalised by the last block).  Replaces the
 * temporary the reference allocates for its two-pass reduce
 * (reduce_function.h:1118-1140, CubTensorReduce :938-969). Returns 0 on a bad
 * descriptor; b200phi_reduce then reports the error. */
B200PHI_API size_t b200phi_reduce_workspace(const b200phi_reduce_desc* d);

/* Replaces funcs::ReduceKernel<Tx,Ty,ReduceOp,TransformOp,IsMean>
 * (reduce_function.h:1047-1198) together with phi::Reduce (gpu/reduce.h:32-80);
 * the dtype change the reference does as a separate Cast pass (gpu/reduce.h:48-69)
 * is fused into the load. */
B200PHI_API int b200phi_reduce(const b200phi_reduce_desc* d, void* stream);

/* Diagnostic: which kernel family b200phi_reduce would launch for this descriptor on the current
 * device (the ReduceKind of csrc/reduce_plan.h: 0/1 row, 2 split row, 3 column, 4 scalar index
 * walk, 5/7/8/9 short-row forms, 6 two-stage, 10/11 tiled general patterns), or -1 for a bad
 * descriptor.  Tests and benchmarks use it to assert that a shape runs on the path they mean to
 * measure; the reference exposes the same decision as ReduceConfig::reduce_type
 * (reduce_function.h:410-431).  Pure host logic. */
B200PHI_API int b200phi_reduce_plan_kind(const b200phi_reduce_desc* d);

/* No counterpart in the reference.  The same reductions as b200phi_reduce (any shape: column
 * reduces, few very long rows, two-stage [A,K,B] patterns -- the ones that need scratch -- and
 * all the others), for hosts that cannot hand over a workspace because they may not, or
 * cannot, allocate device memory where the call is made: a C/Go service linking this
 * library with no device allocator of its own, a server whose memory plan is frozen before
 * the first request, a callback that must not take the allocator lock.  The scratch is a
 * fixed B200PHI_NOALLOC_SCRATCH_BYTES array inside the library (every device that loads the
 * library holds it until exit); d->workspace is ignored.  A reduce that needs more scratch is
 * refused with B200PHI_WORKSPACE_TOO_SMALL.  The first call on a device resolves the
 * array's address and should be made before any stream capture on that device.  All calls
 * on a device share the array: calls that overlap on the GPU race, keep them on one stream. */
#define B200PHI_NOALLOC_SCRATCH_BYTES (4u << 20)
B200PHI_API int b200phi_reduce_noalloc(const b200phi_reduce_desc* d, void* stream);

/* ------------------------------------------------------------------------ */
/* argmax / argmin                                                          */
/* ------------------------------------------------------------------------ */

typedef enum { B200PHI_ARGMAX = 0, B200PHI_ARGMIN = 1 } b200phi_arg_op;

/* Stands in for
 *   phi::ArgMaxKernel(ctx, x, Scalar axis, bool keepdims, bool flatten, DataType dtype, out)
 *   (paddle/phi/kernels/arg_min_max_kernel.h:22-38; GPU body
 *    paddle/phi/kernels/gpu/arg_min_max_kernel.cu:58-254).
 * Ties resolve to the lowest index, a slice whose every element equals the
 * identity yields 0 (the CPU kernel's answer, cpu/arg_min_max_kernel.cc:46-59). */
typedef struct {
  int32_t op;          /* b200phi_arg_op */
  int32_t in_dtype;
  int32_t index_dtype; /* INT64 (default when UNDEFINED) or INT32 */
  int32_t rank;
  int64_t dims[B200PHI_MAX_RANK];
  int64_t axis;        /* in [-rank, rank); ignored when flatten */
  int32_t flatten;
  const void* x;
  void* out;
  void* max_out;       /* optional: also write the extreme value (in_dtype); NULL to skip */
  void* workspace;     /* device scratch, >= b200phi_argreduce_workspace() bytes, 256-B aligned; may be NULL if 0 */
  size_t workspace_bytes;
} b200phi_argreduce_desc;

/* Bytes of device scratch b200phi_argreduce needs for this descriptor on the current device: 0
 * when there is a thread-block group per slice (many rows reduced over the last axis); non-zero
 * when the reduced extent is cut across thread blocks -- `flatten` (paddle.argmax's default
 * axis=None: the whole tensor is ONE slice), few long rows, or a non-last axis -- and the block
 * that finishes last combines the (value, index) partials.  ComputeFullArg
 * (gpu/arg_min_max_kernel.cu:85-145) has no counterpart: it gives every slice one block.  The
 * answer does not depend on d->x.  Returns 0 on a bad descriptor; b200phi_argreduce reports it. */
B200PHI_API size_t b200phi_argreduce_workspace(const b200phi_argreduce_desc* d);

B200PHI_API int b200phi_argreduce(const b200phi_argreduce_desc* d, void* stream);

/* ------------------------------------------------------------------------ */
/* Row statistics: mean and variance over the last axis in one read         */
/* ------------------------------------------------------------------------ */

/* Replaces the Python composite paddle.var = mean -> subtract -> pow -> sum ->
 * divide (python/paddle/tensor/stat.py:184-197) and the MeanRawKernel it starts
 * with (kps/reduce_kernel.cu:110-120) for a [rows, cols] view reduced over
 * cols.  var = sum((x-mean)^2) / (cols - unbiased), accumulated in fp32
 * (fp64 for FLOAT64 input). */
typedef struct {
  int32_t in_dtype;  /* FLOAT32, FLOAT64, BFLOAT16, FLOAT16 */
  int32_t out_dtype; /* UNDEFINED => in_dtype; FLOAT32 allowed for 16-bit inputs */
  int64_t rows;
  int64_t cols;
  int32_t unbiased;  /* 1 => divide by cols-1 (paddle.var default) */
  const void* x;
  void* mean;        /* [rows] or NULL */
  void* var;         /* [rows] or NULL */
} b200phi_rowstats_desc;

B200PHI_API int b200phi_rowstats(const b200phi_rowstats_desc* d, void* stream);

/* Row logsumexp: out[r] = log(sum_j exp(x[r][j])) over the last (merged trailing) axes, computed
 * as max + log(sum exp(x - max)) in one read of x.  Stands in for phi::LogsumexpKernel
 * (paddle/phi/kernels/gpu/logsumexp_kernel.cu:81-165) when the reduced axes are the trailing
 * ones (or all of them): the caller flattens x to [rows, cols].  For other axis sets the host
 * composes max -> exp(x - max) -> sum -> log + max from b200phi_reduce / b200phi_map, which is
 * the reference's own fallback (:49-78).  fp16/bf16 accumulate in fp32. */
typedef struct {
  int32_t dtype; /* FLOAT32, FLOAT64, BFLOAT16, FLOAT16; out has the same dtype */
  int64_t rows;
  int64_t cols;
  const void* x;
  void* out;     /* [rows] */
} b200phi_logsumexp_desc;

B200PHI_API int b200phi_logsumexp(const b200phi_logsumexp_desc* d, void* stream);

/* ------------------------------------------------------------------------ */
/* n-ary broadcast / elementwise engine with fused unary epilogue           */
/* ------------------------------------------------------------------------ */

/* Binary functors of paddle/phi/kernels/funcs/elementwise_functor.h. */
typedef enum {
  B200PHI_ADD = 0,      /* AddFunctor :39-41 */
  B200PHI_SUBTRACT = 1,
  B200PHI_MULTIPLY = 2,
  B200PHI_DIVIDE = 3,
  B200PHI_MAXIMUM = 4,
  B200PHI_MINIMUM = 5
} b200phi_binary_op;

/* Unary functors that can be fused around the binary one, as
 * fused_elemwise_activation's functor_list does
 * (paddle/phi/kernels/funcs/fused_elemwise_activation_functor.h:388-458). */
typedef enum {
  B200PHI_ACT_NONE = 0,
  B200PHI_ACT_RELU = 1,  /* v > 0 ? v : 0   (ReluCPUFunctor, activation_functor.h:1473-1480) */
  B200PHI_ACT_SCALE = 2, /* v * scale       (ScaleFunctor, functors.h) */
  B200PHI_ACT_GELU = 3,  /* 0.5 v (1 + erf(v / sqrt 2)) */
  B200PHI_ACT_SIGMOID = 4,
  B200PHI_ACT_TANH = 5
} b200phi_unary_op;

typedef struct {
  int32_t rank;
  int64_t dims[B200PHI_MAX_RANK];
  const void* data;
} b200phi_tensor_view;

/* One elementwise call.  Stands in for
 *   funcs::BroadcastKernel<T>(ctx, ins, outs, functor, axis)  (broadcast_function.h:758-824)
 *   funcs::ElementwiseKernel                                  (elementwise_base.h:807-845)
 *   phi::AddRawKernel(ctx, x, y, axis, out) and siblings      (impl/elementwise_kernel_impl.h:50-62)
 *   phi::FusedElemwiseAddActivationKernel                     (impl/fused_elemwise_activation_kernel_impl.h:202-222)
 * compound == 0:  out = act(binary(in0, in1 [, in2]))    e.g. ["relu","elementwise_add"]
 * compound == 1:  out = binary(in0, act(in1))            e.g. ["elementwise_add","relu"]
 * With three inputs the functor folds left: binary(binary(in0,in1),in2)
 * (test/cpp/phi/kernels/test_ternary_broadcast.cu:96-120).
 * axis: -1 => numpy right-aligned broadcasting; >= 0 => the legacy left offset
 * of the lower-rank operand (paddle/phi/kernels/funcs/common_shape.h:35-92). */
typedef struct {
  int32_t binary;   /* b200phi_binary_op */
  int32_t act;      /* b200phi_unary_op */
  int32_t compound; /* see above */
  float scale;      /* for B200PHI_ACT_SCALE */
  int32_t dtype;    /* all inputs and the output share it */
  int32_t arity;    /* 1 (act only), 2 or 3 */
  int32_t axis;
  b200phi_tensor_view in[B200PHI_MAX_ARITY];
  int32_t out_rank; /* out dims as InferMeta fixed them (binary.cc:1665-1740) */
  int64_t out_dims[B200PHI_MAX_RANK];
  void* out;
  void* intermediate_out; /* optional save_intermediate_out; NULL to skip */
} b200phi_elementwise_desc;

B200PHI_API int b200phi_elementwise(const b200phi_elementwise_desc* d, void* stream);

/* ------------------------------------------------------------------------ */
/* Functor plug-ins on the broadcast engine          */
/* ------------------------------------------------------------------------ */

/* The other clients of funcs::BroadcastKernel / funcs::ElementwiseKernel: one launch,
 * same broadcasting rules and the same [rows, inner] walk as b200phi_elementwise, but the
 * output dtype may differ from the input dtype (compare / logical write BOOL) and the
 * functor may take three operands.  in[] dtype is in_dtype for every operand. */
typedef enum {
  /* out BOOL; funcs/compare_functors.h:20-60 (kps/compare_kernel.cu).  EQUAL on floating
   * point is the reference's tolerant form: |a-b| < 1e-8 unless either is inf (then ==)
   * or NaN (then false). */
  B200PHI_MAP_EQUAL = 0,
  B200PHI_MAP_NOT_EQUAL = 1,
  B200PHI_MAP_LESS_THAN = 2,
  B200PHI_MAP_LESS_EQUAL = 3,
  B200PHI_MAP_GREATER_THAN = 4,
  B200PHI_MAP_GREATER_EQUAL = 5,
  /* out BOOL, operands by truthiness; funcs/logical_functor.h:20-42 (kps/logical_kernel.cu) */
  B200PHI_MAP_LOGICAL_AND = 16,
  B200PHI_MAP_LOGICAL_OR = 17,
  B200PHI_MAP_LOGICAL_XOR = 18,
  B200PHI_MAP_LOGICAL_NOT = 19, /* arity 1 */
  /* integer and bool dtypes, out = in dtype; funcs/bitwise_functors.h:20-48
   * (kps/bitwise_kernel.cu); on BOOL they are the logical forms, NOT is !a */
  B200PHI_MAP_BITWISE_AND = 32,
  B200PHI_MAP_BITWISE_OR = 33,
  B200PHI_MAP_BITWISE_XOR = 34,
  B200PHI_MAP_BITWISE_NOT = 35, /* arity 1 */
  /* integer dtypes except BOOL; funcs/bitwise_functors.h:49-206 (bitwise_left_shift /
   * bitwise_right_shift with is_arithmetic).  A shift count that is negative or >= the bit width
   * gives 0, except the arithmetic right shift of a signed type, which gives the sign fill. */
  B200PHI_MAP_LEFT_SHIFT_ARITHMETIC = 36,
  B200PHI_MAP_LEFT_SHIFT_LOGIC = 37,
  B200PHI_MAP_RIGHT_SHIFT_ARITHMETIC = 38,
  B200PHI_MAP_RIGHT_SHIFT_LOGIC = 39,
  /* three operands, out = in dtype: the functors the backward kernels feed to the
   * broadcast engine (funcs/elementwise_functor.h:173-179, 478-531; gpu/reduce_kernel.cu:74-119) */
  B200PHI_MAP_EQUAL_MUL = 48,  /* in2 * T(EQUAL(in0, in1))   max_grad / min_grad: (x, out, dout) */
  B200PHI_MAP_MAX_GRAD_X = 49, /* in2 * T(in0 >  in1)        (x, y, dout) */
  B200PHI_MAP_MAX_GRAD_Y = 50, /* in2 * T(in0 <= in1) */
  B200PHI_MAP_MIN_GRAD_X = 51, /* in2 * T(in0 <  in1) */
  B200PHI_MAP_MIN_GRAD_Y = 52, /* in2 * T(in0 >= in1) */
  B200PHI_MAP_DIV_GRAD_Y = 53, /* -in0 * ((in1 / in2) / in2)  (dout, out, y) */
  /* two operands, out = in dtype */
  B200PHI_MAP_HEAVISIDE = 54,  /* in0 == 0 ? in1 : T(in0 > 0)   ElementwiseHeavisideFunctor, elementwise_functor.h:614-619 */
  B200PHI_MAP_EQUAL_T = 56,    /* T(EQUAL(in0, in1)): the mask amax_grad / amin_grad count */
  B200PHI_MAP_SUB_EXP = 57,    /* exp(in0 - in1): logsumexp's second pass (gpu/logsumexp_kernel.cu:64-67) */
  /* the other binary functors registered by kps/elementwise_kernel.cu:183-293 and
   * legacy/kps/elementwise_kernel.cu:139-165 (funcs/elementwise_functor.h:190-290, 540-734, 844-963).
   * 16-bit floats compute in fp32.  Integer REMAINDER / FLOOR_DIVIDE by zero give 0 (the
   * reference asserts there).  FMAX / FMIN on INT32 go through fp32 and on INT64 through fp64, as
   * the reference's specialisations do; integer POW is llrint(pow(double, double)); integer
   * COPYSIGN goes through fp64. */
  B200PHI_MAP_REMAINDER = 58,    /* result takes the divisor's sign */
  B200PHI_MAP_FLOOR_DIVIDE = 59,
  B200PHI_MAP_POW = 60,          /* elementwise_pow */
  B200PHI_MAP_COPYSIGN = 61,
  B200PHI_MAP_FMAX = 62,         /* NaN-ignoring maximum */
  B200PHI_MAP_FMIN = 63,
  /* one operand, floating point, computed in fp32 for 16-bit dtypes (MPType); the Cuda*Functor
   * forms of funcs/activation_functor.h driven through ElementwiseKernel by
   * gpu/activation_kernel.cu.  p0 / p1 are the functor's attributes. */
  B200PHI_MAP_EXP = 64,
  B200PHI_MAP_LOG = 65,
  B200PHI_MAP_SQRT = 66,
  B200PHI_MAP_RSQRT = 67,
  B200PHI_MAP_SQUARE = 68,
  B200PHI_MAP_RECIPROCAL = 69,
  B200PHI_MAP_SIN = 70,
  B200PHI_MAP_COS = 71,
  B200PHI_MAP_TANH = 72,
  B200PHI_MAP_SIGMOID = 73,
  B200PHI_MAP_SILU = 74,
  B200PHI_MAP_LEAKY_RELU = 75, /* p0 = alpha */
  B200PHI_MAP_ELU = 76,        /* p0 = alpha */
  B200PHI_MAP_SOFTPLUS = 77,   /* p0 = beta, p1 = threshold */
  B200PHI_MAP_FLOOR = 78,
  B200PHI_MAP_CEIL = 79,
  B200PHI_MAP_EXPM1 = 80,
  B200PHI_MAP_LOG1P = 81,
  B200PHI_MAP_LOG2 = 82,
  B200PHI_MAP_LOG10 = 83,
  B200PHI_MAP_RELU6 = 84,        /* p0 = threshold */
  B200PHI_MAP_HARD_SIGMOID = 85, /* p0 = slope, p1 = offset */
  B200PHI_MAP_SWISH = 86,        /* p0 = beta */
  B200PHI_MAP_MISH = 87,         /* p0 = threshold */
  B200PHI_MAP_SOFTSIGN = 88,
  B200PHI_MAP_LOG_SIGMOID = 89,
  B200PHI_MAP_ABS = 90,          /* gpu/abs_kernel.cu; integer dtypes allowed */
  B200PHI_MAP_POW_ABS = 91,      /* |x|^p0: p_norm's UnsignedPowFunctor (gpu/p_norm_kernel.cu:43-52) */
  B200PHI_MAP_RELU = 92          /* CudaReluFunctor: x > 0 ? x : 0 */
} b200phi_map_op;

typedef struct {
  int32_t op;        /* b200phi_map_op */
  int32_t in_dtype;  /* every operand */
  int32_t out_dtype; /* UNDEFINED => what the op implies; otherwise it must be that */
  int32_t arity;     /* must be the op's operand count */
  int32_t axis;      /* as in b200phi_elementwise_desc */
  float p0, p1;      /* functor attributes */
  b200phi_tensor_view in[B200PHI_MAX_ARITY];
  int32_t out_rank;
  int64_t out_dims[B200PHI_MAX_RANK];
  void* out;
} b200phi_map_desc;

B200PHI_API int b200phi_map(const b200phi_map_desc* d, void* stream);

/* Mixed-precision add: out (FLOAT32) = x (FLOAT32) + widen(y), y BFLOAT16 or FLOAT16, numpy
 * broadcasting.  The branch phi::AddKernel takes for a float32 x and a 16-bit y
 * (MultiPrecisionAddKernelImpl, paddle/phi/kernels/kps/elementwise_kernel.cu:54-99): the
 * master-gradient accumulation of mixed-precision training, without a cast pass over y. */
typedef struct {
  int32_t y_dtype;        /* BFLOAT16 or FLOAT16; x and out are FLOAT32 */
  b200phi_tensor_view x;
  b200phi_tensor_view y;
  int32_t out_rank;
  int64_t out_dims[B200PHI_MAX_RANK];
  void* out;              /* may alias x.data when x is not broadcast (in-place accumulate) */
} b200phi_mixed_add_desc;

B200PHI_API int b200phi_mixed_add(const b200phi_mixed_add_desc* d, void* stream);

/* ------------------------------------------------------------------------ */
/* Sharded reduction: local reduce + allreduce of the partial               */
/* ------------------------------------------------------------------------ */

/* Communicator helpers.  Replace the lazy ncclGetUniqueId / ncclCommInitRank
 * that CommContextManager does over a TCPStore
 * (paddle/phi/core/distributed/comm_context_manager.cc:75,
 *  nccl_comm_context.cc:37-46).  The caller moves the 128-byte id between
 * processes with whatever rendezvous it has. NCCL is dlopen'ed on first use,
 * as phi::dynload does. */
B200PHI_API int b200phi_comm_unique_id(unsigned char id[128]);
B200PHI_API int b200phi_comm_init_rank(const unsigned char id[128], int nranks, int rank, void** comm);
B200PHI_API int b200phi_comm_destroy(void* comm);

/* Each rank holds a contiguous block of rows of the global tensor (axis 0
 * sharded); d describes the LOCAL shard.  The local partial is reduced into
 * d->out and combined across ranks on the same stream with no host sync:
 * ncclSum for SUM/MEAN (MEAN divides by global_reduce_count afterwards),
 * ncclMax/Min/Prod for MAX/MIN/PROD.  Replaces the reference's
 * SumKernel + ProcessGroupNCCL::AllReduce pair
 * (process_group_nccl.cc:267-298 -> nccl_comm_context.cc:184-206) and the
 * Partial->Replicate reshard (p_to_r_reshard_function.cc:67). With comm ==
 * NULL (single GPU) it is b200phi_reduce and issues no collective. */
B200PHI_API int b200phi_sharded_reduce(const b200phi_reduce_desc* d, int64_t global_reduce_count,
                           void* comm, void* stream);

/* Scratch b200phi_sharded_reduce needs in d->workspace: b200phi_reduce_workspace(d), plus --
 * when the result is a SUM / MEAN into FLOAT16 / BFLOAT16 -- room for the rank's partial in
 * FLOAT32.  Such a partial is the reference's MPType (amp_type_traits.h:25-53): it is all-reduced
 * in fp32 and rounded to the 16-bit type once, after the ranks are combined, where an allreduce
 * in the output dtype would round once per rank and once per NCCL step.  Every argument is
 * checked before anything is enqueued, so a rank that gets an error has launched nothing. */
B200PHI_API size_t b200phi_sharded_reduce_workspace(const b200phi_reduce_desc* d);

/* argmax / argmin over the SHARDED leading axis.  d describes the local shard and must reduce
 * axis 0; row_offset is the global index of the shard's first row.  Each rank reduces its rows to
 * (value, global index) per column, the small pairs are all-gathered (NCCL has no arg-reduce)
 * and every rank selects with the reference's tie rule: better value, else lowest index
 * (cub::ArgMax as used by gpu/arg_min_max_kernel.cu:58-83).  d->out / d->max_out
 * receive the replicated result.  nranks <= 1 or comm == NULL: plain b200phi_argreduce. */
B200PHI_API size_t b200phi_sharded_argreduce_workspace(const b200phi_argreduce_desc* d, int nranks);
B200PHI_API int b200phi_sharded_argreduce(const b200phi_argreduce_desc* d, int64_t row_offset,
                                          int nranks, void* comm, void* workspace,
                                          size_t workspace_bytes, void* stream);

/* ---- without a collective library: local reduce + exchange over NVLink peer memory ---- */

/* Replaces the same reference pair as b200phi_sharded_reduce
 * (SumKernel + ProcessGroupNCCL::AllReduce, process_group_nccl.cc:267-298).  The local reduce
 * (the unchanged kernels of b200phi_reduce) writes the rank's partial -- in the accumulation
 * type, so 16-bit float sums are rounded once -- into the rank's own exchange buffer; a small
 * second kernel, launched with programmatic dependent launch so that its launch latency hides
 * under the reduce, pushes the partial into every peer's buffer (CUDA IPC mapping), publishes an
 * epoch flag, waits for the peers' flags and combines the ranks' slots in rank order, so every
 * rank ends with identical bits.  Any shape / dtype / op b200phi_sharded_reduce takes.
 * Set-up (once): every rank allocates b200phi_p2p_buffer_bytes(...) with b200phi_p2p_alloc,
 * exchanges the 64-byte handles with its rendezvous, maps the peers' buffers with
 * b200phi_p2p_open and fills base[r] (base[rank] = its own pointer, which must come from
 * b200phi_p2p_alloc).
 * Per call: the same epoch on every rank, starting at 1 and growing by 1; calls of one
 * group must be issued in the same order, with the same output shape, on every rank and on one
 * stream per rank; the kernels of all ranks must be able to run at the same time (one GPU per rank).
 * Failure mode: a peer that does not publish within B200PHI_P2P_TIMEOUT_MS (environment, read
 * once; default 4000) makes the waiting rank write NOTHING to d->out for that call and record
 * the epoch in a host-visible status word; the next b200phi_sharded_reduce_p2p on the group, or
 * b200phi_p2p_status(base[rank]) after a stream synchronise, returns B200PHI_PEER_TIMEOUT.  A
 * group that timed out is out of step and must be torn down. */
typedef struct {
  int32_t nranks;        /* 1..8 */
  int32_t rank;
  void* base[8];         /* base[r]: rank r's buffer as mapped into this process */
  int64_t max_out_numel; /* output elements the buffers were sized for */
  int64_t max_tiles;     /* flags per rank: the exchange runs on min(64, max_tiles, what the size needs) CTAs */
} b200phi_p2p_group;

B200PHI_API size_t b200phi_p2p_buffer_bytes(int nranks, int64_t max_out_numel, int64_t max_tiles);
B200PHI_API int b200phi_p2p_alloc(size_t bytes, void** ptr, unsigned char handle[64]);
B200PHI_API int b200phi_p2p_open(const unsigned char handle[64], void** ptr);
B200PHI_API int b200phi_p2p_close(void* ptr);
B200PHI_API int b200phi_p2p_free(void* ptr);
/* B200PHI_OK, or B200PHI_PEER_TIMEOUT if an exchange on this buffer (own_ptr from
 * b200phi_p2p_alloc) gave up waiting for a peer.  Host-side read of a mapped word: no stream
 * synchronisation, so call it after synchronising the stream of the call it should cover. */
B200PHI_API int b200phi_p2p_status(void* own_ptr);
B200PHI_API int b200phi_sharded_reduce_p2p(const b200phi_reduce_desc* d, int64_t global_reduce_count,
                                           const b200phi_p2p_group* g, uint32_t epoch, void* stream);

/* ------------------------------------------------------------------------ */

/* Message of the last non-zero status returned on this host thread. */
B200PHI_API const char* b200phi_last_error(void);

/* "b200phi <version> sm_100a" -- lets a loader check it bound the right file. */
B200PHI_API const char* b200phi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200PHI_H_ */
