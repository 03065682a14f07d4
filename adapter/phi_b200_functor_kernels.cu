// Drop-in adapter, second translation unit: the phi kernels that are functor plug-ins on the
// broadcast engine, implemented on b200phi_map / b200phi_reduce and
// registered under the reference's kernel names and signatures.  Compiled inside a Paddle build
// in place of
//   paddle/phi/kernels/kps/compare_kernel.cu            (less_than ... not_equal)
//   paddle/phi/kernels/kps/logical_kernel.cu            (logical_and/or/xor/not)
//   paddle/phi/kernels/kps/bitwise_kernel.cu            (bitwise_and/or/xor/not)
//   paddle/phi/kernels/gpu/activation_kernel.cu         (the activations listed below)
//   paddle/phi/kernels/gpu/reduce_kernel.cu:42-232      (ReduceSum/Mean/Max/Min/AMax/AMinGradKernel; its
//                                                        `reduce` collective kernel stays where it is)
//   paddle/phi/kernels/gpu/elementwise_grad_kernel.cu   (Add/Subtract/Multiply/Divide/Maximum/MinimumGradKernel:
//                                                        those registrations only)
//   paddle/phi/kernels/gpu/{p_norm,squared_l2_norm,mean_all,reduce_as,logsumexp}_kernel.cu
// Compile-checked against the reference headers (adapter/check_compile.sh).
#include <vector>

#include "b200phi.h"
#include "paddle/phi/backends/gpu/gpu_context.h"
#include "paddle/phi/common/int_array.h"
#include "paddle/phi/core/dense_tensor.h"
#include "paddle/phi/core/kernel_registry.h"
#include "paddle/phi/core/tensor_utils.h"

namespace phi {
namespace b200 {

inline void CheckMap(int status) {
  if (status != B200PHI_OK) {
    PADDLE_THROW(common::errors::InvalidArgument("%s", b200phi_last_error()));
  }
}

inline void View(const DenseTensor& t, b200phi_tensor_view* v) {
  v->rank = t.dims().size();
  for (int i = 0; i < v->rank; ++i) v->dims[i] = t.dims()[i];
  v->data = t.data();
}

// x's dims with the reduced ones set to 1: how ReduceGrad views out and out_grad
// (gpu/reduce_kernel.cu:84-101)
inline void KeepDimView(const DenseTensor& t, const DenseTensor& x, const std::vector<int64_t>& dims,
                        bool reduce_all, b200phi_tensor_view* v) {
  const int rank = x.dims().size();
  v->rank = rank;
  for (int i = 0; i < rank; ++i) v->dims[i] = x.dims()[i];
  if (reduce_all || dims.empty() || static_cast<int>(dims.size()) == rank) {
    for (int i = 0; i < rank; ++i) v->dims[i] = 1;
  } else {
    for (int64_t a : dims) v->dims[a < 0 ? a + rank : a] = 1;
  }
  v->data = t.data();
}

// One b200phi_map launch; out's dims come from InferMeta, its dtype from the registration
// (BOOL for compare / logical) or from x.
template <typename Context>
void Map(const Context& dev_ctx, int op, const std::vector<const DenseTensor*>& ins, int axis,
         float p0, float p1, DenseTensor* out) {
  const bool to_bool = op < B200PHI_MAP_BITWISE_AND;
  dev_ctx.Alloc(out, to_bool ? DataType::BOOL : ins[0]->dtype());
  b200phi_map_desc d = {};
  d.op = op;
  d.in_dtype = static_cast<int>(ins[0]->dtype());
  d.out_dtype = B200PHI_UNDEFINED;
  d.arity = static_cast<int>(ins.size());
  d.axis = axis;
  d.p0 = p0;
  d.p1 = p1;
  for (size_t i = 0; i < ins.size(); ++i) View(*ins[i], &d.in[i]);
  d.out_rank = out->dims().size();
  for (int i = 0; i < d.out_rank; ++i) d.out_dims[i] = out->dims()[i];
  d.out = out->data();
  CheckMap(b200phi_map(&d, dev_ctx.stream()));
}

template <typename Context>
void MinMaxGrad(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& out,
                const DenseTensor& out_grad, const std::vector<int64_t>& dims, bool reduce_all,
                DenseTensor* x_grad) {
  dev_ctx.Alloc(x_grad, x.dtype());
  b200phi_map_desc d = {};
  d.op = B200PHI_MAP_EQUAL_MUL;  // dout * T(Equal(x, out)): one launch, no equal_out temporary
  d.in_dtype = static_cast<int>(x.dtype());
  d.arity = 3;
  d.axis = -1;
  View(x, &d.in[0]);
  KeepDimView(out, x, dims, reduce_all, &d.in[1]);
  KeepDimView(out_grad, x, dims, reduce_all, &d.in[2]);
  d.out_rank = x.dims().size();
  for (int i = 0; i < d.out_rank; ++i) d.out_dims[i] = x.dims()[i];
  d.out = x_grad->data();
  CheckMap(b200phi_map(&d, dev_ctx.stream()));
}

// dx (or dy) of maximum / minimum: the ternary functor, then -- when the operand was
// broadcast -- a sum over the broadcast dims (funcs::GetReduceDim; gpu/elementwise_grad.h:173+).
template <typename Context>
void SelectGrad(const Context& dev_ctx, int op, const DenseTensor& x, const DenseTensor& y,
                const DenseTensor& dout, const DenseTensor& like, DenseTensor* grad) {
  if (grad == nullptr) return;
  if (like.dims() == dout.dims()) {
    grad->Resize(dout.dims());
    Map(dev_ctx, op, {&x, &y, &dout}, -1, 0.f, 0.f, grad);
    return;
  }
  DenseTensor full;
  full.Resize(dout.dims());
  Map(dev_ctx, op, {&x, &y, &dout}, -1, 0.f, 0.f, &full);
  // dims of dout that `like` does not have, or has as 1
  const int r = dout.dims().size(), lr = like.dims().size(), off = r - lr;
  b200phi_reduce_desc d = {};
  d.op = B200PHI_SUM;
  d.in_dtype = d.out_dtype = static_cast<int>(full.dtype());
  d.rank = r;
  for (int i = 0; i < r; ++i) {
    d.dims[i] = dout.dims()[i];
    if (i < off || (like.dims()[i - off] == 1 && dout.dims()[i] != 1)) d.axes[d.naxes++] = i;
  }
  dev_ctx.Alloc(grad, like.dtype());
  d.x = full.data();
  d.out = grad->data();
  DenseTensor scratch;
  const size_t ws = b200phi_reduce_workspace(&d);
  if (ws > 0) {
    scratch.Resize({static_cast<int64_t>(ws)});
    d.workspace = dev_ctx.template Alloc<uint8_t>(&scratch);
    d.workspace_bytes = ws;
  }
  CheckMap(b200phi_reduce(&d, dev_ctx.stream()));
}

}  // namespace b200

// ---- compare: compare_kernel.h:20-33 -----------------------------------------------------------
#define B200_DEFINE_COMPARE(Name, OpId)                                                   \
  template <typename T, typename Context>                                                 \
  void Name##Kernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,   \
                    DenseTensor* out) {                                                   \
    b200::Map(dev_ctx, OpId, {&x, &y}, -1, 0.f, 0.f, out);                                \
  }
B200_DEFINE_COMPARE(LessThan, B200PHI_MAP_LESS_THAN)
B200_DEFINE_COMPARE(LessEqual, B200PHI_MAP_LESS_EQUAL)
B200_DEFINE_COMPARE(GreaterThan, B200PHI_MAP_GREATER_THAN)
B200_DEFINE_COMPARE(GreaterEqual, B200PHI_MAP_GREATER_EQUAL)
B200_DEFINE_COMPARE(Equal, B200PHI_MAP_EQUAL)
B200_DEFINE_COMPARE(NotEqual, B200PHI_MAP_NOT_EQUAL)
// ---- logical: logical_kernel.h:21-36 -------------------------------------------------------------
B200_DEFINE_COMPARE(LogicalAnd, B200PHI_MAP_LOGICAL_AND)
B200_DEFINE_COMPARE(LogicalOr, B200PHI_MAP_LOGICAL_OR)
B200_DEFINE_COMPARE(LogicalXor, B200PHI_MAP_LOGICAL_XOR)
// ---- bitwise: bitwise_kernel.h:21-42 -------------------------------------------------------------
B200_DEFINE_COMPARE(BitwiseAnd, B200PHI_MAP_BITWISE_AND)
B200_DEFINE_COMPARE(BitwiseOr, B200PHI_MAP_BITWISE_OR)
B200_DEFINE_COMPARE(BitwiseXor, B200PHI_MAP_BITWISE_XOR)

// bitwise_kernel.h:44-56
template <typename T, typename Context>
void BitwiseLeftShiftKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,
                            bool is_arithmetic, DenseTensor* out) {
  b200::Map(dev_ctx, is_arithmetic ? B200PHI_MAP_LEFT_SHIFT_ARITHMETIC : B200PHI_MAP_LEFT_SHIFT_LOGIC,
            {&x, &y}, -1, 0.f, 0.f, out);
}
template <typename T, typename Context>
void BitwiseRightShiftKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,
                             bool is_arithmetic, DenseTensor* out) {
  b200::Map(dev_ctx, is_arithmetic ? B200PHI_MAP_RIGHT_SHIFT_ARITHMETIC : B200PHI_MAP_RIGHT_SHIFT_LOGIC,
            {&x, &y}, -1, 0.f, 0.f, out);
}

template <typename T, typename Context>
void LogicalNotKernel(const Context& dev_ctx, const DenseTensor& x, DenseTensor* out) {
  b200::Map(dev_ctx, B200PHI_MAP_LOGICAL_NOT, {&x}, -1, 0.f, 0.f, out);
}
template <typename T, typename Context>
void BitwiseNotKernel(const Context& dev_ctx, const DenseTensor& x, DenseTensor* out) {
  b200::Map(dev_ctx, B200PHI_MAP_BITWISE_NOT, {&x}, -1, 0.f, 0.f, out);
}

// ---- activations: activation_kernel.h:23-110 -----------------------------------------------------
#define B200_DEFINE_ACT(Name, OpId)                                                       \
  template <typename T, typename Context>                                                 \
  void Name##Kernel(const Context& dev_ctx, const DenseTensor& x, DenseTensor* out) {     \
    b200::Map(dev_ctx, OpId, {&x}, -1, 0.f, 0.f, out);                                    \
  }
#define B200_DEFINE_ACT_1(Name, OpId)                                                     \
  template <typename T, typename Context>                                                 \
  void Name##Kernel(const Context& dev_ctx, const DenseTensor& x, float a, DenseTensor* out) { \
    b200::Map(dev_ctx, OpId, {&x}, -1, a, 0.f, out);                                      \
  }
#define B200_DEFINE_ACT_2(Name, OpId)                                                     \
  template <typename T, typename Context>                                                 \
  void Name##Kernel(const Context& dev_ctx, const DenseTensor& x, float a, float b,       \
                    DenseTensor* out) {                                                   \
    b200::Map(dev_ctx, OpId, {&x}, -1, a, b, out);                                        \
  }
B200_DEFINE_ACT(Exp, B200PHI_MAP_EXP)
B200_DEFINE_ACT(Log, B200PHI_MAP_LOG)
B200_DEFINE_ACT(Sqrt, B200PHI_MAP_SQRT)
B200_DEFINE_ACT(Rsqrt, B200PHI_MAP_RSQRT)
B200_DEFINE_ACT(Square, B200PHI_MAP_SQUARE)
B200_DEFINE_ACT(Reciprocal, B200PHI_MAP_RECIPROCAL)
B200_DEFINE_ACT(Sin, B200PHI_MAP_SIN)
B200_DEFINE_ACT(Cos, B200PHI_MAP_COS)
B200_DEFINE_ACT(Tanh, B200PHI_MAP_TANH)
B200_DEFINE_ACT(Sigmoid, B200PHI_MAP_SIGMOID)
B200_DEFINE_ACT(Silu, B200PHI_MAP_SILU)
B200_DEFINE_ACT(Floor, B200PHI_MAP_FLOOR)
B200_DEFINE_ACT(Ceil, B200PHI_MAP_CEIL)
B200_DEFINE_ACT(Expm1, B200PHI_MAP_EXPM1)
B200_DEFINE_ACT(Log1p, B200PHI_MAP_LOG1P)
B200_DEFINE_ACT(Log2, B200PHI_MAP_LOG2)
B200_DEFINE_ACT(Log10, B200PHI_MAP_LOG10)
B200_DEFINE_ACT(Softsign, B200PHI_MAP_SOFTSIGN)
B200_DEFINE_ACT(LogSigmoid, B200PHI_MAP_LOG_SIGMOID)
B200_DEFINE_ACT(Relu, B200PHI_MAP_RELU)
B200_DEFINE_ACT_1(LeakyRelu, B200PHI_MAP_LEAKY_RELU)
B200_DEFINE_ACT_1(Elu, B200PHI_MAP_ELU)
B200_DEFINE_ACT_1(Mish, B200PHI_MAP_MISH)
B200_DEFINE_ACT_2(Softplus, B200PHI_MAP_SOFTPLUS)
B200_DEFINE_ACT_2(HardSigmoid, B200PHI_MAP_HARD_SIGMOID)

template <typename T, typename Context>
void Relu6Kernel(const Context& dev_ctx, const DenseTensor& x, DenseTensor* out) {
  b200::Map(dev_ctx, B200PHI_MAP_RELU6, {&x}, -1, 6.0f, 0.f, out);  // activation_kernel.cu: threshold 6
}
template <typename T, typename Context>
void SwishKernel(const Context& dev_ctx, const DenseTensor& x, DenseTensor* out) {
  b200::Map(dev_ctx, B200PHI_MAP_SWISH, {&x}, -1, 1.0f, 0.f, out);  // beta 1
}

// ---- backward: reduce_max_grad_kernel.h:22-30, elementwise_grad_kernel.h:38-52 -----------------
template <typename T, typename Context>
void ReduceMaxGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& out,
                         const DenseTensor& out_grad, const IntArray& dims, bool keep_dim,
                         bool reduce_all, DenseTensor* x_grad) {
  b200::MinMaxGrad(dev_ctx, x, out, out_grad, dims.GetData(), reduce_all, x_grad);
}
template <typename T, typename Context>
void ReduceMinGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& out,
                         const DenseTensor& out_grad, const IntArray& dims, bool keep_dim,
                         bool reduce_all, DenseTensor* x_grad) {
  b200::MinMaxGrad(dev_ctx, x, out, out_grad, dims.GetData(), reduce_all, x_grad);
}
template <typename T, typename Context>
void MaximumGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,
                       const DenseTensor& dout, DenseTensor* dx, DenseTensor* dy) {
  b200::SelectGrad(dev_ctx, B200PHI_MAP_MAX_GRAD_X, x, y, dout, x, dx);
  b200::SelectGrad(dev_ctx, B200PHI_MAP_MAX_GRAD_Y, x, y, dout, y, dy);
}
template <typename T, typename Context>
void MinimumGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,
                       const DenseTensor& dout, DenseTensor* dx, DenseTensor* dy) {
  b200::SelectGrad(dev_ctx, B200PHI_MAP_MIN_GRAD_X, x, y, dout, x, dx);
  b200::SelectGrad(dev_ctx, B200PHI_MAP_MIN_GRAD_Y, x, y, dout, y, dy);
}

// ---- reduce-side plug-ins: p_norm_kernel.h, squared_l2_norm_kernel.h, mean_all_kernel.h,
// reduce_as_kernel.h, logsumexp_kernel.h -------------------------------------------------------
namespace b200 {

template <typename Context>
void ReduceOp(const Context& dev_ctx, int op, const DenseTensor& x, const std::vector<int64_t>& axes,
              bool reduce_all, DenseTensor* out) {
  dev_ctx.Alloc(out, x.dtype());
  b200phi_reduce_desc d = {};
  d.op = op;
  d.in_dtype = d.out_dtype = static_cast<int>(x.dtype());
  d.rank = x.dims().size();
  for (int i = 0; i < d.rank; ++i) d.dims[i] = x.dims()[i];
  d.naxes = reduce_all ? 0 : static_cast<int>(axes.size());
  for (int i = 0; i < d.naxes; ++i) d.axes[i] = static_cast<int>(axes[i]);
  d.reduce_all = reduce_all ? 1 : 0;
  d.x = x.data();
  d.out = out->data();
  DenseTensor scratch;
  const size_t ws = b200phi_reduce_workspace(&d);
  if (ws > 0) {
    scratch.Resize({static_cast<int64_t>(ws)});
    d.workspace = dev_ctx.template Alloc<uint8_t>(&scratch);
    d.workspace_bytes = ws;
  }
  CheckMap(b200phi_reduce(&d, dev_ctx.stream()));
}

}  // namespace b200

template <typename T, typename Context>
void PNormKernel(const Context& dev_ctx, const DenseTensor& x, float porder, int axis, float epsilon,
                 bool keepdim, bool asvector, DenseTensor* out) {
  const std::vector<int64_t> axes = {axis};
  if (porder == 0) {
    b200::ReduceOp(dev_ctx, B200PHI_COUNT_NONZERO, x, axes, asvector, out);
  } else if (porder == INFINITY) {
    b200::ReduceOp(dev_ctx, B200PHI_MAX_ABS, x, axes, asvector, out);
  } else if (porder == -INFINITY) {
    b200::ReduceOp(dev_ctx, B200PHI_MIN_ABS, x, axes, asvector, out);
  } else if (porder == 1.0f) {
    b200::ReduceOp(dev_ctx, B200PHI_SUM_ABS, x, axes, asvector, out);
  } else if (porder == 2.0f) {
    b200::ReduceOp(dev_ctx, B200PHI_L2_NORM, x, axes, asvector, out);
  } else {
    // |x|^p -> sum -> ^(1/p): the three functors of gpu/p_norm_kernel.cu:84-95
    DenseTensor powed;
    powed.Resize(x.dims());
    b200::Map(dev_ctx, B200PHI_MAP_POW_ABS, {&x}, -1, porder, 0.f, &powed);
    b200::ReduceOp(dev_ctx, B200PHI_SUM, powed, axes, asvector, out);
    DenseTensor in_place = *out;
    b200::Map(dev_ctx, B200PHI_MAP_POW_ABS, {&in_place}, -1, 1.0f / porder, 0.f, out);
  }
}

template <typename T, typename Context>
void SquaredL2NormKernel(const Context& dev_ctx, const DenseTensor& x, DenseTensor* out) {
  b200::ReduceOp(dev_ctx, B200PHI_SUM_SQUARE, x, {}, true, out);
}

template <typename T, typename Context>
void MeanAllKernel(const Context& dev_ctx, const DenseTensor& x, DenseTensor* out) {
  b200::ReduceOp(dev_ctx, B200PHI_MEAN, x, {}, true, out);
}

template <typename T, typename Context>
void ReduceAsKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& target,
                    DenseTensor* out) {
  // funcs::GetReduceDims(x, target): leading extra dims, and dims where target has 1
  const int xr = x.dims().size(), tr = target.dims().size(), off = xr - tr;
  std::vector<int64_t> axes;
  for (int i = 0; i < xr; ++i)
    if (i < off || (target.dims()[i - off] == 1 && x.dims()[i] != 1)) axes.push_back(i);
  if (axes.empty()) axes.push_back(0), out->Resize(x.dims());  // degenerate: reduce_num == 1 copies
  b200::ReduceOp(dev_ctx, B200PHI_SUM, x, axes, false, out);
}

template <typename T, typename Context>
void LogsumexpKernel(const Context& dev_ctx, const DenseTensor& x, const std::vector<int>& axis,
                     bool keepdim, bool reduce_all, DenseTensor* out) {
  const int rank = x.dims().size();
  std::vector<bool> red(rank > 0 ? rank : 1, false);
  if (axis.empty() || reduce_all || static_cast<int>(axis.size()) == rank) {
    for (int i = 0; i < rank; ++i) red[i] = true;
  } else {
    for (int a : axis) red[a < 0 ? a + rank : a] = true;
  }
  int first = 0;
  while (first < rank && !red[first]) ++first;
  bool trailing = true;
  for (int i = first; i < rank; ++i) trailing = trailing && red[i];
  dev_ctx.Alloc(out, x.dtype());
  if (trailing) {  // one launch, one read of x
    b200phi_logsumexp_desc d = {};
    d.dtype = static_cast<int>(x.dtype());
    d.rows = 1;
    d.cols = 1;
    for (int i = 0; i < rank; ++i) (i < first ? d.rows : d.cols) *= x.dims()[i];
    d.x = x.data();
    d.out = out->data();
    b200::CheckMap(b200phi_logsumexp(&d, dev_ctx.stream()));
    return;
  }
  // other axis sets: max -> exp(x - max) -> sum -> log + max (the reference's fallback,
  // gpu/logsumexp_kernel.cu:49-78) on the two engines
  std::vector<int64_t> axes;
  std::vector<int64_t> kept = common::vectorize(x.dims());
  for (int i = 0; i < rank; ++i)
    if (red[i]) axes.push_back(i), kept[i] = 1;
  DenseTensor mx, ex, sm;
  mx.Resize(common::make_ddim(kept));
  b200::ReduceOp(dev_ctx, B200PHI_MAX, x, axes, false, &mx);
  ex.Resize(x.dims());
  b200::Map(dev_ctx, B200PHI_MAP_SUB_EXP, {&x, &mx}, -1, 0.f, 0.f, &ex);
  sm.Resize(common::make_ddim(kept));
  b200::ReduceOp(dev_ctx, B200PHI_SUM, ex, axes, false, &sm);
  DenseTensor lg;
  lg.Resize(common::make_ddim(kept));
  b200::Map(dev_ctx, B200PHI_MAP_LOG, {&sm}, -1, 0.f, 0.f, &lg);
  const auto out_dims = out->dims();
  out->Resize(common::make_ddim(kept));
  b200phi_elementwise_desc e = {};
  e.binary = B200PHI_ADD;
  e.dtype = static_cast<int>(x.dtype());
  e.arity = 2;
  e.axis = -1;
  b200::View(lg, &e.in[0]);
  b200::View(mx, &e.in[1]);
  e.out_rank = static_cast<int>(kept.size());
  for (size_t i = 0; i < kept.size(); ++i) e.out_dims[i] = kept[i];
  e.out = out->data();
  b200::CheckMap(b200phi_elementwise(&e, dev_ctx.stream()));
  out->Resize(out_dims);
}

// ---- backward of the reductions and of add .. divide ---------------------------
namespace b200 {

// One launch of the broadcast engine with a single operand: out = act(broadcast(in)), the unary
// BroadcastKernel call of ReduceGrad (gpu/reduce_grad.h:30-85).
template <typename Context>
void BroadcastUnary(const Context& dev_ctx, const b200phi_tensor_view& in, DataType dtype, int act,
                    float scale, DenseTensor* out) {
  dev_ctx.Alloc(out, dtype);
  b200phi_elementwise_desc e = {};
  e.binary = B200PHI_ADD;
  e.act = act;
  e.scale = scale;
  e.dtype = static_cast<int>(dtype);
  e.arity = 1;
  e.axis = -1;
  e.in[0] = in;
  e.out_rank = out->dims().size();
  for (int i = 0; i < e.out_rank; ++i) e.out_dims[i] = out->dims()[i];
  e.out = out->data();
  CheckMap(b200phi_elementwise(&e, dev_ctx.stream()));
}

template <typename Context>
void Binary(const Context& dev_ctx, int binary, const DenseTensor& x, const DenseTensor& y, int axis,
            DenseTensor* out) {
  dev_ctx.Alloc(out, x.dtype());
  b200phi_elementwise_desc e = {};
  e.binary = binary;
  e.dtype = static_cast<int>(x.dtype());
  e.arity = 2;
  e.axis = axis;
  View(x, &e.in[0]);
  View(y, &e.in[1]);
  e.out_rank = out->dims().size();
  for (int i = 0; i < e.out_rank; ++i) e.out_dims[i] = out->dims()[i];
  e.out = out->data();
  CheckMap(b200phi_elementwise(&e, dev_ctx.stream()));
}

// grad of a (possibly broadcast) operand shaped like `like`: `src` itself when the shapes agree,
// else its sum over the dims funcs::GetReduceDim(like.dims(), src.dims(), axis) names
// (elementwise_base.h; gpu/elementwise_grad.h:29-105 reduces with phi::SumKernel the same way).
template <typename Context>
void SumTo(const Context& dev_ctx, const DenseTensor& src, const DenseTensor& like, int axis,
           DenseTensor* dst) {
  if (dst == nullptr) return;
  if (like.dims() == src.dims()) {
    phi::Copy(dev_ctx, src, dev_ctx.GetPlace(), false, dst);
    return;
  }
  const int r = src.dims().size(), lr = like.dims().size();
  const int off = axis < 0 ? r - lr : axis;
  b200phi_reduce_desc d = {};
  d.op = B200PHI_SUM;
  d.in_dtype = d.out_dtype = static_cast<int>(src.dtype());
  d.rank = r;
  for (int i = 0; i < r; ++i) {
    d.dims[i] = src.dims()[i];
    const bool outside = i < off || i >= off + lr;
    if (outside || (like.dims()[i - off] == 1 && src.dims()[i] != 1)) d.axes[d.naxes++] = i;
  }
  dst->Resize(like.dims());
  dev_ctx.Alloc(dst, src.dtype());
  d.x = src.data();
  d.out = dst->data();
  DenseTensor scratch;
  const size_t ws = b200phi_reduce_workspace(&d);
  if (ws > 0) {
    scratch.Resize({static_cast<int64_t>(ws)});
    d.workspace = dev_ctx.template Alloc<uint8_t>(&scratch);
    d.workspace_bytes = ws;
  }
  CheckMap(b200phi_reduce(&d, dev_ctx.stream()));
}

inline int AxisFor(const DenseTensor& operand, const DenseTensor& other, int axis) {
  return operand.dims().size() < other.dims().size() ? axis : -1;
}

}  // namespace b200

// gpu/reduce_kernel.cu:42-75: out_grad viewed with the reduced dims as 1, broadcast to x's shape
template <typename T, typename Context>
void ReduceSumGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& out_grad,
                         const IntArray& dims, bool keep_dim, bool reduce_all, DenseTensor* x_grad) {
  PADDLE_ENFORCE_EQ(out_grad.dtype(), x.dtype(),
                    common::errors::Unimplemented("sum_grad: out_grad and x must share a dtype here"));
  b200phi_tensor_view g;
  b200::KeepDimView(out_grad, x, dims.GetData(), reduce_all, &g);
  x_grad->Resize(x.dims());
  b200::BroadcastUnary(dev_ctx, g, x.dtype(), B200PHI_ACT_NONE, 0.f, x_grad);
}

// gpu/reduce_kernel.cu:124-158: the same broadcast with kps::DivideFunctor(reduce_num) fused in
template <typename T, typename Context>
void ReduceMeanGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& out_grad,
                          const IntArray& dims, bool keep_dim, bool reduce_all, DenseTensor* x_grad) {
  b200phi_tensor_view g;
  b200::KeepDimView(out_grad, x, dims.GetData(), reduce_all, &g);
  int64_t reduce_num = 1;
  for (int i = 0; i < g.rank; ++i)
    if (g.dims[i] == 1) reduce_num *= x.dims()[i];
  x_grad->Resize(x.dims());
  b200::BroadcastUnary(dev_ctx, g, x.dtype(), B200PHI_ACT_SCALE, 1.0f / static_cast<float>(reduce_num),
                       x_grad);
}

// gpu/reduce_amin_amax_common.h:28-107: the gradient is shared evenly between the positions that
// tie with the extreme: mask = T(x == out); x_grad = mask * (out_grad / sum(mask))
template <typename Context>
void AMaxAMinGrad(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& out,
                  const DenseTensor& out_grad, const std::vector<int64_t>& dims, bool reduce_all,
                  DenseTensor* x_grad) {
  b200phi_tensor_view o, g;
  b200::KeepDimView(out, x, dims, reduce_all, &o);
  b200::KeepDimView(out_grad, x, dims, reduce_all, &g);
  // mask over x's shape
  DenseTensor mask;
  mask.Resize(x.dims());
  dev_ctx.Alloc(&mask, x.dtype());
  b200phi_map_desc m = {};
  m.op = B200PHI_MAP_EQUAL_T;
  m.in_dtype = static_cast<int>(x.dtype());
  m.arity = 2;
  m.axis = -1;
  b200::View(x, &m.in[0]);
  m.in[1] = o;
  m.out_rank = x.dims().size();
  for (int i = 0; i < m.out_rank; ++i) m.out_dims[i] = x.dims()[i];
  m.out = mask.data();
  b200::CheckMap(b200phi_map(&m, dev_ctx.stream()));
  // count of ties per output, kept dims as 1
  DenseTensor count;
  std::vector<int64_t> kept(g.dims, g.dims + g.rank);
  count.Resize(common::make_ddim(kept));
  dev_ctx.Alloc(&count, x.dtype());
  b200phi_reduce_desc r = {};
  r.op = B200PHI_SUM;
  r.in_dtype = r.out_dtype = static_cast<int>(x.dtype());
  r.rank = x.dims().size();
  for (int i = 0; i < r.rank; ++i) {
    r.dims[i] = x.dims()[i];
    if (g.dims[i] == 1 && x.dims()[i] != 1) r.axes[r.naxes++] = i;
  }
  r.reduce_all = 0;
  r.x = mask.data();
  r.out = count.data();
  DenseTensor scratch;
  if (r.naxes > 0) {
    const size_t ws = b200phi_reduce_workspace(&r);
    if (ws > 0) {
      scratch.Resize({static_cast<int64_t>(ws)});
      r.workspace = dev_ctx.template Alloc<uint8_t>(&scratch);
      r.workspace_bytes = ws;
    }
    b200::CheckMap(b200phi_reduce(&r, dev_ctx.stream()));
  } else {
    phi::Copy(dev_ctx, mask, dev_ctx.GetPlace(), false, &count);
  }
  // share = out_grad / count on the small tensor, then x_grad = share * T(x == out)
  DenseTensor share, gview;
  gview.ShareDataWith(out_grad);
  gview.Resize(common::make_ddim(kept));
  share.Resize(common::make_ddim(kept));
  b200::Binary(dev_ctx, B200PHI_DIVIDE, gview, count, -1, &share);
  x_grad->Resize(x.dims());
  dev_ctx.Alloc(x_grad, x.dtype());
  b200phi_map_desc e = {};
  e.op = B200PHI_MAP_EQUAL_MUL;
  e.in_dtype = static_cast<int>(x.dtype());
  e.arity = 3;
  e.axis = -1;
  b200::View(x, &e.in[0]);
  e.in[1] = o;
  b200::View(share, &e.in[2]);
  e.out_rank = x.dims().size();
  for (int i = 0; i < e.out_rank; ++i) e.out_dims[i] = x.dims()[i];
  e.out = x_grad->data();
  b200::CheckMap(b200phi_map(&e, dev_ctx.stream()));
}
template <typename T, typename Context>
void ReduceAMaxGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& out,
                          const DenseTensor& out_grad, const std::vector<int64_t>& dims, bool keep_dim,
                          bool reduce_all, DenseTensor* x_grad) {
  AMaxAMinGrad(dev_ctx, x, out, out_grad, dims, reduce_all, x_grad);
}
template <typename T, typename Context>
void ReduceAMinGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& out,
                          const DenseTensor& out_grad, const std::vector<int64_t>& dims, bool keep_dim,
                          bool reduce_all, DenseTensor* x_grad) {
  AMaxAMinGrad(dev_ctx, x, out, out_grad, dims, reduce_all, x_grad);
}

// gpu/elementwise_grad.h:29-105,173-304 and elementwise_{add,subtract,multiply,divide}_grad_kernel.h
template <typename T, typename Context>
void AddGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,
                   const DenseTensor& dout, int axis, DenseTensor* dx, DenseTensor* dy) {
  b200::SumTo(dev_ctx, dout, x, b200::AxisFor(x, y, axis), dx);
  b200::SumTo(dev_ctx, dout, y, b200::AxisFor(y, x, axis), dy);
}
template <typename T, typename Context>
void SubtractGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,
                        const DenseTensor& dout, int axis, DenseTensor* dx, DenseTensor* dy) {
  b200::SumTo(dev_ctx, dout, x, b200::AxisFor(x, y, axis), dx);
  if (dy != nullptr) {  // dy = -sum(dout): negate on the small side
    DenseTensor pos;
    b200::SumTo(dev_ctx, dout, y, b200::AxisFor(y, x, axis), &pos);
    b200phi_tensor_view v;
    b200::View(pos, &v);
    dy->Resize(y.dims());
    b200::BroadcastUnary(dev_ctx, v, pos.dtype(), B200PHI_ACT_SCALE, -1.0f, dy);
  }
}
template <typename T, typename Context>
void MultiplyGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,
                        const DenseTensor& dout, int axis, DenseTensor* dx, DenseTensor* dy) {
  if (dx != nullptr) {
    DenseTensor full;
    full.Resize(dout.dims());
    b200::Binary(dev_ctx, B200PHI_MULTIPLY, dout, y, b200::AxisFor(y, dout, axis), &full);
    b200::SumTo(dev_ctx, full, x, b200::AxisFor(x, y, axis), dx);
  }
  if (dy != nullptr) {
    DenseTensor full;
    full.Resize(dout.dims());
    b200::Binary(dev_ctx, B200PHI_MULTIPLY, dout, x, b200::AxisFor(x, dout, axis), &full);
    b200::SumTo(dev_ctx, full, y, b200::AxisFor(y, x, axis), dy);
  }
}
template <typename T, typename Context>
void DivideGradKernel(const Context& dev_ctx, const DenseTensor& x, const DenseTensor& y,
                      const DenseTensor& out, const DenseTensor& dout, int axis, DenseTensor* dx,
                      DenseTensor* dy) {
  if (dx != nullptr) {  // dout / y
    DenseTensor full;
    full.Resize(dout.dims());
    b200::Binary(dev_ctx, B200PHI_DIVIDE, dout, y, b200::AxisFor(y, dout, axis), &full);
    b200::SumTo(dev_ctx, full, x, b200::AxisFor(x, y, axis), dx);
  }
  if (dy != nullptr) {  // -dout * ((out / y) / y): DivGradYFunctor, one ternary launch
    DenseTensor full;
    full.Resize(dout.dims());
    b200::Map(dev_ctx, B200PHI_MAP_DIV_GRAD_Y, {&dout, &out, &y}, b200::AxisFor(y, dout, axis), 0.f, 0.f,
              &full);
    b200::SumTo(dev_ctx, full, y, b200::AxisFor(y, x, axis), dy);
  }
}

}  // namespace phi

using float16 = phi::dtype::float16;
using bfloat16 = phi::dtype::bfloat16;

// dtype lists: kps/compare_kernel.cu:146-165, kps/logical_kernel.cu, kps/bitwise_kernel.cu,
// gpu/activation_kernel.cu:205-330, minus complex (not implemented by this library)
#define B200_REGISTER_COMPARE(name, func)                                                        \
  PD_REGISTER_KERNEL(name, GPU, ALL_LAYOUT, phi::func##Kernel, bool, int, uint8_t, int8_t,       \
                     int16_t, int64_t, float, double, float16, bfloat16) {                       \
    kernel->OutputAt(0).SetDataType(phi::DataType::BOOL);                                        \
  }
B200_REGISTER_COMPARE(less_than, LessThan)
B200_REGISTER_COMPARE(less_equal, LessEqual)
B200_REGISTER_COMPARE(greater_than, GreaterThan)
B200_REGISTER_COMPARE(greater_equal, GreaterEqual)
B200_REGISTER_COMPARE(equal, Equal)
B200_REGISTER_COMPARE(not_equal, NotEqual)
B200_REGISTER_COMPARE(logical_and, LogicalAnd)
B200_REGISTER_COMPARE(logical_or, LogicalOr)
B200_REGISTER_COMPARE(logical_xor, LogicalXor)
B200_REGISTER_COMPARE(logical_not, LogicalNot)

#define B200_REGISTER_BITWISE(name, func)                                                        \
  PD_REGISTER_KERNEL(name, GPU, ALL_LAYOUT, phi::func##Kernel, bool, uint8_t, int8_t, int16_t,   \
                     int, int64_t) {}
B200_REGISTER_BITWISE(bitwise_and, BitwiseAnd)
B200_REGISTER_BITWISE(bitwise_or, BitwiseOr)
B200_REGISTER_BITWISE(bitwise_xor, BitwiseXor)
B200_REGISTER_BITWISE(bitwise_not, BitwiseNot)
PD_REGISTER_KERNEL(bitwise_left_shift, GPU, ALL_LAYOUT, phi::BitwiseLeftShiftKernel, uint8_t, int8_t,
                   int16_t, int, int64_t) {}
PD_REGISTER_KERNEL(bitwise_right_shift, GPU, ALL_LAYOUT, phi::BitwiseRightShiftKernel, uint8_t, int8_t,
                   int16_t, int, int64_t) {}

#define B200_REGISTER_ACT(name, func) \
  PD_REGISTER_KERNEL(name, GPU, ALL_LAYOUT, phi::func##Kernel, float, double, float16, bfloat16) {}
B200_REGISTER_ACT(exp, Exp)
B200_REGISTER_ACT(log, Log)
B200_REGISTER_ACT(sqrt, Sqrt)
B200_REGISTER_ACT(rsqrt, Rsqrt)
B200_REGISTER_ACT(square, Square)
B200_REGISTER_ACT(reciprocal, Reciprocal)
B200_REGISTER_ACT(sin, Sin)
B200_REGISTER_ACT(cos, Cos)
B200_REGISTER_ACT(tanh, Tanh)
B200_REGISTER_ACT(sigmoid, Sigmoid)
B200_REGISTER_ACT(silu, Silu)
B200_REGISTER_ACT(floor, Floor)
B200_REGISTER_ACT(ceil, Ceil)
B200_REGISTER_ACT(expm1, Expm1)
B200_REGISTER_ACT(log1p, Log1p)
B200_REGISTER_ACT(log2, Log2)
B200_REGISTER_ACT(log10, Log10)
B200_REGISTER_ACT(softsign, Softsign)
B200_REGISTER_ACT(logsigmoid, LogSigmoid)
B200_REGISTER_ACT(relu, Relu)
B200_REGISTER_ACT(leaky_relu, LeakyRelu)
B200_REGISTER_ACT(elu, Elu)
B200_REGISTER_ACT(mish, Mish)
B200_REGISTER_ACT(softplus, Softplus)
B200_REGISTER_ACT(hardsigmoid, HardSigmoid)
B200_REGISTER_ACT(relu6, Relu6)
B200_REGISTER_ACT(swish, Swish)

PD_REGISTER_KERNEL(max_grad, GPU, ALL_LAYOUT, phi::ReduceMaxGradKernel, float, double, int, int64_t,
                   float16, bfloat16) {}
PD_REGISTER_KERNEL(min_grad, GPU, ALL_LAYOUT, phi::ReduceMinGradKernel, float, double, int, int64_t,
                   float16, bfloat16) {}
// gpu/reduce_kernel.cu:317-388 and gpu/elementwise_grad_kernel.cu:320-445, minus complex
// (and, for divide_grad, minus the 8/16-bit integer and bool instantiations)
PD_REGISTER_KERNEL(sum_grad, GPU, ALL_LAYOUT, phi::ReduceSumGradKernel, bool, float, double, float16,
                   bfloat16, int8_t, uint8_t, int16_t, int, int64_t) {
  kernel->OutputAt(0).SetDataType(phi::DataType::UNDEFINED);
}
PD_REGISTER_KERNEL(mean_grad, GPU, ALL_LAYOUT, phi::ReduceMeanGradKernel, bool, float, double, float16,
                   bfloat16, int, int64_t) {}
PD_REGISTER_KERNEL(amax_grad, GPU, ALL_LAYOUT, phi::ReduceAMaxGradKernel, float, double, int, int64_t) {}
PD_REGISTER_KERNEL(amin_grad, GPU, ALL_LAYOUT, phi::ReduceAMinGradKernel, float, double, int, int64_t) {}
PD_REGISTER_KERNEL(add_grad, GPU, ALL_LAYOUT, phi::AddGradKernel, float, double, int, int64_t, float16,
                   bfloat16) {}
PD_REGISTER_KERNEL(subtract_grad, GPU, ALL_LAYOUT, phi::SubtractGradKernel, float, double, int, int64_t,
                   float16, bfloat16) {}
PD_REGISTER_KERNEL(multiply_grad, GPU, ALL_LAYOUT, phi::MultiplyGradKernel, float, float16, double, int,
                   int64_t, bool, bfloat16) {}
PD_REGISTER_KERNEL(divide_grad, GPU, ALL_LAYOUT, phi::DivideGradKernel, float, float16, bfloat16, double,
                   int, int64_t) {}
PD_REGISTER_KERNEL(maximum_grad, GPU, ALL_LAYOUT, phi::MaximumGradKernel, float, double, int,
                   int64_t, float16, bfloat16) {}
PD_REGISTER_KERNEL(minimum_grad, GPU, ALL_LAYOUT, phi::MinimumGradKernel, float, double, int,
                   int64_t, float16, bfloat16) {}

PD_REGISTER_KERNEL(p_norm, GPU, ALL_LAYOUT, phi::PNormKernel, float, double, float16, bfloat16) {}
PD_REGISTER_KERNEL(squared_l2_norm, GPU, ALL_LAYOUT, phi::SquaredL2NormKernel, float, double, float16,
                   bfloat16) {}
PD_REGISTER_KERNEL(mean_all, GPU, ALL_LAYOUT, phi::MeanAllKernel, float, double, float16) {}
PD_REGISTER_KERNEL(reduce_as, GPU, ALL_LAYOUT, phi::ReduceAsKernel, bool, float, double, float16,
                   bfloat16, int16_t, int, int64_t, uint8_t, int8_t) {}
PD_REGISTER_KERNEL(logsumexp, GPU, ALL_LAYOUT, phi::LogsumexpKernel, float, double, float16,
                   bfloat16) {}
