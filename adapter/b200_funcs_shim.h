// The three C++ template entry points other phi kernels call directly,
//   phi::funcs::ReduceKernel<Tx, Ty, ReduceOp, TransformOp, IsMean>    funcs/reduce_function.h:1047-1198
//   phi::funcs::BroadcastKernel<OutT, Functor, NumOuts>                funcs/broadcast_function.h:758-824
//   phi::funcs::ElementwiseKernel<OutT, Functor, NumOuts>              funcs/elementwise_base.h:807-845
// re-implemented on libb200phi.so's C ABI with the reference's exact signatures, so that the ~80
// client kernels (gpu/elementwise_grad.h, gpu/p_norm_kernel.cu, kps/compare_kernel.cu, ...) compile
// unmodified and land on the B200 kernels.
//
// How a C++ functor TYPE becomes a C-ABI op id:
//   * binary / ternary functors of funcs/{elementwise,compare,logical,bitwise}_functor.h: a trait
//     specialisation per functor (FunctorOp<>); a functor with no specialisation is a compile-time
//     error (static_assert) naming this header -- add the trait, or keep that client on the
//     reference's own kernels;
//   * ReduceOp (template template parameter): specialisations for the kps:: functors; any other
//     reducer (client TUs define their own, e.g. BitwiseAdd in kps/compare_kernel.cu:36-45) is
//     classified by its identity element initial(): 0 sum, 1 product (true on bool: and),
//     lowest() max, max() min;
//   * unary functors (TransformOp, and the one-operand Broadcast / Elementwise calls): client TUs
//     define these locally too (NonzeroFunctor / AbsFunctor / UnsignedPowFunctor in
//     gpu/p_norm_kernel.cu:25-52), so they are classified by what they DO on a few probe values on
//     the host (they are HOSTDEVICE): identity, |x|, x*x, x != 0, -x, k*x, |x|^p.  Anything else
//     raises Unimplemented at run time.
//
// This header is included through adapter/funcs_shim/paddle/phi/kernels/funcs/*.h, which shadow
// the three reference headers: each includes the reference's own header with the entry point's
// name macro-renamed (so every helper in it stays available) and then this file.
#pragma once

// The shadow headers rename the reference's entry points while they parse its headers, and those
// headers include each other: this file may be reached while such a rename is active.  Its own
// definitions must keep the real names.
#pragma push_macro("ReduceKernel")
#pragma push_macro("BroadcastKernel")
#pragma push_macro("ElementwiseKernel")
#undef ReduceKernel
#undef BroadcastKernel
#undef ElementwiseKernel

#include <cmath>
#include <limits>
#include <type_traits>
#include <vector>

#include "b200phi.h"
#include "paddle/phi/backends/gpu/gpu_context.h"
#include "paddle/phi/common/amp_type_traits.h"
#include "paddle/phi/core/dense_tensor.h"
#include "paddle/phi/core/enforce.h"
#include "paddle/phi/kernels/funcs/bitwise_functors.h"
#include "paddle/phi/kernels/funcs/compare_functors.h"
#include "paddle/phi/kernels/funcs/elementwise_functor.h"
#include "paddle/phi/kernels/funcs/function_traits.h"
#include "paddle/phi/kernels/funcs/logical_functor.h"
#include "paddle/phi/kernels/primitive/functor_primitives.h"

namespace phi {
namespace b200shim {

inline void Check(int status) {
  if (status != B200PHI_OK) {
    PADDLE_THROW(common::errors::InvalidArgument("%s", b200phi_last_error()));
  }
}

inline void View(const DenseTensor& t, b200phi_tensor_view* v) {
  v->rank = t.dims().size();
  for (int i = 0; i < v->rank; ++i) v->dims[i] = t.dims()[i];
  v->data = t.data();
}

// ---- unary functors: classified by behaviour ------------------------------------------------------
enum UnaryClass { kIdentity, kAbs, kSquare, kNonzero, kNegate, kScale, kPowAbs, kUnknownUnary };
struct UnaryInfo {
  int cls;
  double param;  // k of k*x, p of |x|^p
};

template <typename Tx, typename F>
UnaryInfo ClassifyUnary(const F& f) {
  auto ev = [&](double v) {
    return static_cast<double>(static_cast<float>(f(static_cast<Tx>(static_cast<float>(v)))));
  };
  if constexpr (std::is_same<Tx, bool>::value) {
    const bool t = static_cast<bool>(f(true)), z = static_cast<bool>(f(false));
    return UnaryInfo{(t && !z) ? kIdentity : kUnknownUnary, 0.0};
  } else {
    constexpr bool kSigned = !std::is_unsigned<Tx>::value;
    const double a = ev(kSigned ? -2.0 : 2.0), z = ev(0.0), b = ev(3.0);
    const double sa = kSigned ? -2.0 : 2.0;
    auto near = [](double u, double v) { return std::fabs(u - v) <= 0.02 * std::fabs(v) + 1e-30; };
    if (z == 0.0 && a == sa && b == 3.0) return UnaryInfo{kIdentity, 0.0};
    if (z == 0.0 && a == 2.0 && b == 3.0) return UnaryInfo{kAbs, 0.0};
    if (z == 0.0 && a == 4.0 && b == 9.0) return UnaryInfo{kSquare, 0.0};
    if (z == 0.0 && a == 1.0 && b == 1.0) return UnaryInfo{kNonzero, 0.0};
    if (z == 0.0 && a == -sa && b == -3.0) return UnaryInfo{kNegate, 0.0};
    if (z == 0.0 && b != 0.0 && near(a / sa, b / 3.0)) return UnaryInfo{kScale, b / 3.0};
    if (a > 0.0 && b > 0.0) {  // |x|^p: p from the probe at 2, confirmed at 3
      const double p = std::log2(a);
      if (p != 0.0 && near(std::pow(3.0, p), b) && (p < 0.0 || z == 0.0)) return UnaryInfo{kPowAbs, p};
    }
    return UnaryInfo{kUnknownUnary, 0.0};
  }
}

// ---- reducers -------------------------------------------------------------------------------------
template <template <typename> class ReduceOp>
struct KnownReduceOp {
  static constexpr int op = -1;
};
#define B200SHIM_REDUCE_OP(Functor, Id) \
  template <>                           \
  struct KnownReduceOp<Functor> {       \
    static constexpr int op = Id;       \
  };
B200SHIM_REDUCE_OP(kps::AddFunctor, B200PHI_SUM)
B200SHIM_REDUCE_OP(kps::MulFunctor, B200PHI_PROD)
B200SHIM_REDUCE_OP(kps::MaxFunctor, B200PHI_MAX)
B200SHIM_REDUCE_OP(kps::MinFunctor, B200PHI_MIN)
B200SHIM_REDUCE_OP(kps::LogicalOrFunctor, B200PHI_ANY)
B200SHIM_REDUCE_OP(kps::LogicalAndFunctor, B200PHI_ALL)
#undef B200SHIM_REDUCE_OP

template <template <typename> class ReduceOp, typename T>
int ReduceOpId() {
  if constexpr (KnownReduceOp<ReduceOp>::op >= 0) {
    return KnownReduceOp<ReduceOp>::op;
  } else {  // a client's own reducer: its identity element tells which of the family it is
    const T init = ReduceOp<T>().initial();
    if constexpr (std::is_same<T, bool>::value) {
      return init ? B200PHI_ALL : B200PHI_ANY;
    } else {
      const double v = static_cast<double>(static_cast<float>(init));
      if (v == 0.0) return B200PHI_SUM;
      if (v == 1.0) return B200PHI_PROD;
      if (v < 0.0) return B200PHI_MAX;
      return B200PHI_MIN;
    }
  }
}

inline void RunReduce(const GPUContext& dev_ctx, int op, const void* x, int in_dtype, const DDim& xdims,
                      const std::vector<int>& dims, void* out, int out_dtype) {
  b200phi_reduce_desc d = {};
  d.op = op;
  d.in_dtype = in_dtype;
  d.out_dtype = out_dtype;
  d.rank = xdims.size();
  for (int i = 0; i < d.rank; ++i) d.dims[i] = xdims[i];
  d.naxes = static_cast<int>(dims.size());
  for (int i = 0; i < d.naxes && i < B200PHI_MAX_RANK; ++i) d.axes[i] = dims[i];
  if (d.naxes == 0) {
    // no reduced dim: the reference's reduce_num == 1 branch, an element-wise transform
    // (reduce_function.h:1087-1092).  The C ABI reads naxes == 0 as "every axis", so the copy is
    // spelled as a reduce over an appended dim of extent 1.
    PADDLE_ENFORCE_LT(d.rank, B200PHI_MAX_RANK,
                      common::errors::Unimplemented("b200 shim: rank-9 reduce over no axis"));
    d.dims[d.rank] = 1;
    d.axes[0] = d.rank;
    d.rank += 1;
    d.naxes = 1;
  }
  d.reduce_all = 0;
  d.x = x;
  d.out = out;
  DenseTensor scratch;
  const size_t ws = b200phi_reduce_workspace(&d);
  if (ws > 0) {
    scratch.Resize({static_cast<int64_t>(ws)});
    d.workspace = dev_ctx.template Alloc<uint8_t>(&scratch);
    d.workspace_bytes = ws;
  }
  Check(b200phi_reduce(&d, dev_ctx.stream()));
}

// out = act(broadcast(in)) with one operand (act: NONE or SCALE)
inline void RunUnaryBroadcast(const GPUContext& dev_ctx, const b200phi_tensor_view& in, int dtype, int act,
                              float scale, DenseTensor* out) {
  b200phi_elementwise_desc e = {};
  e.binary = B200PHI_ADD;
  e.act = act;
  e.scale = scale;
  e.dtype = dtype;
  e.arity = 1;
  e.axis = -1;
  e.in[0] = in;
  e.out_rank = out->dims().size();
  for (int i = 0; i < e.out_rank; ++i) e.out_dims[i] = out->dims()[i];
  e.out = out->data();
  Check(b200phi_elementwise(&e, dev_ctx.stream()));
}

inline void RunMap(const GPUContext& dev_ctx, int op, const std::vector<const DenseTensor*>& ins, int axis,
                   float p0, DenseTensor* out) {
  b200phi_map_desc d = {};
  d.op = op;
  d.in_dtype = static_cast<int>(ins[0]->dtype());
  d.out_dtype = B200PHI_UNDEFINED;
  d.arity = static_cast<int>(ins.size());
  d.axis = axis;
  d.p0 = p0;
  for (size_t i = 0; i < ins.size(); ++i) View(*ins[i], &d.in[i]);
  d.out_rank = out->dims().size();
  for (int i = 0; i < d.out_rank; ++i) d.out_dims[i] = out->dims()[i];
  d.out = out->data();
  Check(b200phi_map(&d, dev_ctx.stream()));
}

// ---- binary / ternary functors: one trait per reference functor -------------------------------------
// engine 0: b200phi_elementwise binary id; 1: b200phi_map op id; -1: not mapped.  second: the op
// of a two-output functor's second output (NumOuts == 2), else -1.
template <typename Functor>
struct FunctorOp {
  static constexpr int engine = -1, id = 0, second = -1;
};
#define B200SHIM_FUNCTOR(Functor, Engine, Id)                         \
  template <typename T>                                               \
  struct FunctorOp<Functor<T>> {                                      \
    static constexpr int engine = Engine, id = Id, second = -1;       \
  };
B200SHIM_FUNCTOR(funcs::AddFunctor, 0, B200PHI_ADD)
B200SHIM_FUNCTOR(funcs::SubtractFunctor, 0, B200PHI_SUBTRACT)
B200SHIM_FUNCTOR(funcs::MultiplyFunctor, 0, B200PHI_MULTIPLY)
B200SHIM_FUNCTOR(funcs::MaximumFunctor, 0, B200PHI_MAXIMUM)
B200SHIM_FUNCTOR(funcs::MinimumFunctor, 0, B200PHI_MINIMUM)
B200SHIM_FUNCTOR(funcs::RemainderFunctor, 1, B200PHI_MAP_REMAINDER)
B200SHIM_FUNCTOR(funcs::FloorDivideFunctor, 1, B200PHI_MAP_FLOOR_DIVIDE)
B200SHIM_FUNCTOR(funcs::ElementwisePowFunctor, 1, B200PHI_MAP_POW)
B200SHIM_FUNCTOR(funcs::CopySignFunctor, 1, B200PHI_MAP_COPYSIGN)
B200SHIM_FUNCTOR(funcs::FMaxFunctor, 1, B200PHI_MAP_FMAX)
B200SHIM_FUNCTOR(funcs::FMinFunctor, 1, B200PHI_MAP_FMIN)
B200SHIM_FUNCTOR(funcs::ElementwiseHeavisideFunctor, 1, B200PHI_MAP_HEAVISIDE)
B200SHIM_FUNCTOR(funcs::LessThanFunctor, 1, B200PHI_MAP_LESS_THAN)
B200SHIM_FUNCTOR(funcs::LessEqualFunctor, 1, B200PHI_MAP_LESS_EQUAL)
B200SHIM_FUNCTOR(funcs::GreaterThanFunctor, 1, B200PHI_MAP_GREATER_THAN)
B200SHIM_FUNCTOR(funcs::GreaterEqualFunctor, 1, B200PHI_MAP_GREATER_EQUAL)
B200SHIM_FUNCTOR(funcs::LogicalAndFunctor, 1, B200PHI_MAP_LOGICAL_AND)
B200SHIM_FUNCTOR(funcs::LogicalOrFunctor, 1, B200PHI_MAP_LOGICAL_OR)
B200SHIM_FUNCTOR(funcs::LogicalXorFunctor, 1, B200PHI_MAP_LOGICAL_XOR)
B200SHIM_FUNCTOR(funcs::LogicalNotFunctor, 1, B200PHI_MAP_LOGICAL_NOT)
B200SHIM_FUNCTOR(funcs::BitwiseAndFunctor, 1, B200PHI_MAP_BITWISE_AND)
B200SHIM_FUNCTOR(funcs::BitwiseOrFunctor, 1, B200PHI_MAP_BITWISE_OR)
B200SHIM_FUNCTOR(funcs::BitwiseXorFunctor, 1, B200PHI_MAP_BITWISE_XOR)
B200SHIM_FUNCTOR(funcs::BitwiseNotFunctor, 1, B200PHI_MAP_BITWISE_NOT)
B200SHIM_FUNCTOR(funcs::MaxGradXFunctor, 1, B200PHI_MAP_MAX_GRAD_X)
B200SHIM_FUNCTOR(funcs::MaxGradYFunctor, 1, B200PHI_MAP_MAX_GRAD_Y)
B200SHIM_FUNCTOR(funcs::MinGradXFunctor, 1, B200PHI_MAP_MIN_GRAD_X)
B200SHIM_FUNCTOR(funcs::MinGradYFunctor, 1, B200PHI_MAP_MIN_GRAD_Y)
B200SHIM_FUNCTOR(funcs::DivGradYFunctor, 1, B200PHI_MAP_DIV_GRAD_Y)
#undef B200SHIM_FUNCTOR
// functors with extra template parameters
template <typename T, typename E>
struct FunctorOp<funcs::DivideFunctor<T, E>> {
  static constexpr int engine = 0, id = B200PHI_DIVIDE, second = -1;
};
template <typename T, typename E>
struct FunctorOp<funcs::EqualFunctor<T, E>> {
  static constexpr int engine = 1, id = B200PHI_MAP_EQUAL, second = -1;
};
template <typename T, typename E>
struct FunctorOp<funcs::NotEqualFunctor<T, E>> {
  static constexpr int engine = 1, id = B200PHI_MAP_NOT_EQUAL, second = -1;
};
// two outputs in one reference launch (gpu/elementwise_grad.h:38-79): two launches here
template <typename InT, typename OutT>
struct FunctorOp<funcs::MaxGradXYFunctor<InT, OutT>> {
  static constexpr int engine = 1, id = B200PHI_MAP_MAX_GRAD_X, second = B200PHI_MAP_MAX_GRAD_Y;
};
template <typename InT, typename OutT>
struct FunctorOp<funcs::MinGradXYFunctor<InT, OutT>> {
  static constexpr int engine = 1, id = B200PHI_MAP_MIN_GRAD_X, second = B200PHI_MAP_MIN_GRAD_Y;
};

// One call of either engine for an n-ary functor; `axis` already resolved by the caller's rule.
template <typename OutT, typename Functor, int NumOuts>
void RunFunctor(const GPUContext& dev_ctx, const std::vector<const DenseTensor*>& ins,
                std::vector<DenseTensor*>* outs, const Functor& func, int axis) {
  using Traits = phi::funcs::FunctionTraits<Functor>;
  constexpr int kArity = static_cast<int>(Traits::arity);
  PADDLE_ENFORCE_EQ(ins.size(), static_cast<size_t>(kArity),
                    common::errors::InvalidArgument(
                        "The number of inputs is expected to be equal to the arity of functor. But received: "
                        "the number of inputs is %d, the arity of functor is %d.",
                        ins.size(), kArity));
  PADDLE_ENFORCE_EQ(outs->size(), static_cast<size_t>(NumOuts),
                    common::errors::InvalidArgument(
                        "Number of outputs shall equal to number of functions, but number of outputs is %d, "
                        "of functions is %d.",
                        outs->size(), NumOuts));
  for (size_t i = 0; i < outs->size(); ++i) dev_ctx.template Alloc<OutT>((*outs)[i]);
  if constexpr (kArity == 1) {
    static_assert(NumOuts == 1, "b200_funcs_shim.h: a one-operand functor has one output");
    using InT = std::decay_t<std::tuple_element_t<0, typename Traits::ArgsTuple>>;
    const UnaryInfo u = ClassifyUnary<InT>(func);
    const DenseTensor& x = *ins[0];
    DenseTensor* out = (*outs)[0];
    b200phi_tensor_view v;
    View(x, &v);
    const int dt = static_cast<int>(x.dtype());
    PADDLE_ENFORCE_EQ(x.dtype(), out->dtype(),
                      common::errors::Unimplemented("b200 shim: a one-operand functor keeps the dtype"));
    switch (u.cls) {
      case kIdentity:
        RunUnaryBroadcast(dev_ctx, v, dt, B200PHI_ACT_NONE, 0.f, out);
        return;
      case kNegate:
        RunUnaryBroadcast(dev_ctx, v, dt, B200PHI_ACT_SCALE, -1.f, out);
        return;
      case kScale:
        RunUnaryBroadcast(dev_ctx, v, dt, B200PHI_ACT_SCALE, static_cast<float>(u.param), out);
        return;
      default:
        break;
    }
    PADDLE_ENFORCE_EQ(x.dims(), out->dims(),
                      common::errors::Unimplemented("b200 shim: this one-operand functor is not broadcast"));
    if (u.cls == kAbs) return RunMap(dev_ctx, B200PHI_MAP_ABS, ins, -1, 0.f, out);
    if (u.cls == kSquare) return RunMap(dev_ctx, B200PHI_MAP_SQUARE, ins, -1, 0.f, out);
    if (u.cls == kPowAbs) return RunMap(dev_ctx, B200PHI_MAP_POW_ABS, ins, -1, static_cast<float>(u.param), out);
    PADDLE_THROW(common::errors::Unimplemented(
        "b200 shim: this one-operand functor is none of identity, |x|, x*x, -x, k*x, |x|^p"));
  } else {
    using Op = FunctorOp<Functor>;
    static_assert(Op::engine >= 0,
                  "adapter/b200_funcs_shim.h: no FunctorOp<> trait maps this functor to a b200phi op id; "
                  "add one, or keep this client on the reference's own BroadcastKernel");
    static_assert(NumOuts == 1 || Op::second >= 0,
                  "adapter/b200_funcs_shim.h: this two-output functor is not mapped");
    if constexpr (Op::engine == 0) {
      const DenseTensor& x = *ins[0];
      b200phi_elementwise_desc e = {};
      e.binary = Op::id;
      e.dtype = static_cast<int>(x.dtype());
      e.arity = kArity;
      e.axis = axis;
      for (int i = 0; i < kArity; ++i) View(*ins[i], &e.in[i]);
      DenseTensor* out = (*outs)[0];
      e.out_rank = out->dims().size();
      for (int i = 0; i < e.out_rank; ++i) e.out_dims[i] = out->dims()[i];
      e.out = out->data();
      Check(b200phi_elementwise(&e, dev_ctx.stream()));
    } else {
      RunMap(dev_ctx, Op::id, ins, axis, 0.f, (*outs)[0]);
      if constexpr (NumOuts == 2) RunMap(dev_ctx, Op::second, ins, axis, 0.f, (*outs)[1]);
    }
  }
}

}  // namespace b200shim

namespace funcs {

// funcs/reduce_function.h:1047-1198
template <typename Tx, typename Ty, template <typename> class ReduceOp, typename TransformOp,
          bool IsMean = false>
void ReduceKernel(const KPDevice& dev_ctx, const phi::DenseTensor& x, phi::DenseTensor* y,
                  const TransformOp& transform, const std::vector<int>& origin_reduce_dims) {
  PADDLE_ENFORCE_GT(x.numel(), 0,
                    common::errors::InvalidArgument("Tensor need be reduced must not empty."));
  dev_ctx.template Alloc<Ty>(y);
  using MPType = typename phi::dtype::MPTypeTrait<Ty>::Type;
  const int base = b200shim::ReduceOpId<ReduceOp, MPType>();
  const b200shim::UnaryInfo u = b200shim::ClassifyUnary<Tx>(transform);
  const int in_dt = static_cast<int>(x.dtype()), out_dt = static_cast<int>(y->dtype());
  int op = -1;
  switch (u.cls) {
    case b200shim::kIdentity:
      op = (base == B200PHI_SUM && IsMean) ? B200PHI_MEAN : base;
      break;
    case b200shim::kAbs:
      op = base == B200PHI_SUM ? B200PHI_SUM_ABS
                               : (base == B200PHI_MAX ? B200PHI_MAX_ABS : (base == B200PHI_MIN ? B200PHI_MIN_ABS : -1));
      break;
    case b200shim::kSquare:
      op = base == B200PHI_SUM ? B200PHI_SUM_SQUARE : -1;
      break;
    case b200shim::kNonzero:
      op = base == B200PHI_SUM ? B200PHI_COUNT_NONZERO : -1;
      break;
    default:
      break;
  }
  if (op >= 0 && !(IsMean && op != B200PHI_MEAN)) {
    b200shim::RunReduce(dev_ctx, op, x.data(), in_dt, x.dims(), origin_reduce_dims, y->data(), out_dt);
    return;
  }
  if (base == B200PHI_SUM && !IsMean && in_dt == out_dt &&
      (u.cls == b200shim::kScale || u.cls == b200shim::kNegate)) {
    // sum(k * x) = k * sum(x): reduce, then scale the small result in place
    b200shim::RunReduce(dev_ctx, B200PHI_SUM, x.data(), in_dt, x.dims(), origin_reduce_dims, y->data(), out_dt);
    b200phi_tensor_view v;
    b200shim::View(*y, &v);
    const float k = u.cls == b200shim::kNegate ? -1.f : static_cast<float>(u.param);
    b200shim::RunUnaryBroadcast(dev_ctx, v, out_dt, B200PHI_ACT_SCALE, k, y);
    return;
  }
  if (base == B200PHI_SUM && !IsMean && in_dt == out_dt && u.cls == b200shim::kPowAbs) {
    // sum(|x|^p): the functor as one map launch into a temporary, then the sum
    // (the general-porder branch of gpu/p_norm_kernel.cu:90-99)
    DenseTensor tmp;
    tmp.Resize(x.dims());
    dev_ctx.template Alloc<Tx>(&tmp);
    b200shim::RunMap(dev_ctx, B200PHI_MAP_POW_ABS, {&x}, -1, static_cast<float>(u.param), &tmp);
    b200shim::RunReduce(dev_ctx, B200PHI_SUM, tmp.data(), in_dt, x.dims(), origin_reduce_dims, y->data(), out_dt);
    return;
  }
  PADDLE_THROW(common::errors::Unimplemented(
      "b200 shim: this (ReduceOp, TransformOp) pair has no b200phi_reduce op (reducer %d, transform class %d)",
      base, u.cls));
}

// funcs/broadcast_function.h:758-824
template <typename OutT, typename Functor, int NumOuts = 1>
void BroadcastKernel(const KPDevice& ctx, const std::vector<const DenseTensor*>& ins,
                     std::vector<DenseTensor*>* outs, Functor func, int axis = -1) {
  int max_rank = 0, min_rank = phi::DDim::kMaxRank;
  for (auto* in : ins) {
    max_rank = std::max(max_rank, in->dims().size());
    min_rank = std::min(min_rank, in->dims().size());
  }
  if (ins.size() == 1) max_rank = std::max(max_rank, (*outs)[0]->dims().size());
  // the C ABI takes the reference's own convention: -1 = right-aligned, >= 0 = the legacy offset
  (void)max_rank;
  (void)min_rank;
  b200shim::RunFunctor<OutT, Functor, NumOuts>(ctx, ins, outs, func, axis);
}

// funcs/elementwise_base.h:807-845 (same dims on every operand)
template <typename OutT, typename Functor, int NumOuts = 1>
void ElementwiseKernel(const KPDevice& ctx, const std::vector<const DenseTensor*>& ins,
                       std::vector<DenseTensor*>* outs, Functor func) {
  b200shim::RunFunctor<OutT, Functor, NumOuts>(ctx, ins, outs, func, -1);
}

}  // namespace funcs
}  // namespace phi

#pragma pop_macro("ElementwiseKernel")
#pragma pop_macro("BroadcastKernel")
#pragma pop_macro("ReduceKernel")
