// TEST INFRASTRUCTURE -- reference-compiled oracle for the functor plug-ins of b200phi_map:
// the reference's OWN functors, included from where they lie, applied at the index tuples the
// broadcast visits.
//   compare     phi::funcs::{Equal,NotEqual,LessThan,...}Functor   funcs/compare_functors.h:20-60
//   logical     phi::funcs::Logical{And,Or,Xor,Not}Functor         funcs/logical_functor.h:20-42
//   bitwise     phi::funcs::Bitwise{And,Or,Xor,Not}Functor         funcs/bitwise_functors.h:20-48
//   backward    Max/MinGrad{X,Y}Functor, DivGradYFunctor           funcs/elementwise_functor.h:173-179,478-531
//               Equal then Multiply                                gpu/reduce_kernel.cu:74-119
//   activation  the CPU Eigen functors of funcs/activation_functor.h, run on a 1-D tensor map
//               with Eigen::DefaultDevice as ActivationImpl does   impl/activation_impl.h
#include <math.h>

#include <type_traits>

#include "paddle/phi/kernels/funcs/activation_functor.h"
#include "paddle/phi/kernels/funcs/bitwise_functors.h"
#include "paddle/phi/kernels/funcs/compare_functors.h"
#include "paddle/phi/kernels/funcs/elementwise_functor.h"
#include "paddle/phi/kernels/funcs/logical_functor.h"
#include "ref_common.h"

namespace refphi {

template <typename T, typename F>
void run_act(F f, const T* x, T* y, int64_t n) {
  Eigen::DSizes<Eigen::DenseIndex, 1> d(n);
  CTensor<T, 1> xm(x, d);
  MTensor<T, 1> ym(y, d);
  Eigen::DefaultDevice dev;
  f(dev, xm, ym);
}

// unary ops on a dense vector (no broadcast to do): op ids follow include/b200phi.h
template <typename T>
int act_impl(int op, float p0, float p1, const T* x, T* y, int64_t n) {
  using namespace phi::funcs;  // NOLINT
  if constexpr (std::is_floating_point<T>::value) {
    switch (op) {
      case 64: run_act<T>(ExpFunctor<T>(), x, y, n); return 0;
      case 65: run_act<T>(LogFunctor<T>(), x, y, n); return 0;
      case 66: run_act<T>(SqrtFunctor<T>(), x, y, n); return 0;
      case 67: run_act<T>(RsqrtFunctor<T>(), x, y, n); return 0;
      case 68: run_act<T>(SquareFunctor<T>(), x, y, n); return 0;
      case 69: run_act<T>(ReciprocalFunctor<T>(), x, y, n); return 0;
      case 70: run_act<T>(SinFunctor<T>(), x, y, n); return 0;
      case 71: run_act<T>(CosFunctor<T>(), x, y, n); return 0;
      case 72: run_act<T>(TanhFunctor<T>(), x, y, n); return 0;
      case 73: run_act<T>(SigmoidFunctor<T>(), x, y, n); return 0;
      case 74: run_act<T>(SiluFunctor<T>(), x, y, n); return 0;
      case 75: { LeakyReluFunctor<T> f; f.alpha = p0; run_act<T>(f, x, y, n); return 0; }
      case 76: { ELUFunctor<T> f; f.alpha = p0; run_act<T>(f, x, y, n); return 0; }
      case 77: { SoftplusFunctor<T> f; f.beta = p0; f.threshold = p1; run_act<T>(f, x, y, n); return 0; }
      case 78: run_act<T>(FloorFunctor<T>(), x, y, n); return 0;
      case 79: run_act<T>(CeilFunctor<T>(), x, y, n); return 0;
      case 80: run_act<T>(Expm1Functor<T>(), x, y, n); return 0;
      case 81: run_act<T>(Log1pFunctor<T>(), x, y, n); return 0;
      case 82: run_act<T>(Log2Functor<T>(), x, y, n); return 0;
      case 83: run_act<T>(Log10Functor<T>(), x, y, n); return 0;
      case 84: { Relu6Functor<T> f; f.threshold = p0; run_act<T>(f, x, y, n); return 0; }
      case 85: { HardSigmoidFunctor<T> f; f.slope = p0; f.offset = p1; run_act<T>(f, x, y, n); return 0; }
      case 86: { SwishFunctor<T> f; f.beta = p0; run_act<T>(f, x, y, n); return 0; }
      case 87: { MishFunctor<T> f; f.threshold = p0; run_act<T>(f, x, y, n); return 0; }
      case 88: run_act<T>(SoftsignFunctor<T>(), x, y, n); return 0;
      case 89: run_act<T>(LogSigmoidFunctor<T>(), x, y, n); return 0;
      case 91: {  // cpu/p_norm_kernel.cc: x.abs().pow(porder)
        Eigen::DSizes<Eigen::DenseIndex, 1> d(n);
        CTensor<T, 1> xm(x, d);
        MTensor<T, 1> ym(y, d);
        ym = xm.abs().pow(static_cast<T>(p0));
        return 0;
      }
      default: break;
    }
  }
  switch (op) {
    case 90:  // AbsFunctor (funcs/complex_functors.h): std::abs
      for (int64_t i = 0; i < n; ++i) y[i] = x[i] < T(0) ? -x[i] : x[i];
      return 0;
    case 92:  // ReluCPUFunctor, activation_functor.h:1473-1480
      for (int64_t i = 0; i < n; ++i) y[i] = x[i] > T(0) ? x[i] : T(0);
      return 0;
    default:
      return 2;
  }
}

// binary / ternary ops with broadcasting; in_dims[i] padded to rank
template <typename T>
int map_impl(int op, int rank, const int64_t* out_dims, const int64_t* const* in_dims,
             const void* const* ins, void* out) {
  using namespace phi::funcs;  // NOLINT
  const int arity = (op == 19 || op == 35) ? 1 : ((op >= 48 && op <= 53) ? 3 : 2);
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) numel *= out_dims[i];
  std::vector<int64_t> idx(rank > 0 ? rank : 1, 0);
  bool* ob = static_cast<bool*>(out);
  T* ot = static_cast<T*>(out);
  for (int64_t n = 0; n < numel; ++n) {
    T v[3] = {T(0), T(0), T(0)};
    for (int i = 0; i < arity; ++i) {
      int64_t off = 0;
      for (int k = 0; k < rank; ++k) off = off * in_dims[i][k] + (in_dims[i][k] == 1 ? 0 : idx[k]);
      v[i] = static_cast<const T*>(ins[i])[off];
    }
    const T a = v[0], b = v[1], c = v[2];
    switch (op) {
      case 0: ob[n] = EqualFunctor<T>()(a, b); break;
      case 1: ob[n] = NotEqualFunctor<T>()(a, b); break;
      case 2: ob[n] = LessThanFunctor<T>()(a, b); break;
      case 3: ob[n] = LessEqualFunctor<T>()(a, b); break;
      case 4: ob[n] = GreaterThanFunctor<T>()(a, b); break;
      case 5: ob[n] = GreaterEqualFunctor<T>()(a, b); break;
      case 16: ob[n] = LogicalAndFunctor<T>()(a, b); break;
      case 17: ob[n] = LogicalOrFunctor<T>()(a, b); break;
      case 18: ob[n] = LogicalXorFunctor<T>()(a, b); break;
      case 19: ob[n] = LogicalNotFunctor<T>()(a); break;
      case 48:  // equal_out = Equal(out, x) as T; dx = Multiply(dout, equal_out)
        ot[n] = MultiplyFunctor<T>()(c, static_cast<T>(EqualFunctor<T>()(b, a)));
        break;
      case 49: ot[n] = MaxGradXFunctor<T>()(a, b, c); break;
      case 50: ot[n] = MaxGradYFunctor<T>()(a, b, c); break;
      case 51: ot[n] = MinGradXFunctor<T>()(a, b, c); break;
      case 52: ot[n] = MinGradYFunctor<T>()(a, b, c); break;
      case 53: ot[n] = DivGradYFunctor<T>()(a, b, c); break;
      case 56: ot[n] = static_cast<T>(EqualFunctor<T>()(a, b)); break;
      // the other binary functors of kps/elementwise_kernel.cu:183-293 (elementwise_functor.h)
      case 54: ot[n] = ElementwiseHeavisideFunctor<T>()(a, b); break;
      case 61: ot[n] = CopySignFunctor<T>()(a, b); break;
      case 57:
        if constexpr (std::is_floating_point<T>::value) ot[n] = std::exp(a - b);
        else return 2;
        break;
      default:
        if constexpr (!std::is_same<T, bool>::value) {
          bool done = true;
          switch (op) {
            case 58: ot[n] = RemainderFunctor<T>()(a, b); break;
            case 59: ot[n] = FloorDivideFunctor<T>()(a, b); break;
            case 60: ot[n] = ElementwisePowFunctor<T>()(a, b); break;
            case 62: ot[n] = FMaxFunctor<T>()(a, b); break;
            case 63: ot[n] = FMinFunctor<T>()(a, b); break;
            default: done = false; break;
          }
          if (done) break;
        }
        if constexpr (std::is_integral<T>::value) {
          switch (op) {
            case 32: ot[n] = BitwiseAndFunctor<T>()(a, b); break;
            case 33: ot[n] = BitwiseOrFunctor<T>()(a, b); break;
            case 34: ot[n] = BitwiseXorFunctor<T>()(a, b); break;
            case 35: ot[n] = BitwiseNotFunctor<T>()(a); break;
            case 36: ot[n] = BitwiseLeftShiftArithmeticFunctor<T>()(a, b); break;
            case 37: ot[n] = BitwiseLeftShiftLogicFunctor<T>()(a, b); break;
            case 38: ot[n] = BitwiseRightShiftArithmeticFunctor<T>()(a, b); break;
            case 39: ot[n] = BitwiseRightShiftLogicFunctor<T>()(a, b); break;
            default: return 2;
          }
        } else {
          return 2;
        }
    }
    for (int k = rank - 1; k >= 0; --k) {
      if (++idx[k] < out_dims[k]) break;
      idx[k] = 0;
    }
  }
  return 0;
}

}  // namespace refphi

#define REFPHI_MAP(T, name)                                                                     \
  REFPHI_EXPORT int phi_ref_map_##name(int op, int rank, const int64_t* out_dims,               \
                                       const int64_t* const* in_dims, const void* const* ins,   \
                                       void* out) {                                             \
    return refphi::map_impl<T>(op, rank, out_dims, in_dims, ins, out);                          \
  }                                                                                             \
  REFPHI_EXPORT int phi_ref_act_##name(int op, float p0, float p1, const void* x, void* y,      \
                                       int64_t n) {                                             \
    return refphi::act_impl<T>(op, p0, p1, static_cast<const T*>(x), static_cast<T*>(y), n);    \
  }
REFPHI_MAP(float, float)
REFPHI_MAP(double, double)
REFPHI_MAP(int32_t, int32)
REFPHI_MAP(int64_t, int64)
REFPHI_MAP(int16_t, int16)
REFPHI_MAP(int8_t, int8)
REFPHI_MAP(uint8_t, uint8)
REFPHI_MAP(bool, bool)
