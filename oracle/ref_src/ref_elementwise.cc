// TEST INFRASTRUCTURE -- reference-compiled oracle for the broadcast binary ops and
// the fused add+activation: the reference's own scalar functors
// (paddle/phi/kernels/funcs/elementwise_functor.h:39+, funcs/functors.h:36-64) applied
// at the index pairs ElementwiseCompute(CPUContext) visits
// (elementwise_base.h:366-442; CommonForwardBroadcastCPU :273-307).  Which of its
// three iteration schemes (same dims / RowWise+MidWise / common) the CPU kernel
// picks does not change any value: each output is one functor application.
#include <math.h>

#include "paddle/common/array.h"
#include "paddle/phi/common/data_type.h"
#include "paddle/phi/kernels/funcs/elementwise_functor.h"
#include "paddle/phi/kernels/funcs/functors.h"
#include "ref_common.h"

namespace refphi {

// unary ids follow include/b200phi.h
template <typename T>
T unary(int act, T v, float scale) {
  switch (act) {
    case 1:  // ReluCPUFunctor, activation_functor.h:1473-1480
      return v > static_cast<T>(0) ? v : static_cast<T>(0);
    case 2:
      return phi::funcs::ScaleFunctor<T>(scale)(v);
    case 3:  // GeluFunctor (exact form), activation_functor.h
      return static_cast<T>(0.5) * v * (static_cast<T>(1) + std::erf(v * static_cast<T>(M_SQRT1_2)));
    case 4:
      return static_cast<T>(1) / (static_cast<T>(1) + std::exp(-v));
    case 5:
      return std::tanh(v);
    default:
      return v;
  }
}

template <typename T>
T binary(int op, T a, T b) {
  switch (op) {
    case 0: return phi::funcs::AddFunctor<T>()(a, b);
    case 1: return phi::funcs::SubtractFunctor<T>()(a, b);
    case 2: return phi::funcs::MultiplyFunctor<T>()(a, b);
    case 3: return a / b;  // DivideFunctor; the integer form adds a zero check that throws
    case 4: return phi::funcs::MaximumFunctor<T>()(a, b);
    default: return phi::funcs::MinimumFunctor<T>()(a, b);
  }
}

// in_dims[i] are already padded to out_rank (1 where broadcast)
template <typename T>
int elementwise_impl(int op, int act, int compound, float scale, int arity, int rank,
                     const int64_t* out_dims, const int64_t* const* in_dims,
                     const void* const* ins, void* out, void* inter) {
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) numel *= out_dims[i];
  std::vector<int64_t> idx(rank > 0 ? rank : 1, 0);
  T* o = static_cast<T*>(out);
  T* it = static_cast<T*>(inter);
  for (int64_t n = 0; n < numel; ++n) {
    T v[3] = {T(0), T(0), T(0)};
    for (int i = 0; i < arity; ++i) {
      int64_t off = 0;
      for (int k = 0; k < rank; ++k) off = off * in_dims[i][k] + (in_dims[i][k] == 1 ? 0 : idx[k]);
      v[i] = static_cast<const T*>(ins[i])[off];
    }
    T t, r;
    if (arity == 1) {
      t = v[0];
      r = unary<T>(act, t, scale);
    } else if (compound) {
      t = binary<T>(op, v[0], unary<T>(act, v[1], scale));
      r = t;
    } else {
      t = binary<T>(op, v[0], v[1]);
      if (arity == 3) t = binary<T>(op, t, v[2]);
      r = unary<T>(act, t, scale);
    }
    o[n] = r;
    if (it) it[n] = t;
    for (int k = rank - 1; k >= 0; --k) {  // UpdateElementwiseIndexArray
      if (++idx[k] < out_dims[k]) break;
      idx[k] = 0;
    }
  }
  return 0;
}

}  // namespace refphi

#define REFPHI_EW(T, name)                                                                        \
  REFPHI_EXPORT int phi_ref_elementwise_##name(int op, int act, int compound, float scale,        \
                                               int arity, int rank, const int64_t* out_dims,      \
                                               const int64_t* const* in_dims,                     \
                                               const void* const* ins, void* out, void* inter) {  \
    return refphi::elementwise_impl<T>(op, act, compound, scale, arity, rank, out_dims, in_dims,  \
                                       ins, out, inter);                                          \
  }
REFPHI_EW(float, float)
REFPHI_EW(double, double)
REFPHI_EW(int32_t, int32)
REFPHI_EW(int64_t, int64)

// Mixed-precision add: funcs::MultiPrecisionAddFunctor<float, bfloat16 / float16>
// (elementwise_functor.h:45-50), y given as raw 16-bit patterns, same shape.
#include "paddle/phi/common/bfloat16.h"
#include "paddle/phi/common/float16.h"
REFPHI_EXPORT int phi_ref_mixed_add(int y_is_bf16, const float* x, const uint16_t* y_bits, int64_t n,
                                    float* out) {
  for (int64_t i = 0; i < n; ++i) {
    if (y_is_bf16) {
      phi::dtype::bfloat16 y;
      y.x = y_bits[i];
      out[i] = phi::funcs::MultiPrecisionAddFunctor<float, phi::dtype::bfloat16>()(x[i], y);
    } else {
      phi::dtype::float16 y;
      y.x = y_bits[i];
      out[i] = phi::funcs::MultiPrecisionAddFunctor<float, phi::dtype::float16>()(x[i], y);
    }
  }
  return 0;
}
