// TEST INFRASTRUCTURE -- reference-compiled oracle for the norm reductions: the Eigen
// expressions of the reference's CPU kernels, evaluated on Eigen::DefaultDevice over the same
// (pre, n, post) view:
//   p_norm            paddle/phi/kernels/cpu/p_norm_kernel.cc:70-98
//   squared_l2_norm   paddle/phi/kernels/funcs/squared_l2_norm.h:40-46
//   logsumexp         paddle/phi/kernels/impl/logsumexp_kernel_impl.h (LogsumexpFunctor: max, then
//                     (x - max).exp().sum().log() + max)
#include <math.h>

#include "paddle/phi/kernels/impl/logsumexp_kernel_impl.h"
#include "ref_common.h"

namespace refphi {

template <typename T>
int pnorm_impl(const T* x, int pre, int n, int post, float porder, T* out) {
  Eigen::DSizes<int, 3> shape(pre, n, post);
  Eigen::DSizes<int, 2> norm_shape(pre, post);
  Eigen::TensorMap<Eigen::Tensor<const T, 3, Eigen::RowMajor, int>> xr(x, shape);
  Eigen::TensorMap<Eigen::Tensor<T, 2, Eigen::RowMajor, int>> norm(out, norm_shape);
  Eigen::DefaultDevice place;
  Eigen::DSizes<int, 1> rdim(1);
  if (porder == 0) {
    norm.device(place) = (xr != xr.constant(0)).template cast<T>().sum(rdim);
  } else if (porder == INFINITY) {
    norm.device(place) = xr.abs().maximum(rdim);
  } else if (porder == -INFINITY) {
    norm.device(place) = xr.abs().minimum(rdim);
  } else {
    norm.device(place) = xr.abs().pow(porder).sum(rdim).pow(1.0f / porder);
  }
  return 0;
}

template <typename T>
int sql2_impl(const T* x, int64_t numel, T* out) {
  Eigen::DSizes<Eigen::DenseIndex, 1> d(numel);
  CTensor<T, 1> input(x, d);
  MScalar<T> output(out);  // the reference maps y as a 1-element vector; same evaluation
  Eigen::DefaultDevice place;
  output.device(place) = input.square().sum();
  return 0;
}

// phi::LogsumexpFunctor<T> itself (impl/logsumexp_kernel_impl.h:33-61) on the (pre, n, post)
// view, reducing n -- the call HANDLE_DIM(3, 1) makes through funcs::ReduceFunctor
template <typename T>
int lse_impl(const T* x, int pre, int n, int post, T* out) {
  Eigen::DSizes<Eigen::DenseIndex, 3> shape(pre, n, post);
  Eigen::DSizes<Eigen::DenseIndex, 2> oshape(pre, post);
  CTensor<T, 3> xr(x, shape);
  MTensor<T, 2> y(out, oshape);
  Eigen::DefaultDevice place;
  auto rdim = Eigen::array<int, 1>({{1}});
  phi::LogsumexpFunctor<T>()(place, &xr, &y, rdim);
  return 0;
}

}  // namespace refphi

#define REFPHI_NORM(T, name)                                                                    \
  REFPHI_EXPORT int phi_ref_pnorm_##name(const void* x, int pre, int n, int post, float porder, \
                                         void* out) {                                           \
    return refphi::pnorm_impl<T>(static_cast<const T*>(x), pre, n, post, porder,                \
                                 static_cast<T*>(out));                                         \
  }                                                                                             \
  REFPHI_EXPORT int phi_ref_squared_l2_norm_##name(const void* x, int64_t numel, void* out) {   \
    return refphi::sql2_impl<T>(static_cast<const T*>(x), numel, static_cast<T*>(out));         \
  }                                                                                             \
  REFPHI_EXPORT int phi_ref_logsumexp_##name(const void* x, int pre, int n, int post,           \
                                             void* out) {                                       \
    return refphi::lse_impl<T>(static_cast<const T*>(x), pre, n, post, static_cast<T*>(out));   \
  }
REFPHI_NORM(float, float)
REFPHI_NORM(double, double)
