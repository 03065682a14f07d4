// TEST INFRASTRUCTURE -- reference-compiled oracle for argmax/argmin: the Eigen
// expression the CPU kernel evaluates (paddle/phi/kernels/cpu/arg_min_max_kernel.cc:35-66,
// rank switch :108-140, flatten handling :96-106).
#include "ref_common.h"

namespace refphi {

template <typename T, int Rank>
void arg_rank(const T* in, const int64_t* dims, int64_t axis, bool is_max, int64_t* out) {
  Eigen::DSizes<Eigen::DenseIndex, Rank> xd;
  for (int i = 0; i < Rank; ++i) xd[i] = dims[i];
  CTensor<T, Rank> x(in, xd);
  Eigen::DefaultDevice dev;
  Eigen::DSizes<Eigen::DenseIndex, Rank - 1> od;
  int k = 0;
  for (int i = 0; i < Rank; ++i)
    if (i != axis) od[k++] = dims[i];
  MTensor<int64_t, Rank - 1> o(out, od);
  if (is_max)
    o.device(dev) = x.argmax(axis).template cast<int64_t>();
  else
    o.device(dev) = x.argmin(axis).template cast<int64_t>();
}

template <typename T>
int arg_impl(const T* in, int rank, const int64_t* dims, int64_t axis, int flatten, int is_max,
             int64_t* out) {
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) numel *= dims[i];
  if (numel <= 0) return 1;
  if (rank == 0) {  // set_constant(out, 0)
    out[0] = 0;
    return 0;
  }
  if (flatten) {
    int64_t d1[1] = {numel};
    arg_rank<T, 1>(in, d1, 0, is_max != 0, out);
    return 0;
  }
  if (axis < 0) axis += rank;
  switch (rank) {
    case 1: arg_rank<T, 1>(in, dims, axis, is_max != 0, out); return 0;
    case 2: arg_rank<T, 2>(in, dims, axis, is_max != 0, out); return 0;
    case 3: arg_rank<T, 3>(in, dims, axis, is_max != 0, out); return 0;
    case 4: arg_rank<T, 4>(in, dims, axis, is_max != 0, out); return 0;
    case 5: arg_rank<T, 5>(in, dims, axis, is_max != 0, out); return 0;
    case 6: arg_rank<T, 6>(in, dims, axis, is_max != 0, out); return 0;
    default: return 2;  // "doesn't supports tensors whose ranks are greater than 6"
  }
}

}  // namespace refphi

#define REFPHI_ARG(T, name)                                                                     \
  REFPHI_EXPORT int phi_ref_arg_##name(const void* in, int rank, const int64_t* dims,           \
                                       int64_t axis, int flatten, int is_max, int64_t* out) {   \
    return refphi::arg_impl<T>(static_cast<const T*>(in), rank, dims, axis, flatten, is_max, out); \
  }
REFPHI_ARG(float, float)
REFPHI_ARG(double, double)
REFPHI_ARG(int32_t, int32)
REFPHI_ARG(int64_t, int64)
