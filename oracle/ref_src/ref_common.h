// TEST INFRASTRUCTURE -- not product code.
// Glue that drives the REFERENCE's own CPU arithmetic for the hot path, compiled
// from the headers where they lie under /root/reference (never copied):
//   phi::funcs::{Sum,Mean,Max,Min,Prod,Any,All}Functor   paddle/phi/kernels/funcs/reduce_functor.h:52-105
//   on EigenTensor maps with Eigen::DefaultDevice, driven exactly as
//   ReduceFunctor / ReduceKernelImpl / HandleLargeDim do   reduce_function.h:1223-1402
//   Eigen argmax/argmin                                    cpu/arg_min_max_kernel.cc:35-66
//   phi::funcs::{Add,...}Functor                           elementwise_functor.h:39+
// Output: oracle/_ref/libphi_ref.so (git-ignored, travels to the GPU box).
#pragma once
#include <stdint.h>

#include <memory>
#include <vector>

#include "paddle/phi/kernels/funcs/reduce_functor.h"

namespace refphi {

template <typename T, int D>
using CTensor = Eigen::TensorMap<Eigen::Tensor<const T, D, Eigen::RowMajor, Eigen::DenseIndex>>;
template <typename T, int D>
using MTensor = Eigen::TensorMap<Eigen::Tensor<T, D, Eigen::RowMajor, Eigen::DenseIndex>>;
template <typename T>
using MScalar =
    Eigen::TensorMap<Eigen::TensorFixedSize<T, Eigen::Sizes<>, Eigen::RowMajor, Eigen::DenseIndex>>;

// ReduceFunctor<Context, T, D, R_D, Functor>  (reduce_function.h:1223-1259)
template <typename T, int D, int R_D, typename Functor>
void reduce_functor(const T* in, const int64_t* dims, T* out, const int64_t* red_dims) {
  Eigen::DSizes<Eigen::DenseIndex, D> xd;
  for (int i = 0; i < D; ++i) xd[i] = dims[i];
  CTensor<T, D> x(in, xd);
  auto reduce_dim = Eigen::array<int, R_D>();
  bool is_red[D > 0 ? D : 1] = {false};
  for (int i = 0; i < R_D; ++i) {
    int64_t r = red_dims[i];
    if (r < 0) r += D;
    reduce_dim[i] = static_cast<int>(r);
    is_red[r] = true;
  }
  Eigen::DefaultDevice place;
  Functor functor;
  if constexpr (D == 1) {
    MScalar<T> o(out);
    functor(place, &x, &o, reduce_dim);
  } else {
    Eigen::DSizes<Eigen::DenseIndex, D - R_D> od;
    int k = 0;
    for (int i = 0; i < D; ++i)
      if (!is_red[i]) od[k++] = dims[i];
    MTensor<T, D - R_D> o(out, od);
    functor(place, &x, &o, reduce_dim);
  }
}

// ReduceKernelImpl (reduce_function.h:1354-1402) incl. HandleLargeDim (:1316-1349)
template <typename T, typename Functor>
int reduce_impl(const T* in, int rank, const int64_t* dims, int naxes, const int64_t* axes,
                bool reduce_all, T* out) {
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) numel *= dims[i];
  if (numel <= 0) return 1;
  {  // phi::Reduce: "if the dims has full dim, set the reduce_all is True" (cpu/reduce.h:33-46)
    std::vector<bool> seen(rank > 0 ? rank : 1, false);
    for (int i = 0; i < naxes; ++i) {
      const int64_t r = axes[i] < 0 ? axes[i] + rank : axes[i];
      if (r >= 0 && r < rank) seen[r] = true;
    }
    bool full = true;
    for (int i = 0; i < rank; ++i) full = full && seen[i];
    if (full || naxes == 0) reduce_all = true;  // kernel_utils.h:405-416 (empty dims, 0-D)
  }
  if (reduce_all) {
    int64_t d1[1] = {numel};
    int64_t r1[1] = {0};
    reduce_functor<T, 1, 1, Functor>(in, d1, out, r1);
    return 0;
  }
  if (rank > 6) {
    // GetShuffledInput: reduced dims go last, then a 2-D reduce over axis 1
    std::vector<bool> is_red(rank, false);
    std::vector<int> perm(rank);
    std::vector<int64_t> sd(rank);
    for (int i = 0; i < naxes; ++i) {
      int64_t r = axes[i] < 0 ? axes[i] + rank : axes[i];
      sd[rank - naxes + i] = dims[r];
      perm[rank - naxes + i] = static_cast<int>(r);
      is_red[r] = true;
    }
    int off = 0;
    for (int i = 0; i < rank; ++i)
      if (!is_red[i]) {
        perm[off] = i;
        sd[off++] = dims[i];
      }
    // TransposeNormal: plain permutation copy
    std::vector<int64_t> in_stride(rank), out_stride(rank);
    in_stride[rank - 1] = out_stride[rank - 1] = 1;
    for (int i = rank - 2; i >= 0; --i) {
      in_stride[i] = in_stride[i + 1] * dims[i + 1];
      out_stride[i] = out_stride[i + 1] * sd[i + 1];
    }
    std::unique_ptr<T[]> shuffled(new T[numel]);
    for (int64_t o = 0; o < numel; ++o) {
      int64_t rem = o, src = 0;
      for (int i = 0; i < rank; ++i) {
        const int64_t c = rem / out_stride[i];
        rem -= c * out_stride[i];
        src += c * in_stride[perm[i]];
      }
      shuffled[o] = in[src];
    }
    int64_t reduced = 1;
    for (int i = 0; i < naxes; ++i) reduced *= sd[rank - naxes + i];
    int64_t d2[2] = {numel / reduced, reduced};
    int64_t r2[1] = {1};
    reduce_functor<T, 2, 1, Functor>(shuffled.get(), d2, out, r2);
    return 0;
  }
#define REFPHI_HANDLE(NDIM, RDIM)                                         \
  if (rank == NDIM && naxes == RDIM) {                                    \
    reduce_functor<T, NDIM, RDIM, Functor>(in, dims, out, axes);          \
    return 0;                                                             \
  }
  REFPHI_HANDLE(6, 5) REFPHI_HANDLE(6, 4) REFPHI_HANDLE(6, 3) REFPHI_HANDLE(6, 2)
  REFPHI_HANDLE(6, 1) REFPHI_HANDLE(5, 4) REFPHI_HANDLE(5, 3) REFPHI_HANDLE(5, 2)
  REFPHI_HANDLE(5, 1) REFPHI_HANDLE(4, 3) REFPHI_HANDLE(4, 2) REFPHI_HANDLE(4, 1)
  REFPHI_HANDLE(3, 2) REFPHI_HANDLE(3, 1) REFPHI_HANDLE(2, 1) REFPHI_HANDLE(1, 1)
#undef REFPHI_HANDLE
  return 2;  // rank/naxes pair the reference does not dispatch (e.g. naxes == rank without reduce_all)
}

}  // namespace refphi

#define REFPHI_EXPORT extern "C" __attribute__((visibility("default")))
