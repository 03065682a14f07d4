// TEST INFRASTRUCTURE -- reference-compiled oracle, op 'mean' (see ref_common.h).
#include "ref_common.h"
REFPHI_EXPORT int phi_ref_mean_float(const void* in, int rank, const int64_t* dims, int naxes,
    const int64_t* axes, int reduce_all, void* out) {
  return refphi::reduce_impl<float, phi::funcs::MeanFunctor>(static_cast<const float*>(in), rank, dims, naxes, axes,
                                                       reduce_all != 0, static_cast<float*>(out));
}
REFPHI_EXPORT int phi_ref_mean_double(const void* in, int rank, const int64_t* dims, int naxes,
    const int64_t* axes, int reduce_all, void* out) {
  return refphi::reduce_impl<double, phi::funcs::MeanFunctor>(static_cast<const double*>(in), rank, dims, naxes, axes,
                                                       reduce_all != 0, static_cast<double*>(out));
}
