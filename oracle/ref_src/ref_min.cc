// TEST INFRASTRUCTURE -- reference-compiled oracle, op 'min' (see ref_common.h).
#include "ref_common.h"
REFPHI_EXPORT int phi_ref_min_float(const void* in, int rank, const int64_t* dims, int naxes,
    const int64_t* axes, int reduce_all, void* out) {
  return refphi::reduce_impl<float, phi::funcs::MinFunctor>(static_cast<const float*>(in), rank, dims, naxes, axes,
                                                       reduce_all != 0, static_cast<float*>(out));
}
REFPHI_EXPORT int phi_ref_min_double(const void* in, int rank, const int64_t* dims, int naxes,
    const int64_t* axes, int reduce_all, void* out) {
  return refphi::reduce_impl<double, phi::funcs::MinFunctor>(static_cast<const double*>(in), rank, dims, naxes, axes,
                                                       reduce_all != 0, static_cast<double*>(out));
}
REFPHI_EXPORT int phi_ref_min_int32(const void* in, int rank, const int64_t* dims, int naxes,
    const int64_t* axes, int reduce_all, void* out) {
  return refphi::reduce_impl<int32_t, phi::funcs::MinFunctor>(static_cast<const int32_t*>(in), rank, dims, naxes, axes,
                                                       reduce_all != 0, static_cast<int32_t*>(out));
}
REFPHI_EXPORT int phi_ref_min_int64(const void* in, int rank, const int64_t* dims, int naxes,
    const int64_t* axes, int reduce_all, void* out) {
  return refphi::reduce_impl<int64_t, phi::funcs::MinFunctor>(static_cast<const int64_t*>(in), rank, dims, naxes, axes,
                                                       reduce_all != 0, static_cast<int64_t*>(out));
}
