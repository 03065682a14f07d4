// TEST INFRASTRUCTURE -- reference-compiled oracle, op 'any' (see ref_common.h).
#include "ref_common.h"
REFPHI_EXPORT int phi_ref_any_bool(const void* in, int rank, const int64_t* dims, int naxes,
    const int64_t* axes, int reduce_all, void* out) {
  return refphi::reduce_impl<bool, phi::funcs::AnyFunctor>(static_cast<const bool*>(in), rank, dims, naxes, axes,
                                                       reduce_all != 0, static_cast<bool*>(out));
}
