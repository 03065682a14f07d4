/* TEST INFRASTRUCTURE -- stand-in for the header CMake generates from
 * third_party/openblas/lapack-netlib/CBLAS/include/cblas_mangling_with_flags.h.in: the reference
 * CBLAS header wants the Fortran name-mangling macros; ref_gpu_bench never calls BLAS. */
#ifndef CBLAS_MANGLING_H
#define CBLAS_MANGLING_H
#define F77_GLOBAL(lcname, UCNAME) lcname##_
#endif
