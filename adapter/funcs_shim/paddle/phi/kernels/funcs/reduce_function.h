// Shadow of paddle/phi/kernels/funcs/reduce_function.h for a Paddle build that links libb200phi.so:
// put adapter/funcs_shim in front of the include path.  The reference's own header is included
// unmodified from where it lies (every helper in it stays available: ReduceConfig, GetReduceDim,
// ...), with its entry point renamed while it is parsed; phi::funcs::ReduceKernel itself then comes
// from adapter/b200_funcs_shim.h.
#pragma once
#define ReduceKernel ReduceKernel_reference
#include_next "paddle/phi/kernels/funcs/reduce_function.h"
#undef ReduceKernel
#include "b200_funcs_shim.h"
