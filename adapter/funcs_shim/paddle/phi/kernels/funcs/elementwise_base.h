// Shadow of paddle/phi/kernels/funcs/elementwise_base.h (see reduce_function.h next to this file):
// the reference header with ElementwiseKernel renamed while it is parsed, then the C-ABI entry point.
#pragma once
#define ElementwiseKernel ElementwiseKernel_reference
#include_next "paddle/phi/kernels/funcs/elementwise_base.h"
#undef ElementwiseKernel
#include "b200_funcs_shim.h"
