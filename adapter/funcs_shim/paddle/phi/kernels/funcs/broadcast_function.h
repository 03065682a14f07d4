// Shadow of paddle/phi/kernels/funcs/broadcast_function.h (see reduce_function.h next to this file):
// the reference header with BroadcastKernel renamed while it is parsed, then the C-ABI entry point.
#pragma once
#define BroadcastKernel BroadcastKernel_reference
#include_next "paddle/phi/kernels/funcs/broadcast_function.h"
#undef BroadcastKernel
#include "b200_funcs_shim.h"
