// The tiled kernels (general keep/reduce patterns, reduce_tiled.cuh) of this op family, in a
// translation unit of their own so that the build stays parallel.
#define B200PHI_TILED_TU 1
#include "reduce_minmax.cu"
