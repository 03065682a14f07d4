// Reductions over ANY alternating keep/reduce pattern of merged dims -- the job of the
// reference's ReduceAnyKernel + IndexCalculator (paddle/phi/kernels/funcs/reduce_function.h:573-699,
// funcs/index_calculator.h:53-107) -- as two tiled kernels:
//   tile_keep_kernel  the innermost merged dim is KEPT ([R,K,R,K], [K,R,K,R,K], [R,K] with a short R ...):
//                     a thread owns one 16-byte vector of outputs and walks the flattened reduce
//                     space; ty and grid.y split that space, ticket finalise as in col_reduce_kernel
//   tile_red_kernel   the innermost merged dim is REDUCED ([K,R,K,R], [R,K,R] with a short inner run ...):
//                     a group of G threads owns one output and walks (outer reduce index, vector
//                     of the contiguous run) as one flat space; grid.y splits it
// Loads are 16 bytes wherever the contiguous dim allows it, U of them in flight per thread.
// Indices are unravelled with 32-bit multiply-shift divides (FastDiv); offsets are 64-bit.  Every
// access to the dims struct uses a compile-time index (fully unrolled loops, uniform predicates):
// a run-time index would make the compiler copy the kernel arguments to local memory.
#pragma once

#include "prim.cuh"

namespace b200phi {

constexpr int kTiledMaxDims = 5;  // rank 9 alternates into at most 5 dims of one role

struct TiledDims {
  int nkeep, nred;
  FastDiv keep_div[kTiledMaxDims];  // innermost last
  long long keep_stride[kTiledMaxDims];
  FastDiv red_div[kTiledMaxDims];
  long long red_stride[kTiledMaxDims];
};

// offset = sum over dims of (coordinate * stride), coordinates from `idx` (innermost dim last)
template <int kWhich>
__device__ __forceinline__ long long tiled_unravel(unsigned idx, const TiledDims& td) {
  long long off = 0;
#pragma unroll
  for (int k = 0; k < kTiledMaxDims; ++k) {
    const int n = kWhich == 0 ? td.nkeep : td.nred;
    const int d = n - 1 - k;  // walk from the innermost dim outwards
    if (d >= 0) {
      // dims are stored right-aligned at [kTiledMaxDims - n, kTiledMaxDims): constant index below
      const FastDiv& f = kWhich == 0 ? td.keep_div[kTiledMaxDims - 1 - k] : td.red_div[kTiledMaxDims - 1 - k];
      const long long s = kWhich == 0 ? td.keep_stride[kTiledMaxDims - 1 - k] : td.red_stride[kTiledMaxDims - 1 - k];
      const unsigned q = fast_div(idx, f);
      off += static_cast<long long>(idx - q * f.d) * s;
      idx = q;
    }
  }
  return off;
}

template <typename T, typename Acc, typename Op, int VE>
__device__ __forceinline__ void tiled_fold(const V16& raw, Acc (&acc)[VE]) {
  typename VecOf<T>::U w;
  w.raw = raw;
#pragma unroll
  for (int e = 0; e < VE; ++e) acc[e] = Op::combine(acc[e], load_as<Acc, Op>(w.e[e]));
}

template <typename T, typename Acc, typename Op, int VE, int UU>
__device__ __forceinline__ void tile_keep_step(const T* __restrict__ base, const TiledDims& td, unsigned& j,
                                               unsigned j1, unsigned TY, Acc (&acc)[VE]) {
  for (; j + (UU - 1) * TY < j1; j += UU * TY) {
    if constexpr (VE > 1) {
      V16 buf[UU];
#pragma unroll
      for (int u = 0; u < UU; ++u) buf[u] = ldg_stream16(base + tiled_unravel<1>(j + u * TY, td));
#pragma unroll
      for (int u = 0; u < UU; ++u) tiled_fold<T, Acc, Op, VE>(buf[u], acc);
    } else {
      T buf[UU];
#pragma unroll
      for (int u = 0; u < UU; ++u) buf[u] = ldg_scalar(base + tiled_unravel<1>(j + u * TY, td));
#pragma unroll
      for (int u = 0; u < UU; ++u) acc[0] = Op::combine(acc[0], load_as<Acc, Op>(buf[u]));
    }
  }
}

// Innermost dim kept.  Threads along x cover the flat index q of output VECTORS (the kept dims
// flattened, the innermost one in units of VE elements), ty and grid.y walk the flattened reduce
// space [0, R).  Block = TX x TY.  Shared memory: [TY][TX][VE] Acc.
template <typename T, typename Acc, typename Op, int VE, int U>
__global__ void __launch_bounds__(kBlock)
    tile_keep_kernel(const T* __restrict__ x, const TiledDims td, unsigned nq, int TX, unsigned R,
                     unsigned rows_per_split, Acc* __restrict__ partial, unsigned* __restrict__ tickets,
                     Epilogue ep) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Acc* smem = reinterpret_cast<Acc*>(smem_raw);
  __shared__ unsigned s_ticket;
  const unsigned TY = kBlock / TX;
  const unsigned tx = threadIdx.x % TX, ty = threadIdx.x / TX;
  const unsigned q = blockIdx.x * TX + tx;
  const bool live = q < nq;
  const unsigned j0 = blockIdx.y * rows_per_split;
  unsigned j1 = j0 + rows_per_split;
  if (j1 > R) j1 = R;
  const unsigned splits = gridDim.y;

  Acc acc[VE];
#pragma unroll
  for (int e = 0; e < VE; ++e) acc[e] = Op::identity();
  if (live) {
    const T* __restrict__ base = x + tiled_unravel<0>(q, td);
    unsigned j = j0 + ty;
    tile_keep_step<T, Acc, Op, VE, U>(base, td, j, j1, TY, acc);
    if constexpr (U > 4) tile_keep_step<T, Acc, Op, VE, 4>(base, td, j, j1, TY, acc);
    if constexpr (U > 2) tile_keep_step<T, Acc, Op, VE, 2>(base, td, j, j1, TY, acc);
    if constexpr (U > 1) tile_keep_step<T, Acc, Op, VE, 1>(base, td, j, j1, TY, acc);
  }

  auto tree = [&](void) {
#pragma unroll
    for (int e = 0; e < VE; ++e) smem[(ty * TX + tx) * VE + e] = acc[e];
    __syncthreads();
    for (unsigned half = TY >> 1; half > 0; half >>= 1) {
      if (ty < half) {
#pragma unroll
        for (int e = 0; e < VE; ++e)
          smem[(ty * TX + tx) * VE + e] =
              Op::combine(smem[(ty * TX + tx) * VE + e], smem[((ty + half) * TX + tx) * VE + e]);
      }
      __syncthreads();
    }
  };
  tree();

  const long long out_base = static_cast<long long>(q) * VE;
  if (splits == 1) {
    if (ty == 0 && live) {
#pragma unroll
      for (int e = 0; e < VE; ++e) store_final_16<Op, T>(ep, out_base + e, smem[tx * VE + e]);
    }
    return;
  }
  const long long out_numel = static_cast<long long>(nq) * VE;
  if (ty == 0 && live) {
#pragma unroll
    for (int e = 0; e < VE; ++e)
      partial[static_cast<long long>(blockIdx.y) * out_numel + out_base + e] = smem[tx * VE + e];
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(&tickets[blockIdx.x], 1u);
  __syncthreads();
  if (s_ticket != splits - 1) return;
  __threadfence();
  // last CTA for these outputs: the partials in index order (deterministic)
#pragma unroll
  for (int e = 0; e < VE; ++e) acc[e] = Op::identity();
  if (live) {
    unsigned s = ty;
    for (; s + 7u * TY < splits; s += 8u * TY) {
      Acc b[8][VE];
#pragma unroll
      for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int e = 0; e < VE; ++e)
          b[u][e] = __ldcg(&partial[static_cast<long long>(s + u * TY) * out_numel + out_base + e]);
#pragma unroll
      for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int e = 0; e < VE; ++e) acc[e] = Op::combine(acc[e], b[u][e]);
    }
    for (; s < splits; s += TY) {
#pragma unroll
      for (int e = 0; e < VE; ++e)
        acc[e] = Op::combine(acc[e], __ldcg(&partial[static_cast<long long>(s) * out_numel + out_base + e]));
    }
  }
  tree();
  if (ty == 0 && live) {
#pragma unroll
    for (int e = 0; e < VE; ++e) store_final_16<Op, T>(ep, out_base + e, smem[tx * VE + e]);
  }
  if (threadIdx.x == 0) tickets[blockIdx.x] = 0;
}

constexpr int kFoldLanesTiled = 4;

template <typename T, typename Acc, typename Op, int VE, int G, int UU>
__device__ __forceinline__ void tile_red_step(const T* __restrict__ base, const TiledDims& td,
                                              const FastDiv& lv_div, unsigned& f, unsigned f1,
                                              Acc (&acc)[kFoldLanesTiled]) {
  for (; f + (UU - 1) * G < f1; f += UU * G) {
    long long off[UU];
#pragma unroll
    for (int u = 0; u < UU; ++u) {
      const unsigned fu = f + u * G;
      const unsigned jo = fast_div(fu, lv_div);
      off[u] = tiled_unravel<1>(jo, td) + static_cast<long long>(fu - jo * lv_div.d) * VE;
    }
    if constexpr (VE > 1) {
      V16 buf[UU];
#pragma unroll
      for (int u = 0; u < UU; ++u) buf[u] = ldg_stream16(base + off[u]);
#pragma unroll
      for (int u = 0; u < UU; ++u) fold_vec<Acc, Op, T, kFoldLanesTiled>(buf[u], acc);
    } else {
      T buf[UU];
#pragma unroll
      for (int u = 0; u < UU; ++u) buf[u] = ldg_scalar(base + off[u]);
#pragma unroll
      for (int u = 0; u < UU; ++u) acc[u % kFoldLanesTiled] = Op::combine(acc[u % kFoldLanesTiled], load_as<Acc, Op>(buf[u]));
    }
  }
}

// Innermost dim reduced.  G threads (32: a warp, 256: the CTA) own one output and walk the flat
// space [0, W) of (outer reduce index, unit of the contiguous run); a unit is VE elements.
// grid.x = ceil(outputs / (256 / G)), grid.y = splits of W.
template <typename T, typename Acc, typename Op, int VE, int G, int U>
__global__ void __launch_bounds__(kBlock)
    tile_red_kernel(const T* __restrict__ x, const TiledDims td, const FastDiv lv_div, unsigned W,
                    unsigned chunk, unsigned out_numel, Acc* __restrict__ partial,
                    unsigned* __restrict__ tickets, Epilogue ep) {
  __shared__ Acc smem[kWarps];
  __shared__ unsigned s_ticket;
  constexpr int kOutPerCta = kBlock / G;
  const unsigned g = threadIdx.x % G, grp = threadIdx.x / G;
  const unsigned o = blockIdx.x * kOutPerCta + grp;
  const bool valid = o < out_numel;
  const unsigned f0 = blockIdx.y * chunk;
  unsigned f1 = f0 + chunk;
  if (f1 > W) f1 = W;
  const unsigned splits = gridDim.y;

  Acc acc[kFoldLanesTiled];
#pragma unroll
  for (int i = 0; i < kFoldLanesTiled; ++i) acc[i] = Op::identity();
  if (valid) {
    const T* __restrict__ base = x + tiled_unravel<0>(o, td);
    unsigned f = f0 + g;
    tile_red_step<T, Acc, Op, VE, G, U>(base, td, lv_div, f, f1, acc);
    if constexpr (U > 4) tile_red_step<T, Acc, Op, VE, G, 4>(base, td, lv_div, f, f1, acc);
    if constexpr (U > 2) tile_red_step<T, Acc, Op, VE, G, 2>(base, td, lv_div, f, f1, acc);
    if constexpr (U > 1) tile_red_step<T, Acc, Op, VE, G, 1>(base, td, lv_div, f, f1, acc);
  }
  Acc v = Op::combine(Op::combine(acc[0], acc[1]), Op::combine(acc[2], acc[3]));
  if constexpr (G == 32) {
    v = warp_combine<Acc, Op>(v);
  } else {
    v = block_combine<Acc, Op>(v, smem);
  }
  if (splits == 1) {
    if (g == 0 && valid) store_final_16<Op, T>(ep, o, v);
    return;
  }
  if (g == 0 && valid) partial[static_cast<long long>(blockIdx.y) * out_numel + o] = v;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(&tickets[blockIdx.x], 1u);
  __syncthreads();
  if (s_ticket != splits - 1) return;
  __threadfence();
  // last CTA for these outputs: group `grp` folds output o's partials in index order
  Acc r = Op::identity();
  if (valid) {
    for (unsigned s = g; s < splits; s += G)
      r = Op::combine(r, __ldcg(&partial[static_cast<long long>(s) * out_numel + o]));
  }
  if constexpr (G == 32) {
    r = warp_combine<Acc, Op>(r);
  } else {
    r = block_combine<Acc, Op>(r, smem);
  }
  if (g == 0 && valid) store_final_16<Op, T>(ep, o, r);
  if (threadIdx.x == 0) tickets[blockIdx.x] = 0;
}

}  // namespace b200phi
