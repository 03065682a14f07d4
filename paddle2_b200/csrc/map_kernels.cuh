// Kernels of b200phi_map: the [rows, inner] broadcast walk of elementwise.cu with an output
// type that may be narrower than the input type (compare / logical -> BOOL) and three
// operands.  One thread step = one 16-byte vector of every streamed operand ->
// VE outputs, stored as one VE*sizeof(TO)-byte word (16 B when TO == TI, 4 B for fp32->bool).
#pragma once

#include <string>

#include "bcast_plan.h"
#include "map_functors.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);

template <typename T>
struct MapCompute {
  using type = T;
};
template <>
struct MapCompute<__nv_bfloat16> {
  using type = float;
};
template <>
struct MapCompute<__half> {
  using type = float;
};
template <>
struct MapCompute<int16_t> {
  using type = int;
};
template <>
struct MapCompute<int8_t> {
  using type = int;
};
template <>
struct MapCompute<uint8_t> {
  using type = int;
};
template <>
struct MapCompute<Bool8> {
  using type = int;
};

template <typename TO, typename R>
__device__ __forceinline__ TO map_store_cast(R v) {
  if constexpr (std::is_same<TO, Bool8>::value) {
    Bool8 b;
    b.v = v != R(0) ? 1 : 0;
    return b;
  } else if constexpr (std::is_same<TO, __nv_bfloat16>::value) {
    return __float2bfloat16_rn(static_cast<float>(v));
  } else if constexpr (std::is_same<TO, __half>::value) {
    return __float2half_rn(static_cast<float>(v));
  } else {
    return static_cast<TO>(v);
  }
}

// N-byte store of a packed group of outputs (N = 1, 2, 4, 8 or 16)
template <int N>
__device__ __forceinline__ void store_packed(void* dst, const void* src) {
  if constexpr (N == 16) {
    stg_stream16(reinterpret_cast<V16*>(dst), *reinterpret_cast<const V16*>(src));
  } else if constexpr (N == 8) {
    *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(src);
  } else if constexpr (N == 4) {
    *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(src);
  } else if constexpr (N == 2) {
    *reinterpret_cast<uint16_t*>(dst) = *reinterpret_cast<const uint16_t*>(src);
  } else {
    *reinterpret_cast<uint8_t*>(dst) = *reinterpret_cast<const uint8_t*>(src);
  }
}

template <int Arity>
struct MapRowsParams {
  const void* in[Arity];
  void* out;
  int nouter;
  long long oext[B200PHI_MAX_RANK];
  long long ostride[Arity][B200PHI_MAX_RANK];
  long long inner;
  int istride[Arity];  // 0 or 1
  long long chunk_v;   // input vectors per chunk
  long long chunks;    // chunks per row
};

// 32-bit divide when both operands fit
__device__ __forceinline__ void divmod_ll(long long idx, long long e, long long* q, long long* rem) {
  if ((static_cast<unsigned long long>(idx) | static_cast<unsigned long long>(e)) >> 32) {
    *q = idx / e;
    *rem = idx - *q * e;
  } else {
    const unsigned qi = static_cast<unsigned>(idx) / static_cast<unsigned>(e);
    *q = qi;
    *rem = static_cast<unsigned>(idx) - qi * static_cast<unsigned>(e);
  }
}

// kMask >= 0: bit i <=> operand i is streamed along the row (compile time, so only streamed
// operands get load buffers); -1: decided at run time.
template <typename TI, typename TO, typename F, int Arity, int kMask, int U>
__global__ void __launch_bounds__(kBlock) map_rows_kernel(MapRowsParams<Arity> p, F f) {
  using C = typename MapCompute<TI>::type;
  constexpr int VE = VecOf<TI>::kElems;
  constexpr int kOutBytes = VE * static_cast<int>(sizeof(TO));
  const long long row = blockIdx.x / p.chunks, chunk = blockIdx.x % p.chunks;
  auto streamed = [&](int i) -> bool {
    if constexpr (kMask >= 0) return (kMask >> i) & 1;
    return p.istride[i] != 0;
  };

  const TI* src[Arity];
  {
    long long off[Arity];
#pragma unroll
    for (int i = 0; i < Arity; ++i) off[i] = 0;
    long long idx = row;
    for (int k = p.nouter - 1; k >= 0; --k) {
      long long q, rem;
      divmod_ll(idx, p.oext[k], &q, &rem);
#pragma unroll
      for (int i = 0; i < Arity; ++i) off[i] += rem * p.ostride[i][k];
      idx = q;
    }
#pragma unroll
    for (int i = 0; i < Arity; ++i) src[i] = static_cast<const TI*>(p.in[i]) + off[i];
  }
  TO* __restrict__ dst = static_cast<TO*>(p.out) + row * p.inner;

  C splat[Arity];
  bool aligned = reinterpret_cast<uintptr_t>(dst) % kOutBytes == 0;
#pragma unroll
  for (int i = 0; i < Arity; ++i) {
    splat[i] = C(0);
    if (!streamed(i)) {
      splat[i] = Cvt<C, TI>::to(ldg_scalar(src[i]));
    } else {
      aligned = aligned && (reinterpret_cast<uintptr_t>(src[i]) % 16 == 0);
    }
  }

  const long long nvec_row = p.inner / VE;
  const long long v0 = chunk * p.chunk_v;
  const bool last_chunk = (chunk == p.chunks - 1);

  auto scalar_at = [&](long long i) {
    C a[3] = {C(0), C(0), C(0)};
#pragma unroll
    for (int k = 0; k < Arity; ++k)
      a[k] = streamed(k) ? Cvt<C, TI>::to(ldg_scalar(src[k] + i)) : splat[k];
    dst[i] = map_store_cast<TO>(f(a[0], a[1], a[2]));
  };

  if (aligned) {
    long long v1 = v0 + p.chunk_v;
    if (v1 > nvec_row) v1 = nvec_row;
    for (long long vb = v0 + threadIdx.x; vb < v1; vb += static_cast<long long>(U) * kBlock) {
      V16 raw[Arity][U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long v = vb + static_cast<long long>(u) * kBlock;
        if (v < v1) {
#pragma unroll
          for (int i = 0; i < Arity; ++i)
            if (streamed(i)) raw[i][u] = ldg_stream16(reinterpret_cast<const V16*>(src[i]) + v);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long v = vb + static_cast<long long>(u) * kBlock;
        if (v < v1) {
          typename VecOf<TI>::U w[Arity];
          struct alignas(kOutBytes) {
            TO e[VE];
          } o;
#pragma unroll
          for (int i = 0; i < Arity; ++i)
            if (streamed(i)) w[i].raw = raw[i][u];
#pragma unroll
          for (int e = 0; e < VE; ++e) {
            C a[3] = {C(0), C(0), C(0)};
#pragma unroll
            for (int i = 0; i < Arity; ++i) a[i] = streamed(i) ? Cvt<C, TI>::to(w[i].e[e]) : splat[i];
            o.e[e] = map_store_cast<TO>(f(a[0], a[1], a[2]));
          }
          store_packed<kOutBytes>(reinterpret_cast<unsigned char*>(dst) + v * kOutBytes, &o);
        }
      }
    }
    if (last_chunk) {
      for (long long i = nvec_row * VE + threadIdx.x; i < p.inner; i += kBlock) scalar_at(i);
    }
  } else {
    const long long e0 = v0 * VE;
    long long e1 = last_chunk ? p.inner : (v0 + p.chunk_v) * VE;
    if (e1 > p.inner) e1 = p.inner;
    for (long long i = e0 + threadIdx.x; i < e1; i += kBlock) scalar_at(i);
  }
}

template <int Arity>
struct MapGenericParams {
  const void* in[Arity];
  void* out;
  int ndim;
  long long extent[B200PHI_MAX_RANK];
  long long stride[Arity][B200PHI_MAX_RANK];
  long long numel;
};

template <typename TI, typename TO, typename F, int Arity>
__global__ void __launch_bounds__(kBlock) map_generic_kernel(MapGenericParams<Arity> p, F f) {
  using C = typename MapCompute<TI>::type;
  for (long long o = static_cast<long long>(blockIdx.x) * kBlock + threadIdx.x; o < p.numel;
       o += static_cast<long long>(gridDim.x) * kBlock) {
    long long off[Arity];
#pragma unroll
    for (int i = 0; i < Arity; ++i) off[i] = 0;
    long long idx = o;
    for (int k = p.ndim - 1; k >= 0; --k) {
      long long q, rem;
      divmod_ll(idx, p.extent[k], &q, &rem);
#pragma unroll
      for (int i = 0; i < Arity; ++i) off[i] += rem * p.stride[i][k];
      idx = q;
    }
    C a[3] = {C(0), C(0), C(0)};
#pragma unroll
    for (int i = 0; i < Arity; ++i)
      a[i] = Cvt<C, TI>::to(ldg_scalar(static_cast<const TI*>(p.in[i]) + off[i]));
    static_cast<TO*>(p.out)[o] = map_store_cast<TO>(f(a[0], a[1], a[2]));
  }
}

// ---- launch ---------------------------------------------------------------------------------
template <typename TI, typename TO, typename F, int Arity>
int launch_map(const b200phi_map_desc* d, const BcastPlan& bp, F f, cudaStream_t stream) {
  constexpr int VE = VecOf<TI>::kElems;
  // The row kernel walks a lone operand as a stream (no run-time mask is built for arity 1); a
  // unary op whose operand is broadcast along the innermost dim goes through the generic walk.
  const bool unary_splat = Arity == 1 && bp.stride[0][bp.ndim - 1] == 0 && bp.inner > 1;
  if (bp.use_rows_kernel && !unary_splat) {
    MapRowsParams<Arity> p;
    for (int i = 0; i < Arity; ++i) p.in[i] = d->in[i].data;
    p.out = d->out;
    p.nouter = bp.ndim - 1;
    for (int k = 0; k < bp.ndim - 1; ++k) {
      p.oext[k] = bp.extent[k];
      for (int i = 0; i < Arity; ++i) p.ostride[i][k] = bp.stride[i][k];
    }
    p.inner = bp.inner;
    int mask = 0;
    for (int i = 0; i < Arity; ++i) {
      p.istride[i] = bp.stride[i][bp.ndim - 1] ? 1 : 0;
      mask |= p.istride[i] << i;
    }
    const long long nvec = (bp.inner + VE - 1) / VE;
    const long long max_chunk = static_cast<long long>(kBlock) * 8;  // 8 vectors per thread
    p.chunks = (nvec + max_chunk - 1) / max_chunk;
    if (p.chunks < 1) p.chunks = 1;
    p.chunk_v = (nvec + p.chunks - 1) / p.chunks;
    const long long blocks = bp.rows * p.chunks;
    if (blocks > 0x7fffffffLL) return set_error(B200PHI_UNSUPPORTED, "map: too many rows");
    const unsigned g = static_cast<unsigned>(blocks);
    constexpr int kAll = (1 << Arity) - 1;
    if constexpr (Arity == 1) {  // a lone operand is always walked as a stream
      map_rows_kernel<TI, TO, F, 1, 1, 8><<<g, kBlock, 0, stream>>>(p, f);
    } else {
      if (mask == kAll)  // three streamed operands: 2 vectors each, or the buffers cost two CTAs per SM
        map_rows_kernel<TI, TO, F, Arity, kAll, (Arity == 3 ? 2 : 4)><<<g, kBlock, 0, stream>>>(p, f);
      else if (mask == 1)
        map_rows_kernel<TI, TO, F, Arity, 1, 8><<<g, kBlock, 0, stream>>>(p, f);
      else
        map_rows_kernel<TI, TO, F, Arity, -1, 4><<<g, kBlock, 0, stream>>>(p, f);
    }
  } else {
    MapGenericParams<Arity> p;
    for (int i = 0; i < Arity; ++i) p.in[i] = d->in[i].data;
    p.out = d->out;
    p.ndim = bp.ndim;
    for (int k = 0; k < bp.ndim; ++k) {
      p.extent[k] = bp.extent[k];
      for (int i = 0; i < Arity; ++i) p.stride[i][k] = bp.stride[i][k];
    }
    p.numel = bp.out_numel;
    long long blocks = (bp.out_numel + kBlock - 1) / kBlock;
    if (blocks > 148LL * 8 * 64) blocks = 148LL * 8 * 64;
    map_generic_kernel<TI, TO, F, Arity><<<static_cast<unsigned>(blocks), kBlock, 0, stream>>>(p, f);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

// one op, one input dtype: the output type and operand count follow from the op
template <int kOp, typename TI>
int launch_map_op(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t stream) {
  using TO = typename std::conditional<map_out_is_bool(kOp), Bool8, TI>::type;
  MapF<kOp, TI> f{d->p0, d->p1};
  return launch_map<TI, TO, MapF<kOp, TI>, map_arity(kOp)>(d, bp, f, stream);
}

// per-family dispatchers (one translation unit each)
int map_compare(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);
int map_logical(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);
int map_bitwise(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);
int map_ternary(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);
int map_unary(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);

}  // namespace b200phi
