// Kernels of b200phi_map: the [rows, inner] broadcast walk of elementwise.cu with an output
// type that may be narrower than the input type (compare / logical -> BOOL) and three
// operands.  One thread step = one 16-byte vector of every streamed operand ->
// VE outputs, stored as one VE*sizeof(TO)-byte word (16 B when TO == TI, 4 B for fp32->bool).
#pragma once

#include <algorithm>
#include <string>

#include "bcast_plan.h"
#include "map_functors.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();

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

// Flat walk for outputs whose rows are short (an innermost merged dim of a few to ~1000
// elements: feature maps, hidden sizes, per-channel operands).  The row kernel above spends a
// whole CTA on one row; here the output is one flat array of units -- a unit is a 16-byte
// vector (kVec: rows are whole, 16-byte aligned vectors) or one element (any shape at all: this
// variant is also the fallback a per-element unravel kernel used to be) -- and thread t of the
// grid takes units t, t + 256, ... U at a time, all loads issued before the first use.  Per
// unit: one divide for (row, position in row); an operand that has the output's own shape
// needs nothing more (its offset is the unit index), an operand that is one row for every
// output row needs only the position, any other operand unravels the row over the outer dims.
template <int Arity>
struct MapFlatParams {
  const void* in[Arity];
  void* out;
  int nouter;
  FastDiv oext[B200PHI_MAX_RANK];
  unsigned ostride[Arity][B200PHI_MAX_RANK];
  FastDiv upr;     // units per row
  unsigned total;  // units in the output (< 2^31, as is every element offset)
  int istride[Arity];  // 0 or 1 along the row
  int kind[Arity];     // 0: same shape as the output, 1: one row shared by all rows, 2: general
};

template <typename TI, typename TO, typename F, int Arity, bool kVec, int U>
__global__ void __launch_bounds__(kBlock) map_flat_kernel(MapFlatParams<Arity> p, F f) {
  using C = typename MapCompute<TI>::type;
  constexpr int VE = kVec ? VecOf<TI>::kElems : 1;
  constexpr int kOutBytes = VE * static_cast<int>(sizeof(TO));
  bool need_row = false;
#pragma unroll
  for (int i = 0; i < Arity; ++i) need_row = need_row || p.kind[i] == 2;
  for (unsigned base = blockIdx.x * (kBlock * U) + threadIdx.x; base < p.total;
       base += gridDim.x * (kBlock * U)) {
    unsigned off[U][Arity];  // in elements
    bool live[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      live[u] = id < p.total;
      if (!live[u]) continue;
      const unsigned row = fast_div(id, p.upr);
      const unsigned col = id - row * p.upr.d;
      unsigned gen[Arity];
#pragma unroll
      for (int i = 0; i < Arity; ++i) gen[i] = 0;
      if (need_row) {
        unsigned idx = row;
        for (int k = p.nouter - 1; k >= 0; --k) {
          const unsigned q = fast_div(idx, p.oext[k]);
          const unsigned rem = idx - q * p.oext[k].d;
#pragma unroll
          for (int i = 0; i < Arity; ++i) gen[i] += rem * p.ostride[i][k];
          idx = q;
        }
      }
#pragma unroll
      for (int i = 0; i < Arity; ++i) {
        if (p.kind[i] == 0)
          off[u][i] = id * VE;
        else if (p.kind[i] == 1)
          off[u][i] = col * VE;
        else
          off[u][i] = gen[i] + (p.istride[i] ? col * VE : 0);
      }
    }
    if constexpr (kVec) {
      V16 raw[U][Arity];
      C splat[U][Arity];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!live[u]) continue;
#pragma unroll
        for (int i = 0; i < Arity; ++i) {
          const TI* src = static_cast<const TI*>(p.in[i]) + off[u][i];
          if (p.istride[i])  // only a full-shape operand is read once; the others repeat across rows
            raw[u][i] = p.kind[i] == 0 ? ldg_stream16(src) : ldg_cached16(src);
          else
            splat[u][i] = Cvt<C, TI>::to(ldg_scalar(src));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!live[u]) continue;
        struct alignas(kOutBytes) {
          TO e[VE];
        } o;
        typename VecOf<TI>::U w[Arity];
#pragma unroll
        for (int i = 0; i < Arity; ++i) w[i].raw = raw[u][i];
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          C a[3] = {C(0), C(0), C(0)};
#pragma unroll
          for (int i = 0; i < Arity; ++i) a[i] = p.istride[i] ? Cvt<C, TI>::to(w[i].e[e]) : splat[u][i];
          o.e[e] = map_store_cast<TO>(f(a[0], a[1], a[2]));
        }
        const unsigned id = base + u * kBlock;
        store_packed<kOutBytes>(reinterpret_cast<unsigned char*>(p.out) +
                                    static_cast<size_t>(id) * kOutBytes, &o);
      }
    } else {
      C a[U][3];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        a[u][0] = a[u][1] = a[u][2] = C(0);
        if (!live[u]) continue;
#pragma unroll
        for (int i = 0; i < Arity; ++i)
          a[u][i] = Cvt<C, TI>::to(ldg_scalar(static_cast<const TI*>(p.in[i]) + off[u][i]));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!live[u]) continue;
        const unsigned id = base + u * kBlock;
        static_cast<TO*>(p.out)[id] = map_store_cast<TO>(f(a[u][0], a[u][1], a[u][2]));
      }
    }
  }
}

// Ragged rows (a row is not a whole number of 16-byte vectors): the output and every operand
// with the output's shape are still contiguous aligned arrays and move as 16-byte units; only
// the broadcast operands are gathered per element (cached loads), the operand kind tested once
// per unit.  Needs inner >= VE (a unit touches at most two rows).  MapFlatParams::upr holds the
// ELEMENTS per row here, total the element count.
template <typename TI, typename TO, typename F, int Arity, int U>
__global__ void __launch_bounds__(kBlock) map_ragged_kernel(MapFlatParams<Arity> p, F f) {
  using C = typename MapCompute<TI>::type;
  constexpr int VE = VecOf<TI>::kElems;
  constexpr int kOutBytes = VE * static_cast<int>(sizeof(TO));
  const unsigned inner = p.upr.d;
  const unsigned units = (p.total + VE - 1) / VE;
  bool need_row = false;
#pragma unroll
  for (int i = 0; i < Arity; ++i) need_row = need_row || p.kind[i] == 2;
  for (unsigned base = blockIdx.x * (kBlock * U) + threadIdx.x; base < units;
       base += gridDim.x * (kBlock * U)) {
    V16 raw[U][Arity];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      if (id < units && (id + 1) * VE <= p.total) {
#pragma unroll
        for (int i = 0; i < Arity; ++i)
          if (p.kind[i] == 0) raw[u][i] = ldg_stream16(reinterpret_cast<const V16*>(p.in[i]) + id);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      if (id >= units) continue;
      const bool whole = (id + 1) * VE <= p.total;
      const unsigned e0 = id * VE;
      const unsigned row0 = fast_div(e0, p.upr);
      const unsigned col0 = e0 - row0 * inner;
      unsigned gen[2][Arity];
#pragma unroll
      for (int i = 0; i < Arity; ++i) gen[0][i] = gen[1][i] = 0;
      if (need_row) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          unsigned idx = row0 + h;
          for (int k = p.nouter - 1; k >= 0; --k) {
            const unsigned q = fast_div(idx, p.oext[k]);
            const unsigned rem = idx - q * p.oext[k].d;
#pragma unroll
            for (int i = 0; i < Arity; ++i) gen[h][i] += rem * p.ostride[i][k];
            idx = q;
          }
        }
      }
      typename VecOf<TI>::U w[Arity];
#pragma unroll
      for (int i = 0; i < Arity; ++i) w[i].raw = raw[u][i];
      C vals[Arity][VE];
#pragma unroll
      for (int i = 0; i < Arity; ++i) {
        const TI* src = static_cast<const TI*>(p.in[i]);
        if (p.kind[i] == 0) {
#pragma unroll
          for (int e = 0; e < VE; ++e)
            vals[i][e] = whole ? Cvt<C, TI>::to(w[i].e[e])
                               : (e0 + e < p.total ? Cvt<C, TI>::to(ldg_scalar(src + e0 + e)) : C(0));
        } else if (p.kind[i] == 2 && !p.istride[i]) {
          const C v0 = Cvt<C, TI>::to(ldg_scalar(src + gen[0][i]));
          const C v1 = (col0 + VE > inner && (row0 + 1) * inner < p.total)
                           ? Cvt<C, TI>::to(ldg_scalar(src + gen[1][i]))
                           : v0;
#pragma unroll
          for (int e = 0; e < VE; ++e) vals[i][e] = (col0 + e >= inner) ? v1 : v0;
        } else {
#pragma unroll
          for (int e = 0; e < VE; ++e) {
            unsigned col = col0 + e;
            const int h = col >= inner ? 1 : 0;
            col -= h ? inner : 0;
            const unsigned o_ = (p.kind[i] == 1 ? 0u : gen[h][i]) + (p.istride[i] ? col : 0u);
            vals[i][e] = (whole || e0 + e < p.total) ? Cvt<C, TI>::to(ldg_scalar(src + o_)) : C(0);
          }
        }
      }
      struct alignas(kOutBytes) {
        TO e[VE];
      } o;
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        const C a0 = vals[0][e];
        const C a1 = Arity > 1 ? vals[Arity > 1 ? 1 : 0][e] : C(0);
        const C a2 = Arity > 2 ? vals[Arity > 2 ? 2 : 0][e] : C(0);
        o.e[e] = map_store_cast<TO>(f(a0, a1, a2));
        if (!whole && e0 + e < p.total) static_cast<TO*>(p.out)[e0 + e] = o.e[e];
      }
      if (whole)
        store_packed<kOutBytes>(reinterpret_cast<unsigned char*>(p.out) + static_cast<size_t>(id) * kOutBytes, &o);
    }
  }
}

// The flat cases that matter, as in the elementwise engine (bcast_flat_row_kernel): the first
// operand has the output's shape and the other one or two are ONE ROW shared by every row -- a
// compare against a row vector, max_grad along axis 0 (x against the row of maxima and the row of
// upstream gradients).  No run-time operand-kind tests, no per-element "streamed or splat" selects
// (map_flat_kernel with three operands: 92 registers, 23 % of the warp slots).  kRows: 1 or 2 shared
// rows; kFullIs: which functor argument the full-shape operand is (binary form only).
struct MapFlatRowParams {
  const void* full;
  const void* row[2];
  void* out;
  FastDiv upr;     // vectors per row
  unsigned total;  // vectors in this slice of the output
};
template <typename TI, typename TO, typename F, int kRows, int kFullIs, int U>
__global__ void __launch_bounds__(kBlock, kRows == 2 ? 4 : 5)  // (two rows: 64 registers instead of 70; one row: its own 48)
    map_flat_row_kernel(const __grid_constant__ MapFlatRowParams p, F f) {
  using C = typename MapCompute<TI>::type;
  constexpr int VE = VecOf<TI>::kElems;
  constexpr int kOutBytes = VE * static_cast<int>(sizeof(TO));
  const V16* __restrict__ xf = static_cast<const V16*>(p.full);
  const V16* __restrict__ r0 = static_cast<const V16*>(p.row[0]);
  const V16* __restrict__ r1 = static_cast<const V16*>(p.row[kRows - 1]);
  for (unsigned base = blockIdx.x * (kBlock * U) + threadIdx.x; base < p.total; base += gridDim.x * (kBlock * U)) {
    V16 a[U], b[U], c[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      if (id < p.total) {
        const unsigned col = id - fast_div(id, p.upr) * p.upr.d;
        a[u] = ldg_stream16(xf + id);
        b[u] = ldg_cached16(r0 + col);  // (the rows repeat: L1 serves them)
        if constexpr (kRows == 2) c[u] = ldg_cached16(r1 + col);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      if (id >= p.total) continue;
      typename VecOf<TI>::U wa, wb, wc;
      wa.raw = a[u];
      wb.raw = b[u];
      if constexpr (kRows == 2) wc.raw = c[u];
      struct alignas(kOutBytes) {
        TO e[VE];
      } o;
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        const C xv = Cvt<C, TI>::to(wa.e[e]), rv = Cvt<C, TI>::to(wb.e[e]);
        if constexpr (kRows == 2) {
          o.e[e] = map_store_cast<TO>(f(xv, rv, Cvt<C, TI>::to(wc.e[e])));
        } else {
          o.e[e] = map_store_cast<TO>(kFullIs == 0 ? f(xv, rv, C(0)) : f(rv, xv, C(0)));
        }
      }
      store_packed<kOutBytes>(reinterpret_cast<unsigned char*>(p.out) + static_cast<size_t>(id) * kOutBytes, &o);
    }
  }
}

// ---- launch ---------------------------------------------------------------------------------
// The ragged case that matters, as in the elementwise engine: a full-shape operand against ONE
// VALUE PER ROW ([N, C, 49] vs [C, 1]), binary functor.  No run-time operand-kind test, no run-time
// index into the argument struct, U units per thread with all loads issued first.
struct MapRaggedScalarParams {
  const void* full;
  const void* scalar;
  void* out;
  FastDiv inner;
  unsigned total;
  int nouter;
  FastDiv oext[3];
  unsigned ostride[3];
};

__device__ __forceinline__ unsigned map_ragged_row_offset(unsigned row, const MapRaggedScalarParams& p) {
  unsigned off = 0;
#pragma unroll
  for (int k = 2; k >= 0; --k) {
    if (k >= 3 - p.nouter) {
      const unsigned q = fast_div(row, p.oext[k]);
      off += (row - q * p.oext[k].d) * p.ostride[k];
      row = q;
    }
  }
  return off;
}

template <typename TI, typename TO, typename F, int kScalarIs, int U>
__global__ void __launch_bounds__(kBlock)
    map_ragged_rowscalar_kernel(const __grid_constant__ MapRaggedScalarParams p, F f) {
  using C = typename MapCompute<TI>::type;
  constexpr int VE = VecOf<TI>::kElems;
  constexpr int kOutBytes = VE * static_cast<int>(sizeof(TO));
  const unsigned inner = p.inner.d;
  const unsigned units = p.total / VE;
  const TI* __restrict__ sc = static_cast<const TI*>(p.scalar);
  for (unsigned base = blockIdx.x * (kBlock * U) + threadIdx.x; base < units + 1;
       base += gridDim.x * (kBlock * U)) {
    V16 raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      if (id < units) raw[u] = ldg_stream16(reinterpret_cast<const V16*>(p.full) + id);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      if (id > units) continue;
      const unsigned e0 = id * VE;
      if (e0 >= p.total) continue;
      const unsigned row0 = fast_div(e0, p.inner);
      const unsigned col0 = e0 - row0 * inner;
      const C v0 = Cvt<C, TI>::to(ldg_scalar(sc + map_ragged_row_offset(row0, p)));
      C v1 = v0;
      if (col0 + VE > inner && (row0 + 1) * inner < p.total)
        v1 = Cvt<C, TI>::to(ldg_scalar(sc + map_ragged_row_offset(row0 + 1, p)));
      if (id < units) {
        typename VecOf<TI>::U w;
        w.raw = raw[u];
        struct alignas(kOutBytes) {
          TO e[VE];
        } o;
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const C xv = Cvt<C, TI>::to(w.e[e]);
          const C sv = (col0 + e >= inner) ? v1 : v0;
          o.e[e] = map_store_cast<TO>(kScalarIs == 1 ? f(xv, sv, C(0)) : f(sv, xv, C(0)));
        }
        store_packed<kOutBytes>(static_cast<TO*>(p.out) + static_cast<size_t>(id) * VE, &o);
      } else {
        const TI* __restrict__ fx = static_cast<const TI*>(p.full);
        for (unsigned e = 0; e0 + e < p.total; ++e) {
          const C xv = Cvt<C, TI>::to(ldg_scalar(fx + e0 + e));
          const C sv = (col0 + e >= inner) ? v1 : v0;
          static_cast<TO*>(p.out)[e0 + e] = map_store_cast<TO>(kScalarIs == 1 ? f(xv, sv, C(0)) : f(sv, xv, C(0)));
        }
      }
    }
  }
}

template <typename TI, typename TO, typename F, int Arity>
int launch_map_tile(const b200phi_map_desc* d, const BcastPlan& bp, F f, int mode, cudaStream_t stream) {
  const bool vec = mode == 1;  // mode 0: elements, 1: whole vectors per row, 2: ragged rows
  constexpr int VE = VecOf<TI>::kElems;
  constexpr int U = 4;
  constexpr long long kLimit = 1LL << 31;  // the kernel indexes with 32 bits
  const int no = bp.ndim - 1;              // outer dims
  // the output's own (contiguous) strides over the outer dims, to recognise full-shape operands
  long long full[B200PHI_MAX_RANK];
  long long run = bp.inner;
  for (int k = no - 1; k >= 0; --k) {
    full[k] = run;
    run *= bp.extent[k];
  }
  // slices of the outermost dim, each under 2^31 elements (one slice for all but huge outputs)
  const long long slab = no > 0 ? full[0] : bp.inner;  // elements per index of the outermost dim
  if (slab >= kLimit)
    return set_error(B200PHI_UNSUPPORTED, "map: a short-row output this large is not built");
  const long long outer0 = no > 0 ? bp.extent[0] : 1;
  const long long per_slice = std::max<long long>(1, (kLimit - 1) / slab);
  if (mode == 2 && outer0 > per_slice) mode = 0;  // slices would lose the 16-byte alignment
  for (long long a = 0; a < outer0; a += per_slice) {
    const long long cnt = std::min(per_slice, outer0 - a);
    MapFlatParams<Arity> p;
    p.nouter = no;
    for (int k = 0; k < no; ++k) {
      p.oext[k] = make_fast_div(static_cast<unsigned>(k == 0 ? cnt : bp.extent[k]));
      for (int i = 0; i < Arity; ++i) p.ostride[i][k] = static_cast<unsigned>(bp.stride[i][k]);
    }
    const long long upr = vec ? bp.inner / VE : bp.inner;
    p.upr = make_fast_div(static_cast<unsigned>(upr));
    const long long rows = (no > 0 ? cnt * (bp.rows / outer0) : 1);
    p.total = static_cast<unsigned>(rows * upr);
    for (int i = 0; i < Arity; ++i) {
      const long long shift = no > 0 ? a * bp.stride[i][0] : 0;
      p.in[i] = static_cast<const TI*>(d->in[i].data) + shift;
      p.istride[i] = bp.stride[i][bp.ndim - 1] ? 1 : 0;
      bool same = p.istride[i] == 1, shared = p.istride[i] == 1;
      for (int k = 0; k < no; ++k) {
        same = same && bp.stride[i][k] == full[k];
        shared = shared && bp.stride[i][k] == 0;
      }
      p.kind[i] = same ? 0 : (shared ? 1 : 2);
    }
    p.out = static_cast<TO*>(d->out) + a * slab;
    long long blocks = (static_cast<long long>(p.total) + kBlock * U - 1) / (kBlock * U);
    const long long cap = static_cast<long long>(current_sm_count()) * 8 * 32;  // grid-stride beyond that
    if (blocks > cap) blocks = cap;
    const unsigned g = static_cast<unsigned>(blocks);
    bool rowscalar_done = false;
    if constexpr (Arity == 2) {
      const int fi = p.kind[0] == 0 ? 0 : (p.kind[1] == 0 ? 1 : -1);
      const int si = fi < 0 ? -1 : 1 - fi;
      // (rows that are whole vectors take the same kernel: its units then never straddle two rows)
      if ((mode == 2 || mode == 1) && fi >= 0 && p.kind[si] == 2 && p.istride[si] == 0 && no >= 1 && no <= 3 &&
          reinterpret_cast<uintptr_t>(p.in[fi]) % 16 == 0) {
        MapRaggedScalarParams rp;
        rp.full = p.in[fi];
        rp.scalar = p.in[si];
        rp.out = p.out;
        rp.inner = mode == 2 ? p.upr : make_fast_div(static_cast<unsigned>(bp.inner));  // elements per row
        rp.total = mode == 2 ? p.total : p.total * VE;                                  // elements
        rp.nouter = no;
        for (int k = 0; k < 3; ++k) {
          const int src = no - 3 + k;
          rp.oext[k] = src >= 0 ? p.oext[src] : make_fast_div(1u);
          rp.ostride[k] = src >= 0 ? p.ostride[si][src] : 0u;
        }
        constexpr int RU = 8;
        long long rb = ((static_cast<long long>(rp.total) / VE + 1) + kBlock * RU - 1) / (kBlock * RU);
        if (rb > cap) rb = cap;
        if (si == 1)
          map_ragged_rowscalar_kernel<TI, TO, F, 1, RU><<<static_cast<unsigned>(rb), kBlock, 0, stream>>>(rp, f);
        else
          map_ragged_rowscalar_kernel<TI, TO, F, 0, RU><<<static_cast<unsigned>(rb), kBlock, 0, stream>>>(rp, f);
        rowscalar_done = true;
      }
    }
    static const bool no_flat_row = getenv("B200PHI_NO_FLAT_ROW") != nullptr;
    if constexpr (Arity == 2) {
      const int fi = p.kind[0] == 0 ? 0 : (p.kind[1] == 0 ? 1 : -1);
      const int si = fi < 0 ? -1 : 1 - fi;
      if (!rowscalar_done && vec && fi >= 0 && p.kind[si] == 1 && !no_flat_row) {
        MapFlatRowParams rp;
        rp.full = p.in[fi];
        rp.row[0] = rp.row[1] = p.in[si];
        rp.out = p.out;
        rp.upr = p.upr;
        rp.total = p.total;
        if (fi == 0)
          map_flat_row_kernel<TI, TO, F, 1, 0, U><<<g, kBlock, 0, stream>>>(rp, f);
        else
          map_flat_row_kernel<TI, TO, F, 1, 1, U><<<g, kBlock, 0, stream>>>(rp, f);
        rowscalar_done = true;
      }
    }
    if constexpr (Arity == 3) {
      if (vec && p.kind[0] == 0 && p.kind[1] == 1 && p.kind[2] == 1 && !no_flat_row) {
        MapFlatRowParams rp;
        rp.full = p.in[0];
        rp.row[0] = p.in[1];
        rp.row[1] = p.in[2];
        rp.out = p.out;
        rp.upr = p.upr;
        rp.total = p.total;
        map_flat_row_kernel<TI, TO, F, 2, 0, U><<<g, kBlock, 0, stream>>>(rp, f);
        rowscalar_done = true;
      }
    }
    if (rowscalar_done) {
    } else if (mode == 2) {
      long long rb = ((static_cast<long long>(p.total) + VE - 1) / VE + kBlock * 2 - 1) / (kBlock * 2);
      if (rb > cap) rb = cap;
      map_ragged_kernel<TI, TO, F, Arity, 2><<<static_cast<unsigned>(rb), kBlock, 0, stream>>>(p, f);
    } else if (vec) {
      map_flat_kernel<TI, TO, F, Arity, true, U><<<g, kBlock, 0, stream>>>(p, f);
    } else {
      map_flat_kernel<TI, TO, F, Arity, false, U><<<g, kBlock, 0, stream>>>(p, f);
    }
  }
  return B200PHI_OK;
}

template <typename TI, typename TO, typename F, int Arity>
int launch_map(const b200phi_map_desc* d, const BcastPlan& bp, F f, cudaStream_t stream) {
  constexpr int VE = VecOf<TI>::kElems;
  // The row kernel walks a lone operand as a stream (no run-time mask is built for arity 1); a
  // unary op whose operand is broadcast along the innermost dim goes through the generic walk.
  const bool unary_splat = Arity == 1 && bp.stride[0][bp.ndim - 1] == 0 && bp.inner > 1;
  // Short rows go to the tiled walk: as 16-byte vectors when every row is a whole number of
  // aligned vectors (at most 512 of them: 8 KiB; measured: 1024-vector rows with shared-row operands are faster on the row kernel), as single elements when rows are ragged and short.
  bool whole_vectors = bp.inner % VE == 0 &&
                       reinterpret_cast<uintptr_t>(d->out) % (VE * sizeof(TO)) == 0;
  for (int i = 0; i < Arity; ++i)
    if (bp.stride[i][bp.ndim - 1] != 0)
      whole_vectors = whole_vectors && reinterpret_cast<uintptr_t>(d->in[i].data) % 16 == 0;
  // a full-shape first operand against one or two shared rows has its own flat kernel
  // (map_flat_row_kernel), which also beats the row kernel on longer rows (fp32 4096 columns:
  // less_than 5.6 -> 6.3 TB/s, max_grad along axis 0 5.6 -> 6.3)
  bool flat_row_pattern = Arity >= 2 && bp.ndim >= 2;
  if (flat_row_pattern) {
    int nfull = 0, nshared = 0, full_idx = -1;
    for (int i = 0; i < Arity; ++i) {
      bool same = bp.stride[i][bp.ndim - 1] == 1, shared = same;
      long long r = bp.inner;
      for (int k = bp.ndim - 2; k >= 0; --k) {
        same = same && bp.stride[i][k] == r;
        shared = shared && bp.stride[i][k] == 0;
        r *= bp.extent[k];
      }
      nfull += same ? 1 : 0;
      nshared += shared ? 1 : 0;
      if (same) full_idx = i;
    }
    flat_row_pattern = nfull == 1 && nshared == Arity - 1 && (Arity == 2 || full_idx == 0);
  }
  // ... and so does a full-shape operand against one value per row, up to 1024 vectors per row
  bool row_scalar_pattern = Arity == 2 && bp.ndim >= 2 && bp.ndim <= 4;
  if (row_scalar_pattern) {
    int nfull = 0, nscalar = 0;
    for (int i = 0; i < Arity; ++i) {
      bool same = bp.stride[i][bp.ndim - 1] == 1;
      long long r = bp.inner;
      for (int k = bp.ndim - 2; k >= 0; --k) {
        same = same && bp.stride[i][k] == r;
        r *= bp.extent[k];
      }
      nfull += same ? 1 : 0;
      nscalar += bp.stride[i][bp.ndim - 1] == 0 ? 1 : 0;
    }
    row_scalar_pattern = nfull == 1 && nscalar == 1;
  }
  const long long tile_vec_max = flat_row_pattern ? (1LL << 16) : (row_scalar_pattern ? 1024 : 512);
  const bool tile_vec = whole_vectors && bp.rows > 1 && bp.inner / VE <= tile_vec_max;
  const bool tile_scalar = !tile_vec && bp.rows > 1 && !whole_vectors && bp.inner <= 1024;
  // ragged rows: the output and the full-shape operands as aligned units, the rest gathered
  bool tile_ragged = tile_scalar && Arity > 1 && bp.inner >= VE &&
                     reinterpret_cast<uintptr_t>(d->out) % (VE * sizeof(TO)) == 0;
  if (tile_ragged) {
    bool any_full = false;
    for (int i = 0; i < Arity; ++i) {
      bool same = bp.stride[i][bp.ndim - 1] == 1;
      long long r = bp.inner;
      for (int k = bp.ndim - 2; k >= 0; --k) {
        same = same && bp.stride[i][k] == r;
        r *= bp.extent[k];
      }
      if (same) {
        any_full = true;
        tile_ragged = tile_ragged && reinterpret_cast<uintptr_t>(d->in[i].data) % 16 == 0;
      }
    }
    tile_ragged = tile_ragged && any_full;
  }
  if (bp.use_rows_kernel && !unary_splat && !tile_vec && !tile_scalar) {
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
    const int rc = launch_map_tile<TI, TO, F, Arity>(d, bp, f, tile_vec ? 1 : (tile_ragged ? 2 : 0), stream);
    if (rc != B200PHI_OK) return rc;
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
int map_binary(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);
int map_unary(const b200phi_map_desc* d, const BcastPlan& bp, cudaStream_t s);

}  // namespace b200phi
