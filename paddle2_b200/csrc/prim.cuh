// Device building blocks shared by the reduce / arg / broadcast kernels.
// Plays the role of the reference's KPS layer
// (paddle/phi/kernels/primitive/{compute,datamover,functor}_primitives.h) but is
// written for sm_100a only: 16-byte streaming loads that bypass L1, xor-shuffle
// warp combines, and one shared-memory hop per block combine.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <limits>
#include <type_traits>

namespace b200phi {

constexpr int kBlock = 256;  // threads per CTA for every streaming kernel
constexpr int kWarps = kBlock / 32;
constexpr int kMaxRank = 9;  // common/ddim.h:71

// bool tensors are bytes; a distinct type keeps the `!= 0` rule in the loader.
struct Bool8 {
  uint8_t v;
};

struct alignas(16) V16 {
  uint32_t w[4];
};

// Streaming 16-byte load: read-only path, no L1 allocation (each byte is used once).
__device__ __forceinline__ V16 ldg_stream16(const void* p) {
  V16 r;
  asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
      : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3])
      : "l"(p));
  return r;
}

// Cached 16-byte load: for an operand that many threads re-read (a broadcast row): L1 serves the
// repeats instead of one L2 slice taking every request.
__device__ __forceinline__ V16 ldg_cached16(const void* p) {
  V16 r;
  asm("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];"
      : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3])
      : "l"(p));
  return r;
}

__device__ __forceinline__ void stg_stream16(void* p, const V16& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.w[0]),
               "r"(v.w[1]), "r"(v.w[2]), "r"(v.w[3])
               : "memory");
}

template <typename T>
__device__ __forceinline__ T ldg_scalar(const T* p) {
  return __ldg(p);
}
template <>
__device__ __forceinline__ __nv_bfloat16 ldg_scalar(const __nv_bfloat16* p) {
  unsigned short u = __ldg(reinterpret_cast<const unsigned short*>(p));
  return __ushort_as_bfloat16(u);
}
template <>
__device__ __forceinline__ __half ldg_scalar(const __half* p) {
  unsigned short u = __ldg(reinterpret_cast<const unsigned short*>(p));
  return __ushort_as_half(u);
}
template <>
__device__ __forceinline__ __nv_fp8_e4m3 ldg_scalar(const __nv_fp8_e4m3* p) {
  __nv_fp8_e4m3 v;
  v.__x = __ldg(reinterpret_cast<const unsigned char*>(p));
  return v;
}
template <>
__device__ __forceinline__ __nv_fp8_e5m2 ldg_scalar(const __nv_fp8_e5m2* p) {
  __nv_fp8_e5m2 v;
  v.__x = __ldg(reinterpret_cast<const unsigned char*>(p));
  return v;
}
template <>
__device__ __forceinline__ Bool8 ldg_scalar(const Bool8* p) {
  Bool8 b;
  b.v = __ldg(reinterpret_cast<const uint8_t*>(p));
  return b;
}

// ---- element conversion: storage type -> accumulation type ------------------
// Mirrors "cast x to out_dtype, then accumulate in MPType"
// (gpu/reduce.h:48-69, amp_type_traits.h:25-53) without the extra pass.
template <typename Acc, typename T>
struct Cvt {
  static __device__ __forceinline__ Acc to(T v) { return static_cast<Acc>(v); }
};
template <typename Acc>
struct Cvt<Acc, __nv_bfloat16> {
  static __device__ __forceinline__ Acc to(__nv_bfloat16 v) {
    return static_cast<Acc>(__bfloat162float(v));
  }
};
template <typename Acc>
struct Cvt<Acc, __half> {
  static __device__ __forceinline__ Acc to(__half v) { return static_cast<Acc>(__half2float(v)); }
};
template <typename Acc>
struct Cvt<Acc, __nv_fp8_e4m3> {
  static __device__ __forceinline__ Acc to(__nv_fp8_e4m3 v) { return static_cast<Acc>(static_cast<float>(v)); }
};
template <typename Acc>
struct Cvt<Acc, __nv_fp8_e5m2> {
  static __device__ __forceinline__ Acc to(__nv_fp8_e5m2 v) { return static_cast<Acc>(static_cast<float>(v)); }
};
template <typename Acc>
struct Cvt<Acc, Bool8> {
  static __device__ __forceinline__ Acc to(Bool8 v) { return static_cast<Acc>(v.v != 0 ? 1 : 0); }
};

// Truthiness used by any/all: the reference casts to bool first (cpu/reduce.h:90-98).
template <typename T>
__device__ __forceinline__ int truthy(T v) {
  return v != static_cast<T>(0) ? 1 : 0;
}
template <>
__device__ __forceinline__ int truthy(__nv_bfloat16 v) {
  return __bfloat162float(v) != 0.f ? 1 : 0;
}
template <>
__device__ __forceinline__ int truthy(__half v) {
  return __half2float(v) != 0.f ? 1 : 0;
}
template <>
__device__ __forceinline__ int truthy(Bool8 v) {
  return v.v != 0 ? 1 : 0;
}

// ---- reduce operators (functor_primitives.h:140-209) -------------------------
template <typename Acc>
struct SumOp {
  static constexpr bool kLogical = false;
  static __device__ __forceinline__ Acc identity() { return Acc(0); }
  static __device__ __forceinline__ Acc combine(Acc a, Acc b) { return a + b; }
};
template <typename Acc>
struct ProdOp {
  static constexpr bool kLogical = false;
  static __device__ __forceinline__ Acc identity() { return Acc(1); }
  static __device__ __forceinline__ Acc combine(Acc a, Acc b) { return a * b; }
};
template <typename Acc>
struct Limits {
  static __device__ __forceinline__ Acc lowest() { return std::numeric_limits<Acc>::lowest(); }
  static __device__ __forceinline__ Acc highest() { return std::numeric_limits<Acc>::max(); }
};
template <>
struct Limits<float> {
  static __device__ __forceinline__ float lowest() { return -3.402823466e+38f; }
  static __device__ __forceinline__ float highest() { return 3.402823466e+38f; }
};
template <>
struct Limits<double> {
  static __device__ __forceinline__ double lowest() { return -1.7976931348623157e+308; }
  static __device__ __forceinline__ double highest() { return 1.7976931348623157e+308; }
};
template <>
struct Limits<long long> {
  static __device__ __forceinline__ long long lowest() { return (-9223372036854775807LL - 1); }
  static __device__ __forceinline__ long long highest() { return 9223372036854775807LL; }
};
template <typename Acc>
struct MaxOp {  // (b > a) ? b : a, seed lowest()
  static constexpr bool kLogical = false;
  static __device__ __forceinline__ Acc identity() { return Limits<Acc>::lowest(); }
  static __device__ __forceinline__ Acc combine(Acc a, Acc b) { return (b > a) ? b : a; }
};
template <typename Acc>
struct MinOp {  // (b < a) ? b : a, seed max()
  static constexpr bool kLogical = false;
  static __device__ __forceinline__ Acc identity() { return Limits<Acc>::highest(); }
  static __device__ __forceinline__ Acc combine(Acc a, Acc b) { return (b < a) ? b : a; }
};
template <typename Acc>
struct AnyOp {
  static constexpr bool kLogical = true;
  static __device__ __forceinline__ Acc identity() { return Acc(0); }
  static __device__ __forceinline__ Acc combine(Acc a, Acc b) { return a | b; }
};
template <typename Acc>
struct AllOp {
  static constexpr bool kLogical = true;
  static __device__ __forceinline__ Acc identity() { return Acc(1); }
  static __device__ __forceinline__ Acc combine(Acc a, Acc b) { return a & b; }
};

// Reduce operators that transform each element first -- the TransformOp slot of
// funcs::ReduceKernel<Tx, Ty, ReduceOp, TransformOp> (reduce_function.h:1047), as p_norm and
// squared_l2_norm fill it (gpu/p_norm_kernel.cu:26-52,84-95, gpu/squared_l2_norm_kernel.cu:31).
// The combine stays the base operator's; only the element load changes.
enum { kPreNone = 0, kPreAbs = 1, kPreSquare = 2, kPreNonzero = 3 };
template <typename Op>
struct OpPre {
  static constexpr int value = kPreNone;
};
template <typename Acc>
struct SumAbsOp : SumOp<Acc> {};
template <typename Acc>
struct SumSquareOp : SumOp<Acc> {};
template <typename Acc>
struct L2NormOp : SumOp<Acc> {};  // sum of squares, sqrt at the final store
template <typename Acc>
struct CountNonzeroOp : SumOp<Acc> {};
template <typename Acc>
struct MaxAbsOp : MaxOp<Acc> {};
template <typename Acc>
struct MinAbsOp : MinOp<Acc> {};
template <typename Acc>
struct OpPre<SumAbsOp<Acc>> {
  static constexpr int value = kPreAbs;
};
template <typename Acc>
struct OpPre<SumSquareOp<Acc>> {
  static constexpr int value = kPreSquare;
};
template <typename Acc>
struct OpPre<L2NormOp<Acc>> {
  static constexpr int value = kPreSquare;
};
template <typename Acc>
struct OpPre<CountNonzeroOp<Acc>> {
  static constexpr int value = kPreNonzero;
};
template <typename Acc>
struct OpPre<MaxAbsOp<Acc>> {
  static constexpr int value = kPreAbs;
};
template <typename Acc>
struct OpPre<MinAbsOp<Acc>> {
  static constexpr int value = kPreAbs;
};

// Load one element as the accumulation type, honouring the any/all truthiness rule and the
// operator's element transform.
// The operator's element transform on a value that is already of the accumulation type.
template <typename Acc, typename Op>
__device__ __forceinline__ Acc pre_as(Acc a) {
  if constexpr (OpPre<Op>::value == kPreAbs) {
    return a < Acc(0) ? -a : a;
  } else if constexpr (OpPre<Op>::value == kPreSquare) {
    return a * a;
  } else if constexpr (OpPre<Op>::value == kPreNonzero) {
    return a != Acc(0) ? Acc(1) : Acc(0);
  } else {
    return a;
  }
}
template <typename Acc, typename Op, typename T>
__device__ __forceinline__ Acc load_as(T v) {
  if constexpr (Op::kLogical) {
    return static_cast<Acc>(truthy(v));
  } else {
    return pre_as<Acc, Op>(Cvt<Acc, T>::to(v));
  }
}

// +1 for the max family, -1 for the min family, 0 otherwise
template <typename Op>
struct OpMinMax {
  static constexpr int value = 0;
};
template <typename Acc>
struct OpMinMax<MaxOp<Acc>> {
  static constexpr int value = 1;
};
template <typename Acc>
struct OpMinMax<MaxAbsOp<Acc>> {
  static constexpr int value = 1;
};
template <typename Acc>
struct OpMinMax<MinOp<Acc>> {
  static constexpr int value = -1;
};
template <typename Acc>
struct OpMinMax<MinAbsOp<Acc>> {
  static constexpr int value = -1;
};

// ---- 16-byte vector of T ------------------------------------------------------
template <typename T>
struct VecOf {
  static constexpr int kElems = 16 / sizeof(T);
  union U {
    V16 raw;
    T e[kElems];
    __device__ U() {}
  };
};

// Unpack a 16-byte vector into VE values of the accumulation type.  bf16 pairs are widened
// with one ALU op per element (hi = w & 0xffff0000, lo = w << 16): conversion cost is what
// makes 16-bit row statistics instruction-bound.
template <typename T, typename Acc>
__device__ __forceinline__ void unpack_vec(const V16& raw, Acc* dst) {
  typename VecOf<T>::U u;
  u.raw = raw;
#pragma unroll
  for (int e = 0; e < VecOf<T>::kElems; ++e) dst[e] = Cvt<Acc, T>::to(u.e[e]);
}
template <>
__device__ __forceinline__ void unpack_vec<__nv_bfloat16, float>(const V16& raw, float* dst) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    dst[2 * k] = __uint_as_float(raw.w[k] << 16);
    dst[2 * k + 1] = __uint_as_float(raw.w[k] & 0xffff0000u);
  }
}

// Fold a 16-byte vector into kLanes independent accumulators (shorter add chains).
// 16-bit floats take two shortcuts, because their kernels are bound by instruction issue before
// they are bound by memory (ncu: 77-87 % issue-slot use on short bf16 rows):
//   * max / min (and their |x| forms): the vector's eight values are reduced as packed pairs
//     (three HMNMX2), only the surviving pair is widened -- 7 instructions instead of ~28.  The
//     result is the same value: widening is exact and max / min pick one of the inputs.
//   * bf16 otherwise: a pair is widened with one ALU op per element (w << 16, w & 0xffff0000).
template <typename Acc, typename Op, typename T, int kLanes>
__device__ __forceinline__ void fold_vec(const V16& raw, Acc (&acc)[kLanes]) {
  constexpr bool k16 = std::is_same<T, __nv_bfloat16>::value || std::is_same<T, __half>::value;
  if constexpr (k16 && std::is_same<Acc, float>::value && OpMinMax<Op>::value != 0 &&
                (OpPre<Op>::value == kPreNone || OpPre<Op>::value == kPreAbs)) {
    using P = typename std::conditional<std::is_same<T, __half>::value, __half2, __nv_bfloat162>::type;
    P p[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint32_t w = raw.w[k];
      if constexpr (OpPre<Op>::value == kPreAbs) w &= 0x7fff7fffu;
      p[k] = *reinterpret_cast<const P*>(&w);
    }
    P m;
    if constexpr (OpMinMax<Op>::value > 0) {
      m = __hmax2(__hmax2(p[0], p[1]), __hmax2(p[2], p[3]));
    } else {
      m = __hmin2(__hmin2(p[0], p[1]), __hmin2(p[2], p[3]));
    }
    float lo, hi;
    if constexpr (std::is_same<T, __half>::value) {
      const float2 f = __half22float2(m);
      lo = f.x, hi = f.y;
    } else {
      const uint32_t w = *reinterpret_cast<const uint32_t*>(&m);
      lo = __uint_as_float(w << 16), hi = __uint_as_float(w & 0xffff0000u);
    }
    acc[0] = Op::combine(acc[0], lo);
    acc[1 % kLanes] = Op::combine(acc[1 % kLanes], hi);
  } else if constexpr (std::is_same<T, __nv_bfloat16>::value && std::is_same<Acc, float>::value &&
                       !Op::kLogical) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float lo = __uint_as_float(raw.w[k] << 16), hi = __uint_as_float(raw.w[k] & 0xffff0000u);
      acc[(2 * k) % kLanes] = Op::combine(acc[(2 * k) % kLanes], pre_as<Acc, Op>(lo));
      acc[(2 * k + 1) % kLanes] = Op::combine(acc[(2 * k + 1) % kLanes], pre_as<Acc, Op>(hi));
    }
  } else {
    typename VecOf<T>::U u;
    u.raw = raw;
#pragma unroll
    for (int e = 0; e < VecOf<T>::kElems; ++e) {
      acc[e % kLanes] = Op::combine(acc[e % kLanes], load_as<Acc, Op>(u.e[e]));
    }
  }
}

// ---- warp / block combines ------------------------------------------------------
template <typename Acc>
__device__ __forceinline__ Acc shfl_xor(Acc v, int m) {
  return __shfl_xor_sync(0xffffffffu, v, m);
}

// Walk elements [i0, i1) of a row that lies in shared memory, in ascending order, handing
// f(index, value widened to float, slot) -- slot is a compile-time 0..7 inside the unrolled rounds
// (for independent accumulators), 0 elsewhere.  16-bit floats are read as aligned 32-bit words, four words
// per round (a 2-byte load moves 64 bytes per warp instruction, half of what the shared-memory
// pipe can deliver: the thread-per-row kernels were bound by exactly that), with one element
// peeled at either end when the range does not start / end on a word.
template <typename T>
struct Is16Float {
  static constexpr bool value = std::is_same<T, __nv_bfloat16>::value || std::is_same<T, __half>::value;
};
template <typename T>
__device__ __forceinline__ void widen_pair(uint32_t w, float& lo, float& hi) {
  if constexpr (std::is_same<T, __half>::value) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
    lo = f.x, hi = f.y;
  } else {
    lo = __uint_as_float(w << 16), hi = __uint_as_float(w & 0xffff0000u);
  }
}
template <typename T, typename F>
__device__ __forceinline__ void smem_walk16(const T* __restrict__ sp, int i0, int i1, F f) {
  static_assert(Is16Float<T>::value, "16-bit floats");
  int i = i0;
  if (i < i1 && (reinterpret_cast<uintptr_t>(sp + i) & 2)) {
    f(i, Cvt<float, T>::to(sp[i]), 0);
    ++i;
  }
  const uint32_t* __restrict__ w = reinterpret_cast<const uint32_t*>(sp + i);
  const int nw = (i1 - i) >> 1;
  int k = 0;
  for (; k + 4 <= nw; k += 4) {
    uint32_t r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) r[j] = w[k + j];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float lo, hi;
      widen_pair<T>(r[j], lo, hi);
      f(i + 2 * (k + j), lo, 2 * j);
      f(i + 2 * (k + j) + 1, hi, 2 * j + 1);
    }
  }
  for (; k < nw; ++k) {
    float lo, hi;
    widen_pair<T>(w[k], lo, hi);
    f(i + 2 * k, lo, 0);
    f(i + 2 * k + 1, hi, 1);
  }
  if ((i1 - i) & 1) f(i1 - 1, Cvt<float, T>::to(sp[i1 - 1]), 0);
}

// The same walk for any element type: 16-bit floats as above (value handed as float), every other
// type element by element in rounds of eight loads (value handed as T).
template <typename T, typename F>
__device__ __forceinline__ void smem_walk_range(const T* __restrict__ sp, int i0, int i1, F f) {
  if constexpr (Is16Float<T>::value) {
    smem_walk16<T>(sp, i0, i1, f);
  } else {
    int i = i0;
    for (; i + 8 <= i1; i += 8) {
      T raw[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) raw[j] = sp[i + j];
#pragma unroll
      for (int j = 0; j < 8; ++j) f(i + j, raw[j], j);
    }
    for (; i < i1; ++i) f(i, sp[i], 0);
  }
}
// Visit a whole row of n elements that is whole, 16-byte aligned vectors, starting `rot` VECTORS
// into it and wrapping around: such rows have a pitch of nv * 16 bytes, so threads that all start at
// vector 0 share 8 / gcd(nv, 8) of the 8 bank groups a 16-byte shared load is served from; started
// at staggered vectors (rows_staged.cuh) the 8 threads of a load phase hit 8 different groups.  One
// LDS.128 per vector and one wrap test per vector; every thread runs the same loop, so threads
// with different starts do not diverge.  rot < 0: the plain ascending walk of [i0, i1).
template <typename T, typename F>
__device__ __forceinline__ void smem_visit(const T* __restrict__ sp, int n, int i0, int i1, int rot, F f) {
  if (rot < 0) {
    smem_walk_range<T>(sp, i0, i1, f);
    return;
  }
  constexpr int VE = 16 / static_cast<int>(sizeof(T));
  const V16* __restrict__ vp = reinterpret_cast<const V16*>(sp);
  const int nv = n / VE;
  int v = rot;
  for (int k = 0; k < nv; ++k) {
    const V16 raw = vp[v];
    const int e0 = v * VE;
    if constexpr (Is16Float<T>::value) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float lo, hi;
        widen_pair<T>(raw.w[j], lo, hi);
        f(e0 + 2 * j, lo, 2 * j);
        f(e0 + 2 * j + 1, hi, 2 * j + 1);
      }
    } else {
      typename VecOf<T>::U u;
      u.raw = raw;
#pragma unroll
      for (int j = 0; j < VE; ++j) f(e0 + j, u.e[j], j);
    }
    ++v;
    if (v >= nv) v = 0;
  }
}

// All-reduce of v over the L consecutive lanes that share a row.
template <int L, typename V, typename F>
__device__ __forceinline__ V group_all(V v, F combine) {
#pragma unroll
  for (int m = L / 2; m > 0; m >>= 1) v = combine(v, __shfl_xor_sync(0xffffffffu, v, m));
  return v;
}

template <typename Acc, typename Op>
__device__ __forceinline__ Acc warp_combine(Acc v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v = Op::combine(v, shfl_xor(v, m));
  return v;
}

// All kBlock threads call; result valid in every thread. `smem` holds kWarps Acc.
template <typename Acc, typename Op>
__device__ __forceinline__ Acc block_combine(Acc v, Acc* smem) {
  v = warp_combine<Acc, Op>(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();  // smem may still be read by the previous round
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  Acc r = smem[0];
#pragma unroll
  for (int w = 1; w < kWarps; ++w) r = Op::combine(r, smem[w]);
  return r;
}

// Barrier over the G threads of one group.  G == kBlock is the plain CTA barrier (immediate
// id 0); smaller groups use named barrier 1 + group (a register id costs a few us per 100 us
// of short-lived CTAs, so the whole-CTA case must not go through it).
template <int G>
__device__ __forceinline__ void group_barrier() {
  if constexpr (G == kBlock) {
    __syncthreads();
  } else {
    // register barrier id; a switch over immediate ids measured slower (4.86 vs 5.18 TB/s on C4)
    asm volatile("bar.sync %0, %1;" ::"r"(1 + static_cast<int>(threadIdx.x) / G), "n"(G) : "memory");
  }
}

// Combine across a group of G threads (G = 32, 64, 128 or 256; groups are G-aligned slices of
// the CTA).  Warps of one group meet on a named barrier (id 1 + group), so groups never wait
// for each other.  `smem` holds kWarps Acc.  Result valid in every thread of the group.
template <typename Acc, typename Op, int G>
__device__ __forceinline__ Acc group_combine(Acc v, Acc* smem) {
  v = warp_combine<Acc, Op>(v);
  if constexpr (G > 32) {
    constexpr int kW = G / 32;
    const int warp = threadIdx.x >> 5, grp = threadIdx.x / G;
    if ((threadIdx.x & 31) == 0) smem[warp] = v;
    group_barrier<G>();
    Acc r = smem[grp * kW];
#pragma unroll
    for (int w = 1; w < kW; ++w) r = Op::combine(r, smem[grp * kW + w]);
    group_barrier<G>();  // smem reusable
    v = r;
  }
  return v;
}

// Same, but only the group's thread 0 needs the result, and `smem` holds 2*kWarps Acc used
// alternately (`parity` flips per call) so ONE named barrier per call is enough.
template <typename Acc, typename Op, int G>
__device__ __forceinline__ Acc group_combine_leader(Acc v, Acc* smem, int parity) {
  v = warp_combine<Acc, Op>(v);
  if constexpr (G > 32) {
    constexpr int kW = G / 32;
    const int warp = threadIdx.x >> 5, grp = threadIdx.x / G;
    Acc* buf = smem + parity * kWarps;
    if ((threadIdx.x & 31) == 0) buf[warp] = v;
    group_barrier<G>();
    if (threadIdx.x % G == 0) {
      Acc r = buf[grp * kW];
#pragma unroll
      for (int w = 1; w < kW; ++w) r = Op::combine(r, buf[grp * kW + w]);
      v = r;
    }
  }
  return v;
}

// ---- typed result store -----------------------------------------------------------
struct Epilogue {
  void* out;
  int out_dtype;    // b200phi_dtype
  int is_mean;      // divide by count before the store (reduce_function.h:686-688)
  long long count;  // reduce_num, also pre-converted on the host: an I2F.S64 + RCP chain in
  float count_f;    // every CTA's epilogue cost ~10% on config C1
  double count_d;
};

inline Epilogue make_epilogue(void* out, int out_dtype, bool is_mean, long long count) {
  Epilogue ep;
  ep.out = out;
  ep.out_dtype = out_dtype;
  ep.is_mean = is_mean ? 1 : 0;
  ep.count = count;
  ep.count_f = static_cast<float>(count);
  ep.count_d = static_cast<double>(count);
  return ep;
}

template <typename Acc>
struct EpiTraits;
template <>
struct EpiTraits<float> {
  static constexpr int kNative = 10;
  static __device__ __forceinline__ float count(const Epilogue& ep) { return ep.count_f; }
};
template <>
struct EpiTraits<double> {
  static constexpr int kNative = 11;
  static __device__ __forceinline__ double count(const Epilogue& ep) { return ep.count_d; }
};
template <>
struct EpiTraits<long long> {
  static constexpr int kNative = 9;
  static __device__ __forceinline__ long long count(const Epilogue& ep) { return ep.count; }
};
template <>
struct EpiTraits<int> {
  static constexpr int kNative = 7;
  static __device__ __forceinline__ int count(const Epilogue& ep) { return static_cast<int>(ep.count); }
};

// Every other output dtype: out of line, so the hot path stays a compare and a store
// (inlined, the compiler folded it into one jump table that every CTA's leader walked).
template <typename Acc>
__device__ __noinline__ void store_result_slow(void* out, int out_dtype, long long idx, Acc v) {
  switch (out_dtype) {
    case 1:  // BOOL
      reinterpret_cast<uint8_t*>(out)[idx] = (v != Acc(0)) ? 1 : 0;
      break;
    case 2:
      reinterpret_cast<uint8_t*>(out)[idx] = static_cast<uint8_t>(v);
      break;
    case 3:
      reinterpret_cast<int8_t*>(out)[idx] = static_cast<int8_t>(v);
      break;
    case 5:
      reinterpret_cast<int16_t*>(out)[idx] = static_cast<int16_t>(v);
      break;
    case 7:
      reinterpret_cast<int32_t*>(out)[idx] = static_cast<int32_t>(v);
      break;
    case 9:
      reinterpret_cast<long long*>(out)[idx] = static_cast<long long>(v);
      break;
    case 10:
      reinterpret_cast<float*>(out)[idx] = static_cast<float>(v);
      break;
    case 11:
      reinterpret_cast<double*>(out)[idx] = static_cast<double>(v);
      break;
    case 15:
      reinterpret_cast<__half*>(out)[idx] = __float2half_rn(static_cast<float>(v));
      break;
    case 16:
      reinterpret_cast<__nv_bfloat16*>(out)[idx] = __float2bfloat16_rn(static_cast<float>(v));
      break;
    case 17:  // FLOAT8_E4M3FN: exact, the value came from the same format
      reinterpret_cast<__nv_fp8_e4m3*>(out)[idx] = __nv_fp8_e4m3(static_cast<float>(v));
      break;
    case 18:
      reinterpret_cast<__nv_fp8_e5m2*>(out)[idx] = __nv_fp8_e5m2(static_cast<float>(v));
      break;
    default:
      break;
  }
}

template <typename Acc>
__device__ __forceinline__ void store_result(const Epilogue& ep, long long idx, Acc v) {
  if (ep.is_mean) v = v / EpiTraits<Acc>::count(ep);
  if (ep.out_dtype == EpiTraits<Acc>::kNative) {
    reinterpret_cast<Acc*>(ep.out)[idx] = v;
    return;
  }
  store_result_slow<Acc>(ep.out, ep.out_dtype, idx, v);
}

// The store every reduce kernel ends with: the operator's own finishing step, if it has one
// (sqrt for the l2 norm: the second functor of gpu/p_norm_kernel.cu:90-95), then the typed store.
template <typename Op>
struct OpPost {
  static constexpr bool kSqrt = false;
};
template <typename Acc>
struct OpPost<L2NormOp<Acc>> {
  static constexpr bool kSqrt = true;
};
template <typename Op, typename Acc>
__device__ __forceinline__ void store_final(const Epilogue& ep, long long idx, Acc v) {
  if constexpr (OpPost<Op>::kSqrt) v = sqrt(v);
  store_result(ep, idx, v);
}

// The same store for kernels that only run on 16-bit inputs: the usual output there is the input
// type, which store_result would reach through its out-of-line switch once per row.
template <typename Op, typename T, typename Acc>
__device__ __forceinline__ void store_final_16(const Epilogue& ep, long long idx, Acc v) {
  if constexpr (std::is_same<Acc, float>::value &&
                (std::is_same<T, __nv_bfloat16>::value || std::is_same<T, __half>::value)) {
    constexpr int kId = std::is_same<T, __half>::value ? 15 : 16;
    if (ep.out_dtype == kId) {
      if constexpr (OpPost<Op>::kSqrt) v = sqrt(v);
      if (ep.is_mean) v = v / ep.count_f;
      if constexpr (std::is_same<T, __half>::value) {
        reinterpret_cast<__half*>(ep.out)[idx] = __float2half_rn(v);
      } else {
        reinterpret_cast<__nv_bfloat16*>(ep.out)[idx] = __float2bfloat16_rn(v);
      }
      return;
    }
  }
  store_final<Op>(ep, idx, v);
}

// Element k of the NE a lane holds of a short row of any alignment (scalar sub-warp kernels):
// which element index it is, and the loads.  4- and 8-byte types: lane g holds elements
// g + k*G.  2-byte types: 4-byte words from the row's aligned-down base instead of 2-byte loads
// (half the load instructions, 128 bytes per warp instruction): word j = g + (k/2)*G holds
// elements 2j - o and 2j + 1 - o, o = 1 when the row starts on an odd element.
// (pairs only for the 32-element-per-lane form: on 101-column rows the extra index arithmetic cost
// more than the halved loads saved, 1.97 vs 2.36 TB/s; on 1001 columns pairs give 1.89 vs 0.99)
template <typename T, int NE>
constexpr bool short_row_pairs() {
  return sizeof(T) == 2 && NE >= 32;
}
template <typename T, int G, int NE>
__device__ __forceinline__ int short_row_index(int k, int g, int odd) {
  if constexpr (short_row_pairs<T, NE>()) {
    return 2 * (g + (k >> 1) * G) - odd + (k & 1);
  } else {
    return g + k * G;
  }
}
template <typename T, int G, int NE>
__device__ __forceinline__ void short_row_load(const T* __restrict__ r, int cols, int g, bool live,
                                               T (&raw)[NE], int* odd_out) {
  if constexpr (short_row_pairs<T, NE>()) {
    static_assert(NE % 2 == 0, "pairs");
    const int odd = static_cast<int>((reinterpret_cast<uintptr_t>(r) >> 1) & 1);
    *odd_out = odd;
    const uint32_t* __restrict__ wbase = reinterpret_cast<const uint32_t*>(r - odd);
#pragma unroll
    for (int j = 0; j < NE / 2; ++j) {
      const int w = g + j * G;
      const int i0 = 2 * w - odd;
      uint32_t bits = 0;
      if (live && i0 + 1 >= 0 && i0 < cols) {
        // a word that straddles the row's first or last element is read as that one element: the
        // other half belongs to the neighbouring row, or lies outside the tensor at its two ends
        const unsigned short* __restrict__ h = reinterpret_cast<const unsigned short*>(r);
        if (i0 < 0)
          bits = static_cast<uint32_t>(__ldg(h)) << 16;
        else if (i0 + 1 >= cols)
          bits = __ldg(h + i0);
        else
          bits = __ldg(wbase + w);
      }
      const T* pe = reinterpret_cast<const T*>(&bits);
      raw[2 * j] = pe[0];
      raw[2 * j + 1] = pe[1];
    }
  } else {
    *odd_out = 0;
#pragma unroll
    for (int k = 0; k < NE; ++k) {
      const int i = g + k * G;
      if (live && i < cols) raw[k] = ldg_scalar(r + i);
    }
  }
}

// Division by a run-time constant for 31-bit numerators (Granlund-Montgomery): with
// l = ceil(log2 d) and m = floor(2^32 (2^l - d) / d) + 1,  n / d = (umulhi(m, n) + n) >> l.
struct FastDiv {
  unsigned d, m, l;
};
inline FastDiv make_fast_div(unsigned d) {
  FastDiv f;
  f.d = d;
  unsigned l = 0;
  while ((1ull << l) < d) ++l;
  f.l = l;
  f.m = static_cast<unsigned>(((1ull << 32) * ((1ull << l) - d)) / d + 1);
  return f;
}
__device__ __forceinline__ unsigned fast_div(unsigned n, const FastDiv& f) {
  return (__umulhi(f.m, n) + n) >> f.l;
}

}  // namespace b200phi
