// Device pieces shared by the argmax / argmin kernels (arg.cu, arg_split.cu): the (value, index)
// pair, the combine rule "better value, else lower index" (cub::ArgMax as used by
// gpu/arg_min_max_kernel.cu:58-83; Eigen's strict > over an in-order scan on the CPU,
// cpu/arg_min_max_kernel.cc:46-59), and the vector scan of a contiguous span.
#pragma once

#include "../../include/b200phi.h"
#include "prim.cuh"

namespace b200phi {

template <typename V>
struct ArgPair {
  V v;
  long long i;
};

// kMax: b replaces a when b.v > a.v, or equal and b.i < a.i.
template <typename V, bool kMax>
struct ArgOp {
  static __device__ __forceinline__ ArgPair<V> identity() {
    ArgPair<V> p;
    p.v = kMax ? Limits<V>::lowest() : Limits<V>::highest();
    p.i = 0;  // an all-identity slice answers 0, like the CPU kernel
    return p;
  }
  static __device__ __forceinline__ bool better(V b, V a) { return kMax ? (b > a) : (b < a); }
  static __device__ __forceinline__ ArgPair<V> combine(ArgPair<V> a, ArgPair<V> b) {
    const bool take = better(b.v, a.v) || (b.v == a.v && b.i < a.i);
    return take ? b : a;
  }
  // in-order visit by one thread: strict comparison keeps the first occurrence
  static __device__ __forceinline__ void visit(ArgPair<V>& a, V v, long long i) {
    if (better(v, a.v)) {
      a.v = v;
      a.i = i;
    }
  }
};

template <typename V>
__device__ __forceinline__ ArgPair<V> shfl_pair(ArgPair<V> p, int m) {
  ArgPair<V> r;
  r.v = __shfl_xor_sync(0xffffffffu, p.v, m);
  r.i = __shfl_xor_sync(0xffffffffu, p.i, m);
  return r;
}

template <typename V, typename Op>
__device__ __forceinline__ ArgPair<V> warp_arg(ArgPair<V> p) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) p = Op::combine(p, shfl_pair(p, m));
  return p;
}

template <typename T>
struct StoreVal {
  template <typename V>
  static __device__ __forceinline__ void put(void* p, long long slot, V v) {
    reinterpret_cast<T*>(p)[slot] = static_cast<T>(v);
  }
};
template <>
struct StoreVal<__nv_bfloat16> {
  static __device__ __forceinline__ void put(void* p, long long slot, float v) {
    reinterpret_cast<__nv_bfloat16*>(p)[slot] = __float2bfloat16_rn(v);  // exact: v came from bf16
  }
};
template <>
struct StoreVal<__half> {
  static __device__ __forceinline__ void put(void* p, long long slot, float v) {
    reinterpret_cast<__half*>(p)[slot] = __float2half_rn(v);
  }
};

template <typename T, typename V>
__device__ __forceinline__ void store_arg(void* out, int index_dtype, void* val_out, long long slot,
                                          const ArgPair<V>& p) {
  if (index_dtype == B200PHI_INT32)
    reinterpret_cast<int*>(out)[slot] = static_cast<int>(p.i);
  else
    reinterpret_cast<long long*>(out)[slot] = p.i;
  if (val_out) StoreVal<T>::put(val_out, slot, p.v);
}

// Extreme value of one 16-byte vector: a short tree instead of VE dependent compare/select
// steps (the per-element chain made config C3b latency-bound at 5.1 TB/s).  16-bit floats use
// the packed min/max instructions on the raw pairs.
template <typename T, typename V, bool kMax>
struct VecExtreme {
  static __device__ __forceinline__ V of(const V16& raw) {
    typename VecOf<T>::U w;
    w.raw = raw;
    V m[VecOf<T>::kElems];
#pragma unroll
    for (int e = 0; e < VecOf<T>::kElems; ++e) m[e] = Cvt<V, T>::to(w.e[e]);
#pragma unroll
    for (int half = VecOf<T>::kElems / 2; half > 0; half >>= 1)
#pragma unroll
      for (int e = 0; e < half; ++e) m[e] = ArgOp<V, kMax>::better(m[e + half], m[e]) ? m[e + half] : m[e];
    return m[0];
  }
};
template <bool kMax>
struct VecExtreme<__nv_bfloat16, float, kMax> {
  static __device__ __forceinline__ float of(const V16& raw) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
    __nv_bfloat162 a = kMax ? __hmax2(h[0], h[1]) : __hmin2(h[0], h[1]);
    __nv_bfloat162 b = kMax ? __hmax2(h[2], h[3]) : __hmin2(h[2], h[3]);
    a = kMax ? __hmax2(a, b) : __hmin2(a, b);
    const float lo = __low2float(a), hi = __high2float(a);
    return kMax ? fmaxf(lo, hi) : fminf(lo, hi);
  }
};
template <bool kMax>
struct VecExtreme<__half, float, kMax> {
  static __device__ __forceinline__ float of(const V16& raw) {
    const __half2* h = reinterpret_cast<const __half2*>(&raw);
    __half2 a = kMax ? __hmax2(h[0], h[1]) : __hmin2(h[0], h[1]);
    __half2 b = kMax ? __hmax2(h[2], h[3]) : __hmin2(h[2], h[3]);
    a = kMax ? __hmax2(a, b) : __hmin2(a, b);
    const float lo = __low2float(a), hi = __high2float(a);
    return kMax ? fmaxf(lo, hi) : fminf(lo, hi);
  }
};

// Visit one vector whose first element has index i0.  Only when the vector's extreme beats
// the thread's best (rare after the first few vectors) is the position inside it looked up;
// scanning downwards leaves the LOWEST matching position.
template <typename T, typename V, bool kMax>
__device__ __forceinline__ void visit_vec(ArgPair<V>& best, const V16& raw, long long i0) {
  const V vm = VecExtreme<T, V, kMax>::of(raw);
  if (ArgOp<V, kMax>::better(vm, best.v)) {
    typename VecOf<T>::U w;
    w.raw = raw;
    int first = 0;
#pragma unroll
    for (int e = VecOf<T>::kElems - 1; e >= 0; --e)
      if (Cvt<V, T>::to(w.e[e]) == vm) first = e;
    best.v = vm;
    best.i = i0 + first;
  }
}

// Batches of UU strided vectors, all loads issued before the first visit.  The caller
// cascades UU = 8, 4, 2, 1 so the remainder of a row never degenerates into one load at a time
// (a 62.5 KiB bf16 row left up to 7 serial loads per thread: 5.8 TB/s on C3b).
template <typename T, typename V, bool kMax, int G, int UU>
__device__ __forceinline__ void arg_step(ArgPair<V>& best, const V16* __restrict__ vp, long long& v,
                                         long long nvec, long long head) {
  constexpr int VE = VecOf<T>::kElems;
  for (; v + static_cast<long long>(UU - 1) * G < nvec; v += static_cast<long long>(UU) * G) {
    V16 buf[UU];
#pragma unroll
    for (int u = 0; u < UU; ++u) buf[u] = ldg_stream16(vp + v + static_cast<long long>(u) * G);
#pragma unroll
    for (int u = 0; u < UU; ++u)
      visit_vec<T, V, kMax>(best, buf[u], head + (v + static_cast<long long>(u) * G) * VE);
  }
}

// Scan `n` contiguous elements with a group of G threads.  kAligned: the row starts on a
// 16-byte boundary and is a whole number of vectors (no head / tail).
template <typename T, typename V, bool kMax, int G, int U, bool kAligned>
__device__ __forceinline__ ArgPair<V> arg_span(const T* __restrict__ row, long long n, int g) {
  using Op = ArgOp<V, kMax>;
  constexpr int VE = VecOf<T>::kElems;
  ArgPair<V> best = Op::identity();
  long long head = 0;
  if constexpr (!kAligned) {
    const uintptr_t addr = reinterpret_cast<uintptr_t>(row);
    head = static_cast<long long>(((16 - (addr & 15)) & 15) / sizeof(T));
    if (head > n) head = n;
    if (g < head) Op::visit(best, Cvt<V, T>::to(ldg_scalar(row + g)), g);
  }
  const long long nvec = (n - head) / VE;
  const long long tail0 = head + nvec * VE;
  const V16* __restrict__ vp = reinterpret_cast<const V16*>(row + head);
  long long v = g;
  arg_step<T, V, kMax, G, U>(best, vp, v, nvec, head);
  if constexpr (U > 4) arg_step<T, V, kMax, G, 4>(best, vp, v, nvec, head);
  if constexpr (U > 2) arg_step<T, V, kMax, G, 2>(best, vp, v, nvec, head);
  if constexpr (U > 1) arg_step<T, V, kMax, G, 1>(best, vp, v, nvec, head);
  if constexpr (!kAligned) {
    for (long long i = tail0 + g; i < n; i += G) Op::visit(best, Cvt<V, T>::to(ldg_scalar(row + i)), i);
  }
  return best;
}

}  // namespace b200phi
