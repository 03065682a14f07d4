// argmax / argmin for sm_100a.  Replaces ArgCUDAKernel + ComputeFullArg
// (paddle/phi/kernels/gpu/arg_min_max_kernel.cu:58-145): instead of one scalar
// strided load per element and a cub::BlockReduce of KeyValuePair<int,T>, rows are
// streamed with 16-byte loads, indices are 64-bit, and the combine rule
// "better value, else lower index" is applied at every level so ties are
// bit-exact with the CPU kernel (Eigen argmax, cpu/arg_min_max_kernel.cc:46-59).
#include <stdlib.h>

#include <string>

#include "../../include/b200phi.h"
#include "prim.cuh"
#include "reduce_plan.h"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();

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

// post == 1: [rows, n], G threads per row (as in row_reduce_kernel).
template <typename T, typename V, bool kMax, int G, bool kAligned, int U = 8>
__global__ void __launch_bounds__(kBlock)
    arg_row_kernel(const T* __restrict__ x, long long rows, long long n, void* out, int index_dtype,
                   void* val_out) {
  using Op = ArgOp<V, kMax>;
  __shared__ V s_v[2 * kWarps];
  __shared__ long long s_i[2 * kWarps];
  constexpr int kRowsPerCta = kBlock / G;
  constexpr int kW = G / 32;
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int parity = 0;
  for (long long row = static_cast<long long>(blockIdx.x) * kRowsPerCta + grp; row < rows;
       row += static_cast<long long>(gridDim.x) * kRowsPerCta) {
    ArgPair<V> p = arg_span<T, V, kMax, G, U, kAligned>(x + row * n, n, g);
    p = warp_arg<V, Op>(p);
    if constexpr (G > 32) {
      V* bv = s_v + parity * kWarps;
      long long* bi = s_i + parity * kWarps;
      parity ^= 1;
      if (lane == 0) {
        bv[warp] = p.v;
        bi[warp] = p.i;
      }
      group_barrier<G>();
      if (g == 0) {
        p.v = bv[grp * kW];
        p.i = bi[grp * kW];
#pragma unroll
        for (int w = 1; w < kW; ++w) {
          ArgPair<V> q;
          q.v = bv[grp * kW + w];
          q.i = bi[grp * kW + w];
          p = Op::combine(p, q);
        }
      }
    }
    if (g == 0) store_arg<T, V>(out, index_dtype, val_out, row, p);
  }
}

// Rows of at most 128 vectors: one warp takes R rows at a time, NV vectors per lane each
// (R * NV = 8 loads in flight), visits each row's vectors in index order and merges the R pairs
// with shared shuffle steps.  One row per warp left a 256-column bf16 row at one load per lane
// (1.7 TB/s).
template <typename T, typename V, bool kMax, int G, int NV, int R>
__global__ void __launch_bounds__(kBlock, 4)
    arg_multi_kernel(const T* __restrict__ x, long long rows, long long nvec, void* out, int index_dtype,
                     void* val_out) {
  // G = 8, 16 or 32 lanes per row: 32 / G rows side by side in a warp, R rounds of them
  using Op = ArgOp<V, kMax>;
  constexpr int VE = VecOf<T>::kElems;
  constexpr int kPerWarp = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  const long long n = nvec * VE;
  for (long long c0 = static_cast<long long>(blockIdx.x) * (kWarps * kPerWarp * R); c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * (kWarps * kPerWarp * R)) {
    const long long w0 = c0 + warp * (kPerWarp * R);
    const long long r0 = w0 + static_cast<long long>(sub) * R;
    V16 buf[R][NV];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const V16* __restrict__ vp = reinterpret_cast<const V16*>(x + (r0 + r) * n);
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long v = g + static_cast<long long>(G) * k;
        if (r0 + r < rows && v < nvec) buf[r][k] = ldg_stream16(vp + v);
      }
    }
    ArgPair<V> best[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      best[r] = Op::identity();
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long v = g + static_cast<long long>(G) * k;
        if (r0 + r < rows && v < nvec) visit_vec<T, V, kMax>(best[r], buf[r][k], v * VE);
      }
    }
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) best[r] = Op::combine(best[r], shfl_pair(best[r], m));
    }
    if (g == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r)
        if (r0 + r < rows) store_arg<T, V>(out, index_dtype, val_out, r0 + r, best[r]);
    }
  }
}

// Short rows of any alignment: G = 8, 16 or 32 lanes per row, NE scalar loads per lane visited in
// index order (strict compare keeps the first occurrence), pairs merged inside the G lanes.
template <typename T, typename V, bool kMax, int G, int NE>
__global__ void __launch_bounds__(kBlock)
    arg_scalar_kernel(const T* __restrict__ x, long long rows, int n, void* out, int index_dtype,
                      void* val_out) {
  using Op = ArgOp<V, kMax>;
  constexpr int kPerWarp = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  // (loop bound on blockIdx alone: the compiler then emits bare shuffles, not checked collectives)
  for (long long c0 = static_cast<long long>(blockIdx.x) * (kWarps * kPerWarp); c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * (kWarps * kPerWarp)) {
    const long long w0 = c0 + warp * kPerWarp;
    const long long row = w0 + sub;
    const bool live = row < rows;
    const T* __restrict__ r = x + row * n;
    T raw[NE];
    int odd;
    short_row_load<T, G, NE>(r, n, g, live, raw, &odd);
    ArgPair<V> best = Op::identity();
#pragma unroll
    for (int k = 0; k < NE; ++k) {  // ascending element index per lane: strict compare keeps the first
      const int i = short_row_index<T, G, NE>(k, g, odd);
      if (live && i >= 0 && i < n) Op::visit(best, Cvt<V, T>::to(raw[k]), i);
    }
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) best = Op::combine(best, shfl_pair(best, m));
    if (live && g == 0) store_arg<T, V>(out, index_dtype, val_out, row, best);
  }
}

// 16-bit rows of 0.5 .. 4 KiB that are not whole aligned vectors (1001-, 2049-column rows): one
// WARP per row, all loads issued together -- up to 7 head elements (one per lane), NV 16-byte
// vectors per lane from the first aligned address, up to 7 tail elements -- and visited in
// ascending index order per lane, so the strict compare keeps the first occurrence.  The scalar
// kernel above reads such a row with 2-byte loads (bf16 1001 columns: 1.5 TB/s), the peeled
// arg_row_kernel in three dependent phases (2.7 TB/s).
template <typename T, typename V, bool kMax, int NV>
__global__ void __launch_bounds__(kBlock)
    arg_warp_peel_kernel(const T* __restrict__ x, long long rows, int n, void* out, int index_dtype,
                         void* val_out) {
  using Op = ArgOp<V, kMax>;
  constexpr int VE = VecOf<T>::kElems;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // (loop bound on blockIdx alone: bare shuffles)
  for (long long c0 = static_cast<long long>(blockIdx.x) * kWarps; c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * kWarps) {
    const long long row = c0 + warp;
    const bool live = row < rows;
    const T* __restrict__ r = x + (live ? row : 0) * n;
    const uintptr_t addr = reinterpret_cast<uintptr_t>(r);
    int head = static_cast<int>(((16 - (addr & 15)) & 15) / sizeof(T));
    if (head > n) head = n;
    const int nvec = (n - head) / VE;
    const int tail0 = head + nvec * VE;
    const V16* __restrict__ vp = reinterpret_cast<const V16*>(r + head);
    const bool has_h = live && lane < head, has_t = live && tail0 + lane < n;
    T hv, tv;
    if (has_h) hv = ldg_scalar(r + lane);
    if (has_t) tv = ldg_scalar(r + tail0 + lane);
    V16 buf[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = lane + 32 * k;
      if (live && v < nvec) buf[k] = ldg_stream16(vp + v);
    }
    ArgPair<V> best = Op::identity();
    if (has_h) Op::visit(best, Cvt<V, T>::to(hv), lane);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = lane + 32 * k;
      if (live && v < nvec) visit_vec<T, V, kMax>(best, buf[k], head + static_cast<long long>(v) * VE);
    }
    if (has_t) Op::visit(best, Cvt<V, T>::to(tv), tail0 + lane);
    best = warp_arg<V, Op>(best);
    if (live && lane == 0) store_arg<T, V>(out, index_dtype, val_out, row, best);
  }
}

// post > 1: [pre, n, post].  Block = TX x TY: tx walks post (contiguous), ty walks n.
template <typename T, typename V, bool kMax>
__global__ void __launch_bounds__(kBlock)
    arg_col_kernel(const T* __restrict__ x, long long n, long long post, int TX, long long tiles,
                   void* out, int index_dtype, void* val_out) {
  using Op = ArgOp<V, kMax>;
  __shared__ V s_v[kBlock];
  __shared__ long long s_i[kBlock];
  const int TY = kBlock / TX;
  const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
  const long long o = blockIdx.x / tiles, tile = blockIdx.x % tiles;
  const long long c = tile * TX + tx;
  const bool live = c < post;
  ArgPair<V> best = Op::identity();
  if (live) {
    const T* __restrict__ base = x + o * n * post + c;
    long long r = ty;
    for (; r + 3LL * TY < n; r += 4LL * TY) {
      T b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) b[u] = ldg_scalar(base + (r + static_cast<long long>(u) * TY) * post);
#pragma unroll
      for (int u = 0; u < 4; ++u) Op::visit(best, Cvt<V, T>::to(b[u]), r + static_cast<long long>(u) * TY);
    }
    for (; r < n; r += TY) Op::visit(best, Cvt<V, T>::to(ldg_scalar(base + r * post)), r);
  }
  s_v[ty * TX + tx] = best.v;
  s_i[ty * TX + tx] = best.i;
  __syncthreads();
  for (int half = TY >> 1; half > 0; half >>= 1) {
    if (ty < half) {
      ArgPair<V> a, b;
      a.v = s_v[ty * TX + tx];
      a.i = s_i[ty * TX + tx];
      b.v = s_v[(ty + half) * TX + tx];
      b.i = s_i[(ty + half) * TX + tx];
      a = Op::combine(a, b);
      s_v[ty * TX + tx] = a.v;
      s_i[ty * TX + tx] = a.i;
    }
    __syncthreads();
  }
  if (ty == 0 && live) {
    ArgPair<V> r;
    r.v = s_v[tx];
    r.i = s_i[tx];
    store_arg<T, V>(out, index_dtype, val_out, o * post + c, r);
  }
}

struct ArgGeom {
  long long pre, n, post;
};

template <typename T, typename V, bool kMax>
static int launch_arg(const T* x, const ArgGeom& g, void* out, int index_dtype, void* val_out,
                      cudaStream_t stream) {
  if (g.post == 1) {
    const long long row_bytes = g.n * static_cast<long long>(sizeof(T));
    // threads per row: about 32 vectors per thread, 8 loads in flight each (measured on C3b,
    // 4000 vectors per row: G=128 reads 6.22 TB/s, G=256 5.98, G=64 5.64).  Long rows of 16-bit
    // elements do better with the whole CTA on one row and 4 loads in flight (40 registers, six
    // CTAs per SM): 6.38 TB/s.
    int G = 32;
    while (G < 256 && static_cast<long long>(G) * 512 < row_bytes) G <<= 1;
    static const int force_g = getenv("B200PHI_ARG_G") ? atoi(getenv("B200PHI_ARG_G")) : 0;
    static const int force_u = getenv("B200PHI_ARG_U") ? atoi(getenv("B200PHI_ARG_U")) : 0;
    int arg_u = 8;
    if (sizeof(T) == 2 && row_bytes > 32768 && !force_g) {
      G = 256;
      arg_u = 4;
    }
    if (force_g == 32 || force_g == 64 || force_g == 128 || force_g == 256) G = force_g;
    if (force_u == 4 || force_u == 8) arg_u = force_u;
    const bool aligned = reinterpret_cast<uintptr_t>(x) % 16 == 0 && row_bytes % 16 == 0;
    static const bool no_multi = getenv("B200PHI_ARG_NO_MULTI") != nullptr;
    const long long nvec_row = row_bytes / 16;
    if (aligned && !no_multi && !force_g && !force_u && nvec_row <= 128 && g.pre >= 64) {
      const int LG = nvec_row <= 32 ? 8 : 16;
      const long long rows_per_warp = (32 / LG) * (nvec_row <= 64 ? 2 : 1);
      long long mc = (g.pre + kWarps * rows_per_warp - 1) / (kWarps * rows_per_warp);
      const long long cap = 148LL * 8 * 8;
      if (mc > cap) mc = cap;
      const unsigned gm = static_cast<unsigned>(mc);
      if (nvec_row <= 32)
        arg_multi_kernel<T, V, kMax, 8, 4, 2><<<gm, kBlock, 0, stream>>>(x, g.pre, nvec_row, out, index_dtype, val_out);
      else if (nvec_row <= 64)
        arg_multi_kernel<T, V, kMax, 16, 4, 2><<<gm, kBlock, 0, stream>>>(x, g.pre, nvec_row, out, index_dtype, val_out);
      else
        arg_multi_kernel<T, V, kMax, 16, 8, 1><<<gm, kBlock, 0, stream>>>(x, g.pre, nvec_row, out, index_dtype, val_out);
      cudaError_t e0 = cudaGetLastError();
      if (e0 != cudaSuccess) return cuda_status(e0);
      return B200PHI_OK;
    }
    // (measured at 1001 columns: fp32 6.5 TB/s against 4.3 for the peeled kernel, bf16 1.5 against
    // 2.7 -- 2-byte scalar loads -- so 16-bit rows stop at 256 columns)
    if (!aligned && g.n < (sizeof(T) == 2 ? 256 : 1024) && g.pre >= 64 && !force_g && !force_u) {
      const int c = static_cast<int>(g.n);
#define B200PHI_ARG_SCALAR(GG, NN)                                                              \
  {                                                                                             \
    long long sc = (g.pre + kWarps * (32 / GG) - 1) / (kWarps * (32 / GG));                     \
    if (sc > 148LL * 8 * 8) sc = 148LL * 8 * 8;                                                 \
    arg_scalar_kernel<T, V, kMax, GG, NN><<<static_cast<unsigned>(sc), kBlock, 0, stream>>>(    \
        x, g.pre, c, out, index_dtype, val_out);                                                \
  }
      if (g.n < 128)
        B200PHI_ARG_SCALAR(8, 16)
      else if (g.n < 256)
        B200PHI_ARG_SCALAR(16, 16)
      else if (g.n < 512)
        B200PHI_ARG_SCALAR(32, 16)
      else
        B200PHI_ARG_SCALAR(32, 32)
#undef B200PHI_ARG_SCALAR
      cudaError_t e1 = cudaGetLastError();
      if (e1 != cudaSuccess) return cuda_status(e1);
      return B200PHI_OK;
    }
    if constexpr (sizeof(T) == 2) {
      static const bool no_peel = getenv("B200PHI_ARG_NO_PEEL") != nullptr;
      // (measured, bf16: 1001 columns 4.89 TB/s against 2.65 for the peeled row kernel; at 4097
      // columns a 16-vector form was slower, 4.78 against 5.39, so rows stop at 4 KiB here)
      if (!aligned && !no_peel && g.n >= 256 && row_bytes < 4112 && g.pre >= 64 && !force_g && !force_u) {
        long long pc = (g.pre + kWarps - 1) / kWarps;
        if (pc > 148LL * 8 * 8) pc = 148LL * 8 * 8;
        const unsigned gp = static_cast<unsigned>(pc);
        const int c = static_cast<int>(g.n);
        if (row_bytes < 2064)
          arg_warp_peel_kernel<T, V, kMax, 4><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
        else
          arg_warp_peel_kernel<T, V, kMax, 8><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
        cudaError_t e2 = cudaGetLastError();
        if (e2 != cudaSuccess) return cuda_status(e2);
        return B200PHI_OK;
      }
    }
    long long ctas = (g.pre + (kBlock / G) - 1) / (kBlock / G);
    if (ctas > 0x7fffffffLL) ctas = 0x7fffffffLL;
    const unsigned grid = static_cast<unsigned>(ctas);
#define B200PHI_ARG_ROW(GG)                                                                          \
  if (aligned && arg_u == 4)                                                                         \
    arg_row_kernel<T, V, kMax, GG, true, 4><<<grid, kBlock, 0, stream>>>(x, g.pre, g.n, out, index_dtype, val_out); \
  else if (aligned)                                                                                  \
    arg_row_kernel<T, V, kMax, GG, true><<<grid, kBlock, 0, stream>>>(x, g.pre, g.n, out, index_dtype, val_out); \
  else                                                                                               \
    arg_row_kernel<T, V, kMax, GG, false><<<grid, kBlock, 0, stream>>>(x, g.pre, g.n, out, index_dtype, val_out);
    switch (G) {
      case 32:
        B200PHI_ARG_ROW(32) break;
      case 64:
        B200PHI_ARG_ROW(64) break;
      case 128:
        B200PHI_ARG_ROW(128) break;
      default:
        B200PHI_ARG_ROW(256) break;
    }
#undef B200PHI_ARG_ROW
  } else {
    int TX = 1;
    while (TX < g.post && TX < kBlock) TX <<= 1;
    const long long tiles = (g.post + TX - 1) / TX;
    const long long blocks = g.pre * tiles;
    if (blocks > 0x7fffffffLL) return set_error(B200PHI_UNSUPPORTED, "argreduce: output too large");
    arg_col_kernel<T, V, kMax><<<static_cast<unsigned>(blocks), kBlock, 0, stream>>>(
        x, g.n, g.post, TX, tiles, out, index_dtype, val_out);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

template <typename T, typename V>
static int launch_arg_op(int op, const void* x, const ArgGeom& g, void* out, int index_dtype,
                         void* val_out, cudaStream_t s) {
  if (op == B200PHI_ARGMAX)
    return launch_arg<T, V, true>(static_cast<const T*>(x), g, out, index_dtype, val_out, s);
  return launch_arg<T, V, false>(static_cast<const T*>(x), g, out, index_dtype, val_out, s);
}

int run_argreduce(const b200phi_argreduce_desc* d, cudaStream_t stream) {
  if (!d) return set_error(B200PHI_INVALID_ARGUMENT, "null descriptor");
  if (d->op != B200PHI_ARGMAX && d->op != B200PHI_ARGMIN)
    return set_error(B200PHI_INVALID_ARGUMENT, "unknown arg op");
  if (d->rank < 0 || d->rank > B200PHI_MAX_RANK)
    return set_error(B200PHI_INVALID_ARGUMENT, "rank must be in [0, 9]");
  int index_dtype = d->index_dtype == B200PHI_UNDEFINED ? B200PHI_INT64 : d->index_dtype;
  if (index_dtype != B200PHI_INT32 && index_dtype != B200PHI_INT64)
    return set_error(B200PHI_INVALID_ARGUMENT, "argmax/argmin index dtype must be int32 or int64");
  long long numel = 1;
  for (int i = 0; i < d->rank; ++i) {
    if (d->dims[i] < 0) return set_error(B200PHI_INVALID_ARGUMENT, "negative extent");
    numel *= d->dims[i];
  }
  if (numel <= 0)
    return set_error(B200PHI_INVALID_ARGUMENT, "argmin/argmax input numel must > 0, bug got 0");
  if (!d->x || !d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  ArgGeom g;
  if (d->flatten || d->rank == 0) {
    g.pre = 1;
    g.n = numel;
    g.post = 1;
  } else {
    long long axis = d->axis;
    if (axis < -d->rank || axis >= d->rank)
      return set_error(B200PHI_INVALID_ARGUMENT, "argmax/argmin axis out of range");
    if (axis < 0) axis += d->rank;
    g.pre = g.post = 1;
    for (int i = 0; i < axis; ++i) g.pre *= d->dims[i];
    g.n = d->dims[axis];
    for (int i = static_cast<int>(axis) + 1; i < d->rank; ++i) g.post *= d->dims[i];
  }
  if (index_dtype == B200PHI_INT32 && g.n > 0x7fffffffLL)
    return set_error(B200PHI_INVALID_ARGUMENT, "int32 indices need the reduced extent <= INT_MAX");
  switch (d->in_dtype) {
    case B200PHI_FLOAT32:
      return launch_arg_op<float, float>(d->op, d->x, g, d->out, index_dtype, d->max_out, stream);
    case B200PHI_FLOAT64:
      return launch_arg_op<double, double>(d->op, d->x, g, d->out, index_dtype, d->max_out, stream);
    case B200PHI_BFLOAT16:
      return launch_arg_op<__nv_bfloat16, float>(d->op, d->x, g, d->out, index_dtype, d->max_out, stream);
    case B200PHI_FLOAT16:
      return launch_arg_op<__half, float>(d->op, d->x, g, d->out, index_dtype, d->max_out, stream);
    case B200PHI_INT64:
      return launch_arg_op<long long, long long>(d->op, d->x, g, d->out, index_dtype, d->max_out, stream);
    case B200PHI_INT32:
      return launch_arg_op<int32_t, int32_t>(d->op, d->x, g, d->out, index_dtype, d->max_out, stream);
    case B200PHI_INT16:
      return launch_arg_op<int16_t, int32_t>(d->op, d->x, g, d->out, index_dtype, d->max_out, stream);
    case B200PHI_UINT8:
      return launch_arg_op<uint8_t, int32_t>(d->op, d->x, g, d->out, index_dtype, d->max_out, stream);
    default:
      return set_error(B200PHI_UNSUPPORTED, "argmax/argmin: dtype not built");
  }
}

}  // namespace b200phi
