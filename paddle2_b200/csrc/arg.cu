// argmax / argmin for sm_100a.  Replaces ArgCUDAKernel + ComputeFullArg
// (paddle/phi/kernels/gpu/arg_min_max_kernel.cu:58-145): instead of one scalar
// strided load per element and a cub::BlockReduce of KeyValuePair<int,T>, rows are
// streamed with 16-byte loads, indices are 64-bit, and the combine rule
// "better value, else lower index" is applied at every level so ties are
// bit-exact with the CPU kernel (Eigen argmax, cpu/arg_min_max_kernel.cc:46-59).
#include <stdio.h>
#include <stdlib.h>

#include <string>

#include "../../include/b200phi.h"
#include "arg_common.cuh"
#include "rows_staged.cuh"
#include "arg_plan.h"
#include "reduce_plan.h"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();
Tunables tunables();
int launch_arg_split(int op, const ArgPlan& p, const void* x, void* out, void* val_out, void* ws,
                     cudaStream_t s);

// probe knobs (environment, read once per process)
struct ArgKnobs {
  int force_g, force_u;
  bool no_multi, no_peel;
};
static const ArgKnobs& arg_knobs() {
  static const ArgKnobs k = {getenv("B200PHI_ARG_G") ? atoi(getenv("B200PHI_ARG_G")) : 0,
                             getenv("B200PHI_ARG_U") ? atoi(getenv("B200PHI_ARG_U")) : 0,
                             getenv("B200PHI_ARG_NO_MULTI") != nullptr, getenv("B200PHI_ARG_NO_PEEL") != nullptr};
  return k;
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
// G < 32: 32 / G rows side by side in a warp (a 201-column bf16 row is 25 vectors); G >= VE - 1 so
// that every possible head / tail element has a lane.
// The row sits in registers, so the search runs in two sweeps of them: first the extreme VALUE
// (packed max tree per vector, a value-only shuffle tree), then, only in the lanes that hold it,
// the lowest position whose value equals it, and an integer min over the lanes.  Carrying (value,
// index) pairs through every vector and every shuffle step cost 2.3x the instructions, and the
// kernel is bound by instruction issue (ncu: 74 % issue-slot use at 3.0 TB/s on 101 bf16 columns).
// "First position equal to the extreme" is what the strict in-order compare of the reference
// returns (cpu/arg_min_max_kernel.cc:46-59); an all-identity row answers 0.
template <typename T, typename V, bool kMax, int NV, int G = 32>
__global__ void __launch_bounds__(kBlock, NV <= 8 ? 4 : 2)
    arg_warp_peel_kernel(const T* __restrict__ x, long long rows, int n, void* out, int index_dtype,
                         void* val_out) {
  using Op = ArgOp<V, kMax>;
  constexpr int VE = VecOf<T>::kElems;
  static_assert(G >= VE - 1, "a lane per possible head / tail element");
  constexpr int kPerWarp = 32 / G;
  constexpr int kPerCta = kWarps * kPerWarp;
  constexpr int kNone = 0x7fffffff;
  const int wl = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lane = wl % G, sub = wl / G;
  // (loop bound on blockIdx alone: bare shuffles; 32-bit row arithmetic on a CTA-uniform base)
  for (long long c0 = static_cast<long long>(blockIdx.x) * kPerCta; c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * kPerCta) {
    const T* __restrict__ xb = x + c0 * n;
    const unsigned xlow = static_cast<unsigned>(reinterpret_cast<uintptr_t>(xb)) & 15u;
    const long long left = rows - c0;
    const int lr = warp * kPerWarp + sub;
    const bool live = lr < (left < kPerCta ? static_cast<int>(left) : kPerCta);
    const unsigned off = static_cast<unsigned>(lr) * static_cast<unsigned>(n);
    const T* __restrict__ r = xb + off;
    int head = static_cast<int>(((16u - ((xlow + off * static_cast<unsigned>(sizeof(T))) & 15u)) & 15u) / sizeof(T));
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
      const int v = lane + G * k;
      if (live && v < nvec) buf[k] = ldg_stream16(vp + v);
    }
    // sweep 1: the extreme value
    const V ident = Op::identity().v;
    const V hf = has_h ? Cvt<V, T>::to(hv) : ident, tf = has_t ? Cvt<V, T>::to(tv) : ident;
    V vm[NV];
    V m = Op::better(tf, hf) ? tf : hf;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = lane + G * k;
      vm[k] = (live && v < nvec) ? VecExtreme<T, V, kMax>::of(buf[k]) : ident;
      m = Op::better(vm[k], m) ? vm[k] : m;
    }
    V best = m;
#pragma unroll
    for (int s = G / 2; s > 0; s >>= 1) {
      const V o = shfl_xor(best, s);
      best = Op::better(o, best) ? o : best;
    }
    // sweep 2: the lowest position that holds it -- head, vectors in ascending order, tail
    int idx = kNone;
    if (m == best) {
      if (has_h && hf == best) idx = lane;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int v = lane + G * k;
        if (idx == kNone && live && v < nvec && vm[k] == best) {
          typename VecOf<T>::U w;
          w.raw = buf[k];
          int first = 0;
#pragma unroll
          for (int e = VE - 1; e >= 0; --e)
            if (Cvt<V, T>::to(w.e[e]) == best) first = e;
          idx = head + v * VE + first;
        }
      }
      if (idx == kNone && has_t && tf == best) idx = tail0 + lane;
    }
#pragma unroll
    for (int s = G / 2; s > 0; s >>= 1) {
      const int o = shfl_xor(idx, s);
      idx = o < idx ? o : idx;
    }
    if (live && lane == 0) {
      ArgPair<V> p;
      p.v = best;
      p.i = idx == kNone ? 0 : idx;
      store_arg<T, V>(out, index_dtype, val_out, c0 + lr, p);
    }
  }
}

// argmax / argmin consumer of rows_staged_kernel (rows_staged.cuh): one thread scans its row in
// ascending order with the strict compare, so it keeps the first occurrence.
template <typename T, typename V, bool kMax>
struct StagedArg {
  struct Params {
    void* out;
    int index_dtype;
    void* val_out;
  };
  template <int L>
  static __device__ __forceinline__ void run(const T* __restrict__ sp, int n, int i0, int i1, int rot,
                                             bool live, long long row, int h, const Params& p) {
    using Op = ArgOp<V, kMax>;
    constexpr int kNone = 0x7fffffff;
    V bv = Op::identity().v;
    int idx = kNone;
    if (rot >= 0) {
      // wrapped visit: not in index order, so equal values are settled by the lower index
      smem_visit<T>(sp, n, i0, i1, rot, [&](int i, auto x, int) {
        const V v = Cvt<V, decltype(x)>::to(x);
        if (Op::better(v, bv) || (v == bv && i < idx)) {
          bv = v;
          idx = i;
        }
      });
      if (idx != kNone && !Op::better(bv, Op::identity().v)) idx = kNone;  // nothing beat the identity
    } else {
      smem_walk_range<T>(sp, i0, i1, [&](int i, auto x, int) {  // ascending: strict compare keeps the first
        const V v = Cvt<V, decltype(x)>::to(x);
        if (Op::better(v, bv)) {
          bv = v;
          idx = i;
        }
      });
    }
#pragma unroll
    for (int m = L / 2; m > 0; m >>= 1) {  // better value, else lower index
      const V ov = __shfl_xor_sync(0xffffffffu, bv, m);
      const int oi = __shfl_xor_sync(0xffffffffu, idx, m);
      const bool take = Op::better(ov, bv) || (ov == bv && oi < idx);
      bv = take ? ov : bv;
      idx = take ? oi : idx;
    }
    if (live && h == 0) {
      ArgPair<V> best;
      best.v = bv;
      best.i = idx == kNone ? 0 : idx;  // an all-identity row answers 0
      store_arg<T, V>(p.out, p.index_dtype, p.val_out, row, best);
    }
  }
};

struct ArgGeom {
  long long pre, n, post;
};

template <typename T, typename V, bool kMax>
static int launch_arg(const T* x, const ArgGeom& g, void* out, int index_dtype, void* val_out,
                      cudaStream_t stream) {
  const long long sm_cap = static_cast<long long>(current_sm_count()) * 8 * 8;  // grid-stride beyond that
  {  // post == 1 (the column and split plans are arg_split.cu's)
    const long long row_bytes = g.n * static_cast<long long>(sizeof(T));
    // threads per row: about 32 vectors per thread, 8 loads in flight each (measured on C3b,
    // 4000 vectors per row: G=128 reads 6.22 TB/s, G=256 5.98, G=64 5.64).  Long rows of 16-bit
    // elements do better with the whole CTA on one row and 4 loads in flight (40 registers, six
    // CTAs per SM): 6.38 TB/s.
    int G = 32;
    while (G < 256 && static_cast<long long>(G) * 512 < row_bytes) G <<= 1;
    const int force_g = arg_knobs().force_g, force_u = arg_knobs().force_u;
    int arg_u = 8;
    if (sizeof(T) == 2 && row_bytes > 32768 && !force_g) {
      G = 256;
      arg_u = 4;
    }
    if (force_g == 32 || force_g == 64 || force_g == 128 || force_g == 256) G = force_g;
    if (force_u == 4 || force_u == 8) arg_u = force_u;
    const bool aligned = reinterpret_cast<uintptr_t>(x) % 16 == 0 && row_bytes % 16 == 0;
    const bool no_multi = arg_knobs().no_multi;
    const long long nvec_row = row_bytes / 16;
    // short rows of any alignment: a thread per row out of shared memory (rows_staged.cuh)
    static const bool no_staged = getenv("B200PHI_NO_ROW_STAGED") != nullptr;
    if (!no_staged && !no_multi && !force_g && !force_u && staged_rows_ok(g.pre, row_bytes, static_cast<int>(sizeof(T)), kStagedArg, reinterpret_cast<uintptr_t>(x) % 16 == 0)) {
      typename StagedArg<T, V, kMax>::Params sp{out, index_dtype, val_out};
      return launch_rows_staged<T, StagedArg<T, V, kMax>>(x, g.pre, static_cast<int>(g.n), sp, stream);
    }
    // (16-bit rows of 33..128 vectors go to the value-sweep + position-sweep kernel below whether
    // aligned or not -- bf16 512 / 1024 columns 7.1 TB/s against 5.8 / 6.4 on arg_multi_kernel, 256
    // columns 5.75 against 6.02; B200PHI_ARG_PEEL_ALIGNED=0 keeps aligned rows on arg_multi_kernel)
    static const bool peel_aligned = !(getenv("B200PHI_ARG_PEEL_ALIGNED") && atoi(getenv("B200PHI_ARG_PEEL_ALIGNED")) == 0);
    const bool to_peel = sizeof(T) == 2 && peel_aligned && nvec_row > 32 && !arg_knobs().no_peel;
    if (aligned && !to_peel && !no_multi && !force_g && !force_u && nvec_row <= 128 && g.pre >= 64) {
      const int LG = nvec_row <= 32 ? 8 : 16;
      const long long rows_per_warp = (32 / LG) * (nvec_row <= 64 ? 2 : 1);
      long long mc = (g.pre + kWarps * rows_per_warp - 1) / (kWarps * rows_per_warp);
      const long long cap = sm_cap;
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
    if constexpr (sizeof(T) == 2) {
      // 16-bit rows of 32..1031 columns that are not whole aligned vectors: 8 or 16 lanes per row,
      // head element + vectors + tail element per lane (the 2-byte-load kernel: 1.6-2.0 TB/s)
      const long long nv_row = g.n / VecOf<T>::kElems;
      if ((!aligned || (peel_aligned && nv_row > 32)) && nv_row >= 4 && nv_row <= 128 && g.pre >= 64 && !force_g && !force_u && !arg_knobs().no_peel) {
        const int lanes = nv_row <= 64 ? 8 : 16;
        long long pc = (g.pre + kWarps * (32 / lanes) - 1) / (kWarps * (32 / lanes));
        if (pc > sm_cap) pc = sm_cap;
        const unsigned gp = static_cast<unsigned>(pc);
        const int c = static_cast<int>(g.n);
        if (nv_row <= 16)
          arg_warp_peel_kernel<T, V, kMax, 2, 8><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
        else if (nv_row <= 32)
          arg_warp_peel_kernel<T, V, kMax, 4, 8><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
        else if (nv_row <= 64)
          arg_warp_peel_kernel<T, V, kMax, 8, 8><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
        else
          arg_warp_peel_kernel<T, V, kMax, 8, 16><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
        cudaError_t e3 = cudaGetLastError();
        if (e3 != cudaSuccess) return cuda_status(e3);
        return B200PHI_OK;
      }
    }
    // (measured at 1001 columns: fp32 6.5 TB/s against 4.3 for the peeled kernel, bf16 1.5 against
    // 2.7 -- 2-byte scalar loads -- so 16-bit rows stop at 256 columns)
    if (!aligned && g.n < (sizeof(T) == 2 ? 256 : 1024) && g.pre >= 64 && !force_g && !force_u) {
      const int c = static_cast<int>(g.n);
#define B200PHI_ARG_SCALAR(GG, NN)                                                              \
  {                                                                                             \
    long long sc = (g.pre + kWarps * (32 / GG) - 1) / (kWarps * (32 / GG));                     \
    if (sc > sm_cap) sc = sm_cap;                                                 \
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
      const bool no_peel = arg_knobs().no_peel;
      // (measured, bf16: 1001 columns 4.89 TB/s against 2.65 for the peeled row kernel; at 4097
      // columns a 16-vector form was slower, 4.78 against 5.39, so rows stop at 4 KiB here)
      // (aligned rows of 2..4 KiB too: B200PHI_ARG_WARP_PEEL_ALIGNED=0 keeps them on arg_row_kernel)
      static const bool warp_peel_aligned =
          !(getenv("B200PHI_ARG_WARP_PEEL_ALIGNED") && atoi(getenv("B200PHI_ARG_WARP_PEEL_ALIGNED")) == 0);
      static const long long warp_peel_max = getenv("B200PHI_ARG_WARP_PEEL_MAX") ? atoll(getenv("B200PHI_ARG_WARP_PEEL_MAX")) : 8208;
      if (((!aligned && row_bytes < warp_peel_max) || (warp_peel_aligned && nvec_row > 128 && row_bytes < 4112)) && !no_peel &&
          g.n >= 256 && g.pre >= 64 && !force_g && !force_u) {
        long long pc = (g.pre + kWarps - 1) / kWarps;
        if (pc > sm_cap) pc = sm_cap;
        const unsigned gp = static_cast<unsigned>(pc);
        const int c = static_cast<int>(g.n);
        if (row_bytes < 2064)
          arg_warp_peel_kernel<T, V, kMax, 4><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
        else if (row_bytes < 4112)
          arg_warp_peel_kernel<T, V, kMax, 8><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
        else
          arg_warp_peel_kernel<T, V, kMax, 16><<<gp, kBlock, 0, stream>>>(x, g.pre, c, out, index_dtype, val_out);
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

// Scratch for this descriptor: the worse of the aligned / unaligned plans, whatever d->x holds.
size_t argreduce_workspace_bytes(const b200phi_argreduce_desc* d) {
  size_t need = 0;
  const Tunables t = tunables();
  for (int k = 0; k < 2; ++k) {
    ArgPlan p;
    std::string err;
    if (plan_arg(d, t, k == 0, &p, &err) != B200PHI_OK) return 0;
    if (p.ws_total_bytes > need) need = p.ws_total_bytes;
  }
  return need;
}

int run_argreduce(const b200phi_argreduce_desc* d, cudaStream_t stream) {
  ArgPlan plan;
  std::string err;
  const bool x_aligned = d && reinterpret_cast<uintptr_t>(d->x) % 16 == 0;
  int rc = plan_arg(d, tunables(), x_aligned, &plan, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  if (!d->x || !d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  const int index_dtype = plan.index_dtype;
  if (plan.kind != kArgRows) {
    if (plan.ws_total_bytes > 0) {
      if (!d->workspace || d->workspace_bytes < plan.ws_total_bytes) {
        char buf[200];
        snprintf(buf, sizeof(buf),
                 "argmax/argmin: workspace too small: need %zu bytes (b200phi_argreduce_workspace), got %zu",
                 plan.ws_total_bytes, d->workspace ? d->workspace_bytes : (size_t)0);
        return set_error(B200PHI_WORKSPACE_TOO_SMALL, buf);
      }
      if (reinterpret_cast<uintptr_t>(d->workspace) % 256 != 0)
        return set_error(B200PHI_INVALID_ARGUMENT, "workspace must be 256-byte aligned");
    }
    return launch_arg_split(d->op, plan, d->x, d->out, d->max_out, d->workspace, stream);
  }
  ArgGeom g;
  g.pre = plan.pre;
  g.n = plan.n;
  g.post = plan.post;
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
