// Row logsumexp: out[r] = max_r + log(sum_j exp(x[r][j] - max_r)), one read of x.
// Replaces the reference's GPU routes (gpu/logsumexp_kernel.cu): a transpose +
// DispatchLogsumexpWarp when the reduced extent is <= 1024, else the five-launch fallback
// max -> subtract -> exp-sum -> log -> add (:49-78), i.e. three reads and one write of x.
//
// One pass: every thread keeps a running (m, s) with s = sum exp(x - m) over what it has seen.
// A 16-byte vector is folded with its own max first: when that beats m (rare once the thread
// has seen a few vectors) s is rescaled by exp(m - vmax), then the VE terms cost one FFMA + one
// MUFU.EX2 + one FADD each.  Pairs merge as M = max(m_a, m_b), S = s_a e^(m_a-M) + s_b e^(m_b-M):
// max-combine then sum-combine over the group with the usual shuffles.
// G threads per row (32..256, ~64 B of the row per thread per step), 8 loads in flight.
#include <string>

#include "../../include/b200phi.h"
#include "prim.cuh"
#include "rows_staged.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();

template <typename A>
struct LseMath;
template <>
struct LseMath<float> {
  static __device__ __forceinline__ float neg_inf() { return -INFINITY; }
  // exp(a - b) for finite b, 0 when a = -inf
  static __device__ __forceinline__ float exp_diff(float a, float b) {
    return exp2f((a - b) * 1.4426950408889634f);
  }
  static __device__ __forceinline__ float log_(float v) { return logf(v); }
  // exp(a - shift) as 2^(a log2e - shift log2e): bias(shift) once per row, then one FFMA and one
  // MUFU.EX2 per element (exp2f() spends five more instructions on range handling per call, which
  // made the short-row kernels issue-bound: 16 instructions per element on bf16 501 columns)
  static __device__ __forceinline__ float bias(float shift) { return shift * 1.4426950408889634f; }
  static __device__ __forceinline__ float term(float a, float b) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaf(a, 1.4426950408889634f, -b)));
    return r;
  }
};
template <>
struct LseMath<double> {
  static __device__ __forceinline__ double neg_inf() { return -INFINITY; }
  static __device__ __forceinline__ double exp_diff(double a, double b) { return exp(a - b); }
  static __device__ __forceinline__ double log_(double v) { return log(v); }
  static __device__ __forceinline__ double bias(double shift) { return shift; }
  static __device__ __forceinline__ double term(double a, double b) { return exp(a - b); }
};

// max of one 16-byte vector; 16-bit floats use the packed max on the raw pairs
template <typename T, typename A>
struct LseVecMax {
  static __device__ __forceinline__ A of(const V16& raw, const A* val) {
    A v = val[0];
#pragma unroll
    for (int e = 1; e < VecOf<T>::kElems; ++e) v = val[e] > v ? val[e] : v;
    return v;
  }
};
template <>
struct LseVecMax<__nv_bfloat16, float> {
  static __device__ __forceinline__ float of(const V16& raw, const float*) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
    __nv_bfloat162 a = __hmax2(__hmax2(h[0], h[1]), __hmax2(h[2], h[3]));
    return fmaxf(__low2float(a), __high2float(a));
  }
};
template <>
struct LseVecMax<__half, float> {
  static __device__ __forceinline__ float of(const V16& raw, const float*) {
    const __half2* h = reinterpret_cast<const __half2*>(&raw);
    __half2 a = __hmax2(__hmax2(h[0], h[1]), __hmax2(h[2], h[3]));
    return fmaxf(__low2float(a), __high2float(a));
  }
};

__device__ __forceinline__ float ex2_approx(float v) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}

// Branch-free on purpose: with a branch on "the vector's max beats m" nearly every step of a
// 32000-column row has some lane of the warp taking it, so both arms ran and the kernel issued
// 18 MUFU per vector instead of 9 (ncu: 17.9 instructions per element, issue-bound at 77 %).
template <typename T, typename A>
__device__ __forceinline__ void lse_fold_vec(const V16& raw, A& m, A& s) {
  constexpr int VE = VecOf<T>::kElems;
  A val[VE];
  unpack_vec<T, A>(raw, val);
  const A vmax = LseVecMax<T, A>::of(raw, val);
  const A mn = vmax > m ? vmax : m;
  if constexpr (sizeof(A) == 4) {
    // exp(x - mn) = 2^(x log2e - mn log2e): one FFMA, one MUFU.EX2, one FADD per element.
    // mn = -inf (nothing but -inf so far): shift by 0, every term is 2^(-inf) = 0.
    const float mb = (mn > -INFINITY) ? mn * 1.4426950408889634f : 0.f;
    const float scale = (m == mn) ? 1.f : ex2_approx(fmaf(m, 1.4426950408889634f, -mb));
    float t[2] = {0.f, 0.f};
#pragma unroll
    for (int e = 0; e < VE; ++e) t[e & 1] += ex2_approx(fmaf(val[e], 1.4426950408889634f, -mb));
    s = fmaf(s, scale, t[0] + t[1]);
  } else {
    const A shift = (mn > LseMath<A>::neg_inf()) ? mn : A(0);
    A t = A(0);
#pragma unroll
    for (int e = 0; e < VE; ++e) t += exp(val[e] - shift);
    s = s * ((m == mn) ? A(1) : exp(m - shift)) + t;
  }
  m = mn;
}

template <typename T, typename A>
__device__ __forceinline__ void lse_fold_scalar(T v, A& m, A& s) {
  const A a = Cvt<A, T>::to(v);
  const A mn = a > m ? a : m;
  const A shift = (mn > LseMath<A>::neg_inf()) ? mn : A(0);
  s = s * ((m == mn) ? A(1) : LseMath<A>::exp_diff(m, shift)) + LseMath<A>::exp_diff(a, shift);
  m = mn;
}

__device__ __forceinline__ void lse_put(void* p, int dtype, long long i, float v) {
  switch (dtype) {
    case B200PHI_FLOAT32:
      reinterpret_cast<float*>(p)[i] = v;
      break;
    case B200PHI_FLOAT16:
      reinterpret_cast<__half*>(p)[i] = __float2half_rn(v);
      break;
    default:
      reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
      break;
  }
}
__device__ __forceinline__ void lse_put(void* p, int dtype, long long i, double v) {
  switch (dtype) {
    case B200PHI_FLOAT32:
      reinterpret_cast<float*>(p)[i] = static_cast<float>(v);
      break;
    case B200PHI_FLOAT64:
      reinterpret_cast<double*>(p)[i] = v;
      break;
    case B200PHI_FLOAT16:
      reinterpret_cast<__half*>(p)[i] = __float2half_rn(static_cast<float>(v));
      break;
    default:
      reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(static_cast<float>(v));
      break;
  }
}

// kAligned: every row starts 16-byte aligned and cols is a multiple of VE.
// U vectors per thread are loaded before the first is folded.  Measured on bf16 [8192,32000]:
// U = 4 (39 registers, 6 CTAs/SM) 5.98 TB/s, U = 8 5.59, U = 16 3.90; issuing the next batch's
// loads before folding the current one (two register buffers) only cost occupancy: 5.30 / 3.85.
template <typename T, typename A, int G, int U, bool kAligned>
__global__ void __launch_bounds__(kBlock)
    lse_rows_kernel(const T* __restrict__ x, long long rows, long long cols, void* out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  __shared__ A smem[2 * kWarps];
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  const long long rows_per_cta = kBlock / G;
  // a group's threads share the row, and groups meet on their own named barrier: no padding
  for (long long row = static_cast<long long>(blockIdx.x) * rows_per_cta + grp; row < rows;
       row += static_cast<long long>(gridDim.x) * rows_per_cta) {
    A m = LseMath<A>::neg_inf(), s = A(0);
    {
      const T* __restrict__ r = x + row * cols;
      if constexpr (kAligned) {
        const V16* __restrict__ vp = reinterpret_cast<const V16*>(r);
        const long long nvec = cols / VE;
        {
          for (long long vb = g; vb < nvec; vb += static_cast<long long>(U) * G) {
            V16 buf[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const long long v = vb + static_cast<long long>(u) * G;
              if (v < nvec) buf[u] = ldg_stream16(vp + v);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const long long v = vb + static_cast<long long>(u) * G;
              if (v < nvec) lse_fold_vec<T, A>(buf[u], m, s);
            }
          }
        }
      } else {
        // any alignment (a 50257-entry vocabulary): scalar head up to the first 16-byte boundary,
        // vectors, scalar tail
        const uintptr_t addr = reinterpret_cast<uintptr_t>(r);
        long long head = static_cast<long long>(((16 - (addr & 15)) & 15) / sizeof(T));
        if (head > cols) head = cols;
        const long long nvec = (cols - head) / VE;
        const long long tail0 = head + nvec * VE;
        if (g < head) lse_fold_scalar<T, A>(ldg_scalar(r + g), m, s);
        const V16* __restrict__ vp = reinterpret_cast<const V16*>(r + head);
        for (long long vb = g; vb < nvec; vb += static_cast<long long>(U) * G) {
          V16 buf[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const long long v = vb + static_cast<long long>(u) * G;
            if (v < nvec) buf[u] = ldg_stream16(vp + v);
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const long long v = vb + static_cast<long long>(u) * G;
            if (v < nvec) lse_fold_vec<T, A>(buf[u], m, s);
          }
        }
        for (long long i = tail0 + g; i < cols; i += G) lse_fold_scalar<T, A>(ldg_scalar(r + i), m, s);
      }
    }
    // merge the group's pairs: max first, then the sums rescaled to it
    const A M = group_combine<A, MaxOp<A>, G>(m, smem);
    const A scaled = (m > LseMath<A>::neg_inf()) ? s * LseMath<A>::exp_diff(m, M) : A(0);
    const A S = group_combine<A, SumOp<A>, G>(scaled, smem + kWarps);
    if (g == 0) {
      // all -inf: max = -inf, sum = 0 -> -inf + log(0) = -inf, as the reference's expression gives
      lse_put(out, out_dtype, row, M + LseMath<A>::log_(S));
    }
  }
}

// Rows of at most 256 vectors (hidden / class counts up to 2048 bf16): one warp takes R rows at
// a time with NV vectors per lane each, R * NV = 8 loads in flight, and the R (max, sum) pairs
// share every shuffle step.  One row per warp left a 256-column bf16 row at one load per lane
// (1.7 TB/s; 1024 columns 4.3 TB/s).
template <typename T, typename A, int G, int NV, int R>
__global__ void __launch_bounds__(kBlock, 4)  // 64 registers: four CTAs per SM
    lse_multi_kernel(const T* __restrict__ x, long long rows, long long nvec, void* out, int out_dtype) {
  // G = 8, 16 or 32 lanes per row: 32 / G rows side by side in a warp, R rounds of them
  constexpr int VE = VecOf<T>::kElems;
  constexpr int kPerWarp = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  const long long cols = nvec * VE;
  // the loop bound depends on blockIdx alone: the body is provably convergent and the shuffles
  // compile to bare SHFL (a per-warp bound gave each one a convergence check + collective path)
  for (long long c0 = static_cast<long long>(blockIdx.x) * (kWarps * kPerWarp * R); c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * (kWarps * kPerWarp * R)) {
    const long long w0 = c0 + warp * (kPerWarp * R);
    const long long r0 = w0 + static_cast<long long>(sub) * R;
    V16 buf[R][NV];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const V16* __restrict__ vp = reinterpret_cast<const V16*>(x + (r0 + r) * cols);
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long v = g + static_cast<long long>(G) * k;
        if (r0 + r < rows && v < nvec) buf[r][k] = ldg_stream16(vp + v);
      }
    }
    A m[R], s[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      m[r] = LseMath<A>::neg_inf();
      s[r] = A(0);
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long v = g + static_cast<long long>(G) * k;
        if (r0 + r < rows && v < nvec) lse_fold_vec<T, A>(buf[r][k], m[r], s[r]);
      }
    }
    A M[R];
#pragma unroll
    for (int r = 0; r < R; ++r) M[r] = m[r];
#pragma unroll
    for (int k = G / 2; k > 0; k >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const A o = shfl_xor(M[r], k);
        M[r] = o > M[r] ? o : M[r];
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r)
      s[r] = (m[r] > LseMath<A>::neg_inf()) ? s[r] * LseMath<A>::exp_diff(m[r], M[r]) : A(0);
#pragma unroll
    for (int k = G / 2; k > 0; k >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) s[r] += shfl_xor(s[r], k);
    }
    if (g == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r)
        if (r0 + r < rows) lse_put(out, out_dtype, r0 + r, M[r] + LseMath<A>::log_(s[r]));
    }
  }
}

// Short rows of any alignment (101, 1001 columns): G = 8, 16 or 32 lanes per row, NE scalar loads
// per lane (a warp still reads one contiguous span), the row kept in registers: max, then the
// sum of exponentials against the row's final max, both trees inside the G lanes.
template <typename T, typename A, int G, int NE>
__global__ void __launch_bounds__(kBlock)
    lse_scalar_kernel(const T* __restrict__ x, long long rows, int cols, void* out, int out_dtype) {
  constexpr int kPerWarp = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  // (loop bound on blockIdx alone: the compiler then emits bare shuffles, not checked collectives)
  for (long long c0 = static_cast<long long>(blockIdx.x) * (kWarps * kPerWarp); c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * (kWarps * kPerWarp)) {
    const long long w0 = c0 + warp * kPerWarp;
    const long long row = w0 + sub;
    const bool live = row < rows;
    const T* __restrict__ r = x + row * cols;
    T raw[NE];
    int odd;
    short_row_load<T, G, NE>(r, cols, g, live, raw, &odd);
    A val[NE];
    A m = LseMath<A>::neg_inf();
#pragma unroll
    for (int k = 0; k < NE; ++k) {
      const int i = short_row_index<T, G, NE>(k, g, odd);
      val[k] = (live && i >= 0 && i < cols) ? Cvt<A, T>::to(raw[k]) : LseMath<A>::neg_inf();
    }
#pragma unroll
    for (int k = 0; k < NE; ++k) m = val[k] > m ? val[k] : m;
#pragma unroll
    for (int k = G / 2; k > 0; k >>= 1) {
      const A o = shfl_xor(m, k);
      m = o > m ? o : m;
    }
    const A shift = (m > LseMath<A>::neg_inf()) ? m : A(0);  // all -inf: every term exp(-inf) = 0
    const A bias = LseMath<A>::bias(shift);
    A s0 = A(0), s1 = A(0);
#pragma unroll
    for (int k = 0; k < NE; ++k) {
      const A t = LseMath<A>::term(val[k], bias);
      if (k & 1)
        s1 += t;
      else
        s0 += t;
    }
    A sum = s0 + s1;
#pragma unroll
    for (int k = G / 2; k > 0; k >>= 1) sum += shfl_xor(sum, k);
    if (live && g == 0) lse_put(out, out_dtype, row, m + LseMath<A>::log_(sum));
  }
}

// 16-bit rows of 32..1031 columns that are not whole aligned vectors: G = 8 or 16 lanes per row,
// 32 / G rows side by side in a warp; a lane loads one head element, NV 16-byte vectors from the
// first aligned address and one tail element together, keeps them in registers for the max sweep
// and the exp-sum sweep (the 2-byte-load kernel read these at 1.6-2.7 TB/s).
template <typename T, typename A, int G, int NV>
__global__ void __launch_bounds__(kBlock, 4)  // (64 registers: an explicit bound of 1 let the compiler take 86)
    lse_peel_kernel(const T* __restrict__ x, long long rows, int cols, void* out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  static_assert(G >= VE - 1, "a lane per possible head / tail element");
  constexpr int kPerWarp = 32 / G;
  constexpr int kPerCta = kWarps * kPerWarp;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  // (loop bound on blockIdx alone: bare shuffles; 32-bit row arithmetic on a CTA-uniform base:
  // the kernel is bound by instruction issue, not by memory)
  for (long long c0 = static_cast<long long>(blockIdx.x) * kPerCta; c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * kPerCta) {
    const T* __restrict__ xb = x + c0 * cols;
    const unsigned xlow = static_cast<unsigned>(reinterpret_cast<uintptr_t>(xb)) & 15u;
    const long long left = rows - c0;
    const int lr = warp * kPerWarp + sub;
    const bool live = lr < (left < kPerCta ? static_cast<int>(left) : kPerCta);
    const unsigned off = static_cast<unsigned>(lr) * static_cast<unsigned>(cols);
    const T* __restrict__ r = xb + off;
    int head = static_cast<int>(((16u - ((xlow + off * static_cast<unsigned>(sizeof(T))) & 15u)) & 15u) / sizeof(T));
    if (head > cols) head = cols;
    const int nvec = (cols - head) / VE;
    const int tail0 = head + nvec * VE;
    const V16* __restrict__ vp = reinterpret_cast<const V16*>(r + head);
    const bool has_h = live && g < head, has_t = live && tail0 + g < cols;
    A hv = LseMath<A>::neg_inf(), tv = LseMath<A>::neg_inf();
    if (has_h) hv = Cvt<A, T>::to(ldg_scalar(r + g));
    if (has_t) tv = Cvt<A, T>::to(ldg_scalar(r + tail0 + g));
    V16 buf[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = g + k * G;
      if (live && v < nvec) buf[k] = ldg_stream16(vp + v);
    }
    A m = hv > tv ? hv : tv;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = g + k * G;
      if (live && v < nvec) {
        A val[VE];
        if constexpr (sizeof(T) != 2) unpack_vec<T, A>(buf[k], val);
        const A vm = LseVecMax<T, A>::of(buf[k], val);  // packed max on 16-bit pairs
        m = vm > m ? vm : m;
      }
    }
#pragma unroll
    for (int k = G / 2; k > 0; k >>= 1) {
      const A o = shfl_xor(m, k);
      m = o > m ? o : m;
    }
    const A shift = (m > LseMath<A>::neg_inf()) ? m : A(0);  // all -inf: every term exp(-inf) = 0
    const A bias = LseMath<A>::bias(shift);
    A s0 = LseMath<A>::term(hv, bias), s1 = LseMath<A>::term(tv, bias);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = g + k * G;
      if (live && v < nvec) {
        A val[VE];
        unpack_vec<T, A>(buf[k], val);
#pragma unroll
        for (int e = 0; e < VE; e += 2) {
          s0 += LseMath<A>::term(val[e], bias);
          s1 += LseMath<A>::term(val[e + 1], bias);
        }
      }
    }
    A sum = s0 + s1;
#pragma unroll
    for (int k = G / 2; k > 0; k >>= 1) sum += shfl_xor(sum, k);
    if (live && g == 0) lse_put(out, out_dtype, c0 + lr, m + LseMath<A>::log_(sum));
  }
}

// logsumexp consumer of rows_staged_kernel (rows_staged.cuh): a max sweep and an exp-sum sweep over
// the thread's own row in shared memory.
template <typename T, typename A>
struct StagedLse {
  struct Params {
    void* out;
    int dtype;
  };
  template <int L>
  static __device__ __forceinline__ void run(const T* __restrict__ sp, int n, int i0, int i1, int rot,
                                             bool live, long long row, int h, const Params& p) {
    A m[4] = {LseMath<A>::neg_inf(), LseMath<A>::neg_inf(), LseMath<A>::neg_inf(), LseMath<A>::neg_inf()};
    smem_visit<T>(sp, n, i0, i1, rot, [&](int, auto x, int slot) {
      // (fmax: one FMNMX where the compare / select pair is two instructions; a NaN still reaches
      // the result through the exponent sweep)
      m[slot % 4] = fmax(m[slot % 4], Cvt<A, decltype(x)>::to(x));
    });
    A mx = m[1] > m[0] ? m[1] : m[0];
    const A my = m[3] > m[2] ? m[3] : m[2];
    mx = my > mx ? my : mx;
    mx = group_all<L>(mx, [](A a, A b) { return b > a ? b : a; });
    const A shift = (mx > LseMath<A>::neg_inf()) ? mx : A(0);  // all -inf: every term exp(-inf) = 0
    const A bias = LseMath<A>::bias(shift);
    A s[4] = {A(0), A(0), A(0), A(0)};
    smem_visit<T>(sp, n, i0, i1, rot, [&](int, auto x, int slot) {
      s[slot % 4] += LseMath<A>::term(Cvt<A, decltype(x)>::to(x), bias);
    });
    const A sum = group_all<L>((s[0] + s[1]) + (s[2] + s[3]), [](A a, A b) { return a + b; });
    if (live && h == 0) lse_put(p.out, p.dtype, row, mx + LseMath<A>::log_(sum));
  }
};

template <typename T, typename A>
static int launch_lse(const b200phi_logsumexp_desc* d, cudaStream_t stream) {
  constexpr int VE = VecOf<T>::kElems;
  const T* x = static_cast<const T*>(d->x);
  const long long cols = d->cols, rows = d->rows;
  const bool aligned = (reinterpret_cast<uintptr_t>(x) % 16 == 0) && (cols % VE == 0);
  const long long row_bytes = cols * static_cast<long long>(sizeof(T));
  // Short rows want few threads per row (the combine costs two barriers once a row spans
  // warps): ~256 B per thread in batches of U = 8, so an 8 KiB row is one warp's work.  Rows
  // that need the whole CTA anyway stream in batches of U = 4, which keeps six CTAs per SM
  // resident.  Measured, bf16 [65536,4096]: G 32 6.09 TB/s, G 64 5.39, G 128 4.15, G 256 3.49;
  // bf16 [8192,32000] (G 256): U 4 5.98 TB/s, U 8 5.59.
  static const int env_u = getenv("B200PHI_LSE_U") ? atoi(getenv("B200PHI_LSE_U")) : 0;
  static const int knob_bpt = getenv("B200PHI_LSE_BPT") ? atoi(getenv("B200PHI_LSE_BPT")) : 256;
  int G = 32;
  while (G < kBlock && static_cast<long long>(G) * knob_bpt < row_bytes) G <<= 1;
  // (after the fold became branch-free: bf16 [8192,32000] U 4 6.20 TB/s, U 8 5.94; fp32 the other
  // way round, 6.88 / 7.01)
  const int knob_u =
      env_u ? env_u : ((G == kBlock && row_bytes > 64LL * kBlock && sizeof(T) == 2) ? 4 : 8);
  const long long nvec_row = cols / VE;
  static const bool no_multi = getenv("B200PHI_LSE_NO_MULTI") != nullptr;
  static const bool no_staged = getenv("B200PHI_NO_ROW_STAGED") != nullptr;
  if (!no_staged && !no_multi && !env_u && staged_rows_ok(rows, row_bytes, static_cast<int>(sizeof(T)), kStagedLse, reinterpret_cast<uintptr_t>(x) % 16 == 0)) {
    // short rows that are not whole aligned vectors: a thread per row out of shared memory
    typename StagedLse<T, A>::Params sp{d->out, d->dtype};
    return launch_rows_staged<T, StagedLse<T, A>>(x, rows, static_cast<int>(cols), sp, stream);
  }
  // aligned 16-bit rows of more than this many vectors (and at most 128) take the peel kernel
  // instead of lse_multi_kernel (B200PHI_LSE_PEEL_ALIGNED_MIN overrides)
  static const long long lse_peel_aligned_min =
      getenv("B200PHI_LSE_PEEL_ALIGNED_MIN") ? atoll(getenv("B200PHI_LSE_PEEL_ALIGNED_MIN")) : 16;  // (bf16 512 columns 5.5 -> 6.6 TB/s)
  // B200PHI_LSE_PEEL_MAX: longest row, in vectors, of the peel kernel (a warp per row from 129)
  static const long long lse_peel_max = getenv("B200PHI_LSE_PEEL_MAX") ? atoll(getenv("B200PHI_LSE_PEEL_MAX")) : 256;
  const bool lse_to_peel = sizeof(T) == 2 && nvec_row > lse_peel_aligned_min && nvec_row <= lse_peel_max;
  if (aligned && !lse_to_peel && !no_multi && !env_u && nvec_row <= 128 && rows >= 64) {
    // several rows per warp (lse_multi_kernel)
    // lanes per row G, vectors per lane NV, rounds R (R * NV = 8 loads in flight)
    const int LG = nvec_row <= 32 ? 8 : 16;
    const long long rows_per_warp = (32 / LG) * (nvec_row <= 64 ? 2 : 1);
    long long ctas = (rows + kWarps * rows_per_warp - 1) / (kWarps * rows_per_warp);
    const long long cap = static_cast<long long>(current_sm_count()) * 8 * 8;
    if (ctas > cap) ctas = cap;
    const unsigned gm = static_cast<unsigned>(ctas);
    if (nvec_row <= 32)
      lse_multi_kernel<T, A, 8, 4, 2><<<gm, kBlock, 0, stream>>>(x, rows, nvec_row, d->out, d->dtype);
    else if (nvec_row <= 64)
      lse_multi_kernel<T, A, 16, 4, 2><<<gm, kBlock, 0, stream>>>(x, rows, nvec_row, d->out, d->dtype);
    else
      lse_multi_kernel<T, A, 16, 8, 1><<<gm, kBlock, 0, stream>>>(x, rows, nvec_row, d->out, d->dtype);
    cudaError_t e0 = cudaGetLastError();
    if (e0 != cudaSuccess) return cuda_status(e0);
    return B200PHI_OK;
  }
  if constexpr (sizeof(T) == 2) {
    static const bool no_peel = getenv("B200PHI_LSE_NO_PEEL") != nullptr;
    static const int peel_min = getenv("B200PHI_LSE_PEEL_MIN") ? atoi(getenv("B200PHI_LSE_PEEL_MIN")) : 4;
    if ((!aligned || lse_to_peel) && !no_peel && !env_u && nvec_row >= peel_min && nvec_row <= lse_peel_max && rows >= 64) {
      const int lanes = nvec_row <= 64 ? 8 : (nvec_row <= 128 ? 16 : 32);
      long long pc = (rows + kWarps * (32 / lanes) - 1) / (kWarps * (32 / lanes));
      const long long cap = static_cast<long long>(current_sm_count()) * 8 * 8;
      if (pc > cap) pc = cap;
      const unsigned gp = static_cast<unsigned>(pc);
      const int c = static_cast<int>(cols);
      if (nvec_row <= 16)
        lse_peel_kernel<T, A, 8, 2><<<gp, kBlock, 0, stream>>>(x, rows, c, d->out, d->dtype);
      else if (nvec_row <= 32)
        lse_peel_kernel<T, A, 8, 4><<<gp, kBlock, 0, stream>>>(x, rows, c, d->out, d->dtype);
      else if (nvec_row <= 64)
        lse_peel_kernel<T, A, 8, 8><<<gp, kBlock, 0, stream>>>(x, rows, c, d->out, d->dtype);
      else if (nvec_row <= 128)
        lse_peel_kernel<T, A, 16, 8><<<gp, kBlock, 0, stream>>>(x, rows, c, d->out, d->dtype);
      else  // (a 16-vector form for 4..8 KiB rows measured 2.5 TB/s on 4097 bf16 columns, lse_rows_kernel 4.5)
        lse_peel_kernel<T, A, 32, 8><<<gp, kBlock, 0, stream>>>(x, rows, c, d->out, d->dtype);
      cudaError_t e2 = cudaGetLastError();
      if (e2 != cudaSuccess) return cuda_status(e2);
      return B200PHI_OK;
    }
  }
  // (measured: at 1001 columns the peeled vector kernel is faster -- 6.6 vs 2.4 TB/s fp32 -- so the
  // scalar form stops at 256 columns)
  if (!aligned && cols < 256 && rows >= 64 && !env_u) {
    const int c = static_cast<int>(cols);
#define B200PHI_LSE_SCALAR(GG, NN)                                                              \
  {                                                                                             \
    long long sc = (rows + kWarps * (32 / GG) - 1) / (kWarps * (32 / GG));                      \
    if (sc > static_cast<long long>(current_sm_count()) * 8 * 8) sc = static_cast<long long>(current_sm_count()) * 8 * 8;                                                 \
    lse_scalar_kernel<T, A, GG, NN><<<static_cast<unsigned>(sc), kBlock, 0, stream>>>(x, rows, c, d->out, d->dtype); \
  }
    if (cols < 128)
      B200PHI_LSE_SCALAR(8, 16)
    else if (cols < 256)
      B200PHI_LSE_SCALAR(16, 16)
    else if (cols <= 512)
      B200PHI_LSE_SCALAR(32, 16)
    else
      B200PHI_LSE_SCALAR(32, 32)
#undef B200PHI_LSE_SCALAR
    cudaError_t e1 = cudaGetLastError();
    if (e1 != cudaSuccess) return cuda_status(e1);
    return B200PHI_OK;
  }
  const long long rows_per_cta = kBlock / G;
  long long ctas = (rows + rows_per_cta - 1) / rows_per_cta;
  const long long cap = static_cast<long long>(current_sm_count()) * 8 * 16;  // grid-stride beyond that
  if (ctas > cap) ctas = cap;
  const unsigned gr = static_cast<unsigned>(ctas);
#define B200PHI_LSE(GG)                                                                              \
  if (!aligned)                                                                                      \
    lse_rows_kernel<T, A, GG, 4, false><<<gr, kBlock, 0, stream>>>(x, rows, cols, d->out, d->dtype); \
  else if (knob_u == 8)                                                                              \
    lse_rows_kernel<T, A, GG, 8, true><<<gr, kBlock, 0, stream>>>(x, rows, cols, d->out, d->dtype);  \
  else                                                                                               \
    lse_rows_kernel<T, A, GG, 4, true><<<gr, kBlock, 0, stream>>>(x, rows, cols, d->out, d->dtype);
  switch (G) {
    case 32:
      B200PHI_LSE(32) break;
    case 64:
      B200PHI_LSE(64) break;
    case 128:
      B200PHI_LSE(128) break;
    default:
      B200PHI_LSE(256) break;
  }
#undef B200PHI_LSE
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

int run_logsumexp(const b200phi_logsumexp_desc* d, cudaStream_t stream) {
  if (!d) return set_error(B200PHI_INVALID_ARGUMENT, "null descriptor");
  if (d->rows < 0 || d->cols < 0) return set_error(B200PHI_INVALID_ARGUMENT, "negative extent");
  if (d->rows == 0 || d->cols == 0)
    return set_error(B200PHI_INVALID_ARGUMENT, "The dims of Input(X) should be greater than 0.");
  if (!d->x || !d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  switch (d->dtype) {
    case B200PHI_FLOAT32:
      return launch_lse<float, float>(d, stream);
    case B200PHI_FLOAT64:
      return launch_lse<double, double>(d, stream);
    case B200PHI_BFLOAT16:
      return launch_lse<__nv_bfloat16, float>(d, stream);
    case B200PHI_FLOAT16:
      return launch_lse<__half, float>(d, stream);
    default:
      return set_error(B200PHI_UNSUPPORTED, "logsumexp: floating-point input only");
  }
}

}  // namespace b200phi

extern "C" int b200phi_logsumexp(const b200phi_logsumexp_desc* d, void* stream) {
  return b200phi::run_logsumexp(d, static_cast<cudaStream_t>(stream));
}
