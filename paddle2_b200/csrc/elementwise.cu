// n-ary broadcast / elementwise engine with a fused unary epilogue, sm_100a.
// Replaces VectorizedBroadcastKernel (broadcast_function.h:284-435),
// VectorizedElementwiseKernel (elementwise_base.h:690-722) and the 2018-era
// FusedElemwiseAndActBroadcast{1,2}CUDAKernel (elementwise_op_function.h:203-350,
// grid = n = 64 CTAs for config C2).
//
// bcast_rows_kernel: the output is viewed as [rows, inner] with `inner` the
// innermost merged dim.  A CTA owns one chunk of one row; the per-input base
// offset of the row is unravelled once per CTA (not per element), then the inner
// loop is pure 16-byte loads -> functor -> 16-byte streaming stores.  An input
// that is broadcast along `inner` is read once per row and splatted.
// bcast_flat_kernel: the output as one flat array of 16-byte vectors (or elements), for
// outputs whose rows are short or ragged (feature maps, hidden sizes, [10,3,1,4] + [10,1,12,1]).
#include <stdlib.h>

#include <algorithm>
#include <type_traits>
#include <string>

#include "../../include/b200phi.h"
#include "bcast_plan.h"
#include "prim.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();
int run_elementwise(const b200phi_elementwise_desc* d, cudaStream_t stream);

// ---- compute type ------------------------------------------------------------------
template <typename T>
struct ComputeT {
  using type = T;
};
template <>
struct ComputeT<__nv_bfloat16> {
  using type = float;
};
template <>
struct ComputeT<__half> {
  using type = float;
};
template <>
struct ComputeT<int16_t> {
  using type = int;
};
template <>
struct ComputeT<int8_t> {
  using type = int;
};
template <>
struct ComputeT<uint8_t> {
  using type = int;
};
template <>
struct ComputeT<Bool8> {
  using type = int;
};

template <typename T, typename C>
__device__ __forceinline__ T from_compute(C v) {
  return static_cast<T>(v);
}
template <>
__device__ __forceinline__ __nv_bfloat16 from_compute<__nv_bfloat16, float>(float v) {
  return __float2bfloat16_rn(v);
}
template <>
__device__ __forceinline__ __half from_compute<__half, float>(float v) {
  return __float2half_rn(v);
}
template <>
__device__ __forceinline__ Bool8 from_compute<Bool8, int>(int v) {
  Bool8 b;
  b.v = v != 0 ? 1 : 0;
  return b;
}

// ---- functors --------------------------------------------------------------------------
template <typename C>
__device__ __forceinline__ C binary_apply(int op, C a, C b) {
  switch (op) {
    case B200PHI_ADD:
      return a + b;
    case B200PHI_SUBTRACT:
      return a - b;
    case B200PHI_MULTIPLY:
      return a * b;
    case B200PHI_DIVIDE:
      if constexpr (std::numeric_limits<C>::is_integer) {
        return b == C(0) ? C(0) : a / b;
      } else {
        return a / b;
      }
    case B200PHI_MAXIMUM:
      return a > b ? a : b;
    default:
      return a < b ? a : b;
  }
}

template <typename C>
__device__ __forceinline__ C unary_apply(int act, C v, float scale) {
  if constexpr (std::numeric_limits<C>::is_integer) {
    return act == B200PHI_ACT_RELU ? (v > C(0) ? v : C(0)) : v;
  } else {
    switch (act) {
      case B200PHI_ACT_RELU:
        return v > C(0) ? v : C(0);
      case B200PHI_ACT_SCALE:
        return v * static_cast<C>(scale);
      case B200PHI_ACT_GELU:
        return C(0.5) * v * (C(1) + erf(v * C(0.70710678118654752440)));
      case B200PHI_ACT_SIGMOID:
        return C(1) / (C(1) + exp(-v));
      case B200PHI_ACT_TANH:
        return tanh(v);
      default:
        return v;
    }
  }
}

// compile-time functor for the hot combinations
template <int kBinary, int kAct>
struct StaticF {
  static constexpr bool kInter = false;  // never asked for save_intermediate_out
  template <typename C>
  __device__ __forceinline__ C pre(C a, C b, C c, int arity) const {
    return binary_apply<C>(kBinary, a, b);
  }
  template <typename C>
  __device__ __forceinline__ C post(C t) const {
    return unary_apply<C>(kAct, t, 0.f);
  }
};

// everything else: functor ids are kernel arguments (uniform branches)
struct RuntimeF {
  static constexpr bool kInter = true;
  int binary, act, compound;
  float scale;
  template <typename C>
  __device__ __forceinline__ C pre(C a, C b, C c, int arity) const {
    if (arity == 1) return a;
    if (compound) return binary_apply<C>(binary, a, unary_apply<C>(act, b, scale));
    C t = binary_apply<C>(binary, a, b);
    if (arity == 3) t = binary_apply<C>(binary, t, c);
    return t;
  }
  template <typename C>
  __device__ __forceinline__ C post(C t) const {
    return compound ? t : unary_apply<C>(act, t, scale);
  }
};

// ---- kernels ----------------------------------------------------------------------------
template <int Arity>
struct RowsParams {
  const void* in[Arity];
  void* out;
  void* inter;  // optional second output: pre() result (save_intermediate_out)
  int nouter;
  long long oext[B200PHI_MAX_RANK];
  long long ostride[Arity][B200PHI_MAX_RANK];
  long long inner;
  int istride[Arity];  // 0 or 1
  long long chunk_v;   // 16-byte vectors per chunk
  long long chunks;    // chunks per row
};

// kMask >= 0: bit i set <=> input i is streamed along the row (stride 1), clear <=> it is
// one value per row (stride 0); known at compile time so only streamed inputs get load
// buffers (with a run-time flag every input got one and the kernel needed 113 registers,
// i.e. 2 CTAs per SM).  kMask == -1: decided at run time (the RuntimeF / ternary path).
template <typename T, typename F, int Arity, int kMask, int U>
__global__ void __launch_bounds__(kBlock) bcast_rows_kernel(RowsParams<Arity> p, F f) {
  using C = typename ComputeT<T>::type;
  constexpr int VE = VecOf<T>::kElems;
  const long long row = blockIdx.x / p.chunks, chunk = blockIdx.x % p.chunks;
  auto streamed = [&](int i) -> bool {
    if constexpr (kMask >= 0) return (kMask >> i) & 1;
    return p.istride[i] != 0;
  };

  // per-row base offsets (uniform over the CTA)
  const T* src[Arity];
  {
    long long off[Arity];
#pragma unroll
    for (int i = 0; i < Arity; ++i) off[i] = 0;
    long long idx = row;
    for (int k = p.nouter - 1; k >= 0; --k) {
      const long long e = p.oext[k];
      long long q, rem;
      if ((static_cast<unsigned long long>(idx) | static_cast<unsigned long long>(e)) >> 32) {
        q = idx / e;
        rem = idx - q * e;
      } else {
        const unsigned qi = static_cast<unsigned>(idx) / static_cast<unsigned>(e);
        q = qi;
        rem = static_cast<unsigned>(idx) - qi * static_cast<unsigned>(e);
      }
#pragma unroll
      for (int i = 0; i < Arity; ++i) off[i] += rem * p.ostride[i][k];
      idx = q;
    }
#pragma unroll
    for (int i = 0; i < Arity; ++i) src[i] = static_cast<const T*>(p.in[i]) + off[i];
  }
  T* __restrict__ dst = static_cast<T*>(p.out) + row * p.inner;
  T* __restrict__ dst2 = nullptr;
  if constexpr (F::kInter) dst2 = p.inter ? static_cast<T*>(p.inter) + row * p.inner : nullptr;

  C splat[Arity];
  bool aligned = (reinterpret_cast<uintptr_t>(dst) % 16 == 0) &&
                 (!dst2 || reinterpret_cast<uintptr_t>(dst2) % 16 == 0);
#pragma unroll
  for (int i = 0; i < Arity; ++i) {
    splat[i] = C(0);
    if (!streamed(i)) {
      splat[i] = Cvt<C, T>::to(ldg_scalar(src[i]));
    } else {
      aligned = aligned && (reinterpret_cast<uintptr_t>(src[i]) % 16 == 0);
    }
  }

  const long long nvec_row = p.inner / VE;
  const long long v0 = chunk * p.chunk_v;
  const bool last_chunk = (chunk == p.chunks - 1);

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
          typename VecOf<T>::U w[Arity], o, o2;
#pragma unroll
          for (int i = 0; i < Arity; ++i)
            if (streamed(i)) w[i].raw = raw[i][u];
#pragma unroll
          for (int e = 0; e < VE; ++e) {
            C a[3] = {C(0), C(0), C(0)};
#pragma unroll
            for (int i = 0; i < Arity; ++i) a[i] = streamed(i) ? Cvt<C, T>::to(w[i].e[e]) : splat[i];
            const C t = f.template pre<C>(a[0], a[1], a[2], Arity);
            o.e[e] = from_compute<T, C>(f.template post<C>(t));
            if constexpr (F::kInter) o2.e[e] = from_compute<T, C>(t);
          }
          stg_stream16(reinterpret_cast<V16*>(dst) + v, o.raw);
          if constexpr (F::kInter) {
            if (dst2) stg_stream16(reinterpret_cast<V16*>(dst2) + v, o2.raw);
          }
        }
      }
    }
    if (last_chunk) {
      for (long long i = nvec_row * VE + threadIdx.x; i < p.inner; i += kBlock) {
        C a[3] = {C(0), C(0), C(0)};
#pragma unroll
        for (int k = 0; k < Arity; ++k)
          a[k] = streamed(k) ? Cvt<C, T>::to(ldg_scalar(src[k] + i)) : splat[k];
        const C t = f.template pre<C>(a[0], a[1], a[2], Arity);
        dst[i] = from_compute<T, C>(f.template post<C>(t));
        if (dst2) dst2[i] = from_compute<T, C>(t);
      }
    }
  } else {
    const long long e0 = v0 * VE;
    long long e1 = last_chunk ? p.inner : (v0 + p.chunk_v) * VE;
    if (e1 > p.inner) e1 = p.inner;
    for (long long i = e0 + threadIdx.x; i < e1; i += kBlock) {
      C a[3] = {C(0), C(0), C(0)};
#pragma unroll
      for (int k = 0; k < Arity; ++k)
        a[k] = streamed(k) ? Cvt<C, T>::to(ldg_scalar(src[k] + i)) : splat[k];
      const C t = f.template pre<C>(a[0], a[1], a[2], Arity);
      dst[i] = from_compute<T, C>(f.template post<C>(t));
      if (dst2) dst2[i] = from_compute<T, C>(t);
    }
  }
}

// Flat walk for outputs whose rows are short (feature maps, hidden sizes, per-channel operands
// against [N,C,H*W]): the row kernel above spends a whole CTA on one row -- relu(x+y) on
// [42799,64,49]+[64,1] ran at 0.16 TB/s.  Here the output is one flat array of units (a 16-byte
// vector when every row is a whole number of aligned vectors, else one element; the element
// form is also the general fallback) and thread t of the grid takes units t, t+256, ... U at a
// time.  Per unit: one multiply-shift divide for (row, position in row); an operand with the
// output's own shape needs nothing more, an operand that is one shared row needs the position,
// any other operand unravels the row over the outer dims.  32-bit indices: the host slices
// outputs of 2^31 elements or more along the outermost dim.
template <int Arity>
struct FlatParams {
  const void* in[Arity];
  void* out;
  void* inter;
  int nouter;
  FastDiv oext[B200PHI_MAX_RANK];
  unsigned ostride[Arity][B200PHI_MAX_RANK];
  FastDiv upr;     // units per row
  unsigned total;  // units in this slice of the output
  int istride[Arity];
  int kind[Arity];  // 0: same shape as the output, 1: one row shared by all rows, 2: general
};

template <typename T, typename F, int Arity, bool kVec, int U>
__global__ void __launch_bounds__(kBlock) bcast_flat_kernel(FlatParams<Arity> p, F f) {
  using C = typename ComputeT<T>::type;
  constexpr int VE = kVec ? VecOf<T>::kElems : 1;
  bool need_row = false;
#pragma unroll
  for (int i = 0; i < Arity; ++i) need_row = need_row || p.kind[i] == 2;
  for (unsigned base = blockIdx.x * (kBlock * U) + threadIdx.x; base < p.total;
       base += gridDim.x * (kBlock * U)) {
    unsigned off[U][Arity];
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
          const T* src = static_cast<const T*>(p.in[i]) + off[u][i];
          if (p.istride[i])  // only a full-shape operand is read once; the others repeat across rows
            raw[u][i] = p.kind[i] == 0 ? ldg_stream16(src) : ldg_cached16(src);
          else
            splat[u][i] = Cvt<C, T>::to(ldg_scalar(src));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!live[u]) continue;
        typename VecOf<T>::U w[Arity], o, o2;
#pragma unroll
        for (int i = 0; i < Arity; ++i) w[i].raw = raw[u][i];
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          C a[3] = {C(0), C(0), C(0)};
#pragma unroll
          for (int i = 0; i < Arity; ++i) a[i] = p.istride[i] ? Cvt<C, T>::to(w[i].e[e]) : splat[u][i];
          const C t = f.template pre<C>(a[0], a[1], a[2], Arity);
          o.e[e] = from_compute<T, C>(f.template post<C>(t));
          if constexpr (F::kInter) o2.e[e] = from_compute<T, C>(t);
        }
        const unsigned id = base + u * kBlock;
        stg_stream16(reinterpret_cast<V16*>(p.out) + id, o.raw);
        if constexpr (F::kInter) {
          if (p.inter) stg_stream16(reinterpret_cast<V16*>(p.inter) + id, o2.raw);
        }
      }
    } else {
      C a[U][3];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        a[u][0] = a[u][1] = a[u][2] = C(0);
        if (!live[u]) continue;
#pragma unroll
        for (int i = 0; i < Arity; ++i)
          a[u][i] = Cvt<C, T>::to(ldg_scalar(static_cast<const T*>(p.in[i]) + off[u][i]));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!live[u]) continue;
        const unsigned id = base + u * kBlock;
        const C t = f.template pre<C>(a[u][0], a[u][1], a[u][2], Arity);
        static_cast<T*>(p.out)[id] = from_compute<T, C>(f.template post<C>(t));
        if constexpr (F::kInter) {
          if (p.inter) static_cast<T*>(p.inter)[id] = from_compute<T, C>(t);
        }
      }
    }
  }
}

// ---- launch -----------------------------------------------------------------------------
// Ragged rows (49- or 196-element feature maps, odd hidden sizes): a row is not a whole number
// of 16-byte vectors, but the output -- and every operand that has the output's shape -- is
// still one contiguous, aligned array.  Units are 16-byte vectors of that flat array; only the
// broadcast operands are gathered element by element (cached loads), each element's
// (row, position) stepped from the unit's first one.  Needs inner >= VE so that a unit touches
// at most two rows.  FlatParams::upr holds the ELEMENTS per row here, total the element count.
template <typename T, typename F, int Arity, int U>
__global__ void __launch_bounds__(kBlock) bcast_ragged_kernel(FlatParams<Arity> p, F f) {
  using C = typename ComputeT<T>::type;
  constexpr int VE = VecOf<T>::kElems;
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
      typename VecOf<T>::U w[Arity], o, o2;
#pragma unroll
      for (int i = 0; i < Arity; ++i) w[i].raw = raw[u][i];
      // operand values for the unit's VE elements; the operand's kind is tested once per unit,
      // not per element (with the test inside the element loop the kernel was instruction-bound)
      C vals[Arity][VE];
#pragma unroll
      for (int i = 0; i < Arity; ++i) {
        const T* src = static_cast<const T*>(p.in[i]);
        if (p.kind[i] == 0) {
#pragma unroll
          for (int e = 0; e < VE; ++e)
            vals[i][e] = whole ? Cvt<C, T>::to(w[i].e[e])
                               : (e0 + e < p.total ? Cvt<C, T>::to(ldg_scalar(src + e0 + e)) : C(0));
        } else if (p.kind[i] == 2 && !p.istride[i]) {
          // one value per row (a per-channel bias): at most two values per unit
          const C v0 = Cvt<C, T>::to(ldg_scalar(src + gen[0][i]));
          const C v1 = (col0 + VE > inner && (row0 + 1) * inner < p.total)
                           ? Cvt<C, T>::to(ldg_scalar(src + gen[1][i]))
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
            vals[i][e] = (whole || e0 + e < p.total) ? Cvt<C, T>::to(ldg_scalar(src + o_)) : C(0);
          }
        }
      }
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        const C a0 = vals[0][e];
        const C a1 = Arity > 1 ? vals[Arity > 1 ? 1 : 0][e] : C(0);
        const C a2 = Arity > 2 ? vals[Arity > 2 ? 2 : 0][e] : C(0);
        const C t = f.template pre<C>(a0, a1, a2, Arity);
        o.e[e] = from_compute<T, C>(f.template post<C>(t));
        if constexpr (F::kInter) o2.e[e] = from_compute<T, C>(t);
        if (!whole && e0 + e < p.total) {
          static_cast<T*>(p.out)[e0 + e] = o.e[e];
          if constexpr (F::kInter) {
            if (p.inter) static_cast<T*>(p.inter)[e0 + e] = o2.e[e];
          }
        }
      }
      if (whole) {
        stg_stream16(reinterpret_cast<V16*>(p.out) + id, o.raw);
        if constexpr (F::kInter) {
          if (p.inter) stg_stream16(reinterpret_cast<V16*>(p.inter) + id, o2.raw);
        }
      }
    }
  }
}

// The ragged case that matters -- a full-shape operand against ONE VALUE PER ROW (a per-channel
// bias on 7x7 / 14x14 feature maps: [N, C, 49] + [C, 1]) -- without any run-time operand-kind
// test and without a run-time index into the argument struct (both made bcast_ragged_kernel issue
// bound: bf16 3.3 TB/s, 80 % issue slots).  kScalarIs: which operand (0 / 1) is the per-row one.
// A thread owns U 16-byte units of the flat output, all U loads issued before the first use; per
// unit one divide finds (row, column), the row's value is looked up through L1 (at most two rows
// per unit since inner >= VE), and the functor runs on the VE elements with a select per element.
struct RaggedScalarParams {
  const void* full;    // the operand shaped like the output (16-byte aligned)
  const void* scalar;  // the per-row operand
  void* out;
  FastDiv inner;       // elements per row
  unsigned total;      // elements
  int nouter;          // <= 3 outer dims, right-aligned in the arrays below
  FastDiv oext[3];
  unsigned ostride[3];  // the scalar operand's element stride per outer dim
};

__device__ __forceinline__ unsigned ragged_row_offset(unsigned row, const RaggedScalarParams& p) {
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

template <typename T, typename F, int kScalarIs, int U>
__global__ void __launch_bounds__(kBlock)
    bcast_ragged_rowscalar_kernel(const __grid_constant__ RaggedScalarParams p, F f) {
  using C = typename ComputeT<T>::type;
  constexpr int VE = VecOf<T>::kElems;
  const unsigned inner = p.inner.d;
  const unsigned units = p.total / VE;  // whole units; the last partial one is handled below
  const T* __restrict__ sc = static_cast<const T*>(p.scalar);
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
      const C v0 = Cvt<C, T>::to(__ldg(sc + ragged_row_offset(row0, p)));
      C v1 = v0;
      if (col0 + VE > inner && (row0 + 1) * inner < p.total)
        v1 = Cvt<C, T>::to(__ldg(sc + ragged_row_offset(row0 + 1, p)));
      if (id < units) {
        typename VecOf<T>::U w, o;
        w.raw = raw[u];
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const C xv = Cvt<C, T>::to(w.e[e]);
          const C sv = (col0 + e >= inner) ? v1 : v0;
          const C t = kScalarIs == 1 ? f.template pre<C>(xv, sv, C(0), 2) : f.template pre<C>(sv, xv, C(0), 2);
          o.e[e] = from_compute<T, C>(f.template post<C>(t));
        }
        stg_stream16(reinterpret_cast<V16*>(p.out) + id, o.raw);
      } else {  // the trailing elements that do not fill a unit
        const T* __restrict__ fx = static_cast<const T*>(p.full);
        for (unsigned e = 0; e0 + e < p.total; ++e) {
          const C xv = Cvt<C, T>::to(ldg_scalar(fx + e0 + e));
          const C sv = (col0 + e >= inner) ? v1 : v0;
          const C t = kScalarIs == 1 ? f.template pre<C>(xv, sv, C(0), 2) : f.template pre<C>(sv, xv, C(0), 2);
          static_cast<T*>(p.out)[e0 + e] = from_compute<T, C>(f.template post<C>(t));
        }
      }
    }
  }
}

// The flat case that matters most -- a bias add: one operand has the output's shape, the other is
// ONE ROW shared by every row ([N, C] + [C]), whole 16-byte vectors.  No run-time operand-kind test
// and no per-element "streamed or splat" select (bcast_flat_kernel carries both: ncu counted 118
// instructions per 16-byte unit on bf16, 65 % of the issue slots at 5.8 TB/s); one multiply-shift
// divide per unit gives the position in the row.  kRowIs: which functor argument the row is.
struct FlatRowParams {
  const void* full;
  const void* row;
  void* out;
  FastDiv upr;     // vectors per row
  unsigned total;  // vectors in this slice of the output
};
template <typename T, typename F, int kRowIs, int U>
__global__ void __launch_bounds__(kBlock) bcast_flat_row_kernel(const __grid_constant__ FlatRowParams p, F f) {
  using C = typename ComputeT<T>::type;
  constexpr int VE = VecOf<T>::kElems;
  const V16* __restrict__ xf = static_cast<const V16*>(p.full);
  const V16* __restrict__ xr = static_cast<const V16*>(p.row);
  for (unsigned base = blockIdx.x * (kBlock * U) + threadIdx.x; base < p.total; base += gridDim.x * (kBlock * U)) {
    V16 a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      if (id < p.total) {
        const unsigned col = id - fast_div(id, p.upr) * p.upr.d;
        a[u] = ldg_stream16(xf + id);
        b[u] = ldg_cached16(xr + col);  // (the row repeats: L1 serves it)
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned id = base + u * kBlock;
      if (id >= p.total) continue;
      typename VecOf<T>::U wa, wb, o;
      wa.raw = a[u];
      wb.raw = b[u];
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        const C xv = Cvt<C, T>::to(wa.e[e]), rv = Cvt<C, T>::to(wb.e[e]);
        const C t = kRowIs == 1 ? f.template pre<C>(xv, rv, C(0), 2) : f.template pre<C>(rv, xv, C(0), 2);
        o.e[e] = from_compute<T, C>(f.template post<C>(t));
      }
      stg_stream16(static_cast<V16*>(p.out) + id, o.raw);
    }
  }
}

template <typename T, typename F, int Arity>
static int launch_flat(const b200phi_elementwise_desc* d, const BcastPlan& bp, F f, void* inter,
                       int mode, cudaStream_t stream) {  // mode 0: elements, 1: whole vectors, 2: ragged
  const bool vec = mode == 1;
  constexpr int VE = VecOf<T>::kElems;
  constexpr int U = 4;
  constexpr long long kLimit = 1LL << 31;  // the kernel indexes with 32 bits
  const int no = bp.ndim - 1;
  long long full[B200PHI_MAX_RANK];  // the output's own strides over the outer dims
  long long run = bp.inner;
  for (int k = no - 1; k >= 0; --k) {
    full[k] = run;
    run *= bp.extent[k];
  }
  const long long slab = no > 0 ? full[0] : bp.inner;  // elements per index of the outermost dim
  if (slab >= kLimit)
    return set_error(B200PHI_UNSUPPORTED, "elementwise: a short-row output this large is not built");
  const long long outer0 = no > 0 ? bp.extent[0] : 1;
  const long long per_slice = std::max<long long>(1, (kLimit - 1) / slab);
  if (mode == 2 && outer0 > per_slice) mode = 0;  // slices would lose the 16-byte alignment
  for (long long a = 0; a < outer0; a += per_slice) {
    const long long cnt = std::min(per_slice, outer0 - a);
    FlatParams<Arity> p;
    p.nouter = no;
    for (int k = 0; k < no; ++k) {
      p.oext[k] = make_fast_div(static_cast<unsigned>(k == 0 ? cnt : bp.extent[k]));
      for (int i = 0; i < Arity; ++i) p.ostride[i][k] = static_cast<unsigned>(bp.stride[i][k]);
    }
    const long long upr = vec ? bp.inner / VE : bp.inner;
    p.upr = make_fast_div(static_cast<unsigned>(upr));
    const long long rows = no > 0 ? cnt * (bp.rows / outer0) : 1;
    p.total = static_cast<unsigned>(rows * upr);
    for (int i = 0; i < Arity; ++i) {
      p.in[i] = static_cast<const T*>(d->in[i].data) + (no > 0 ? a * bp.stride[i][0] : 0);
      p.istride[i] = bp.stride[i][bp.ndim - 1] ? 1 : 0;
      bool same = p.istride[i] == 1, shared = p.istride[i] == 1;
      for (int k = 0; k < no; ++k) {
        same = same && bp.stride[i][k] == full[k];
        shared = shared && bp.stride[i][k] == 0;
      }
      p.kind[i] = same ? 0 : (shared ? 1 : 2);
    }
    p.out = static_cast<T*>(d->out) + a * slab;
    p.inter = inter ? static_cast<T*>(inter) + a * slab : nullptr;
    long long blocks = (static_cast<long long>(p.total) + kBlock * U - 1) / (kBlock * U);
    const long long cap = static_cast<long long>(current_sm_count()) * 8 * 32;  // grid-stride beyond that
    if (blocks > cap) blocks = cap;
    const unsigned g = static_cast<unsigned>(blocks);
    bool rowscalar_done = false;
    if constexpr (Arity == 2 && !F::kInter) {
      // full-shape operand against one value per row: the specialised kernel
      const int fi = p.kind[0] == 0 ? 0 : (p.kind[1] == 0 ? 1 : -1);
      const int si = fi < 0 ? -1 : 1 - fi;
      // (rows that are whole vectors take the same kernel: its units then never straddle two rows;
      // the general flat kernel unravels the row and tests the operand kinds per unit -- relu(x+y)
      // fp32 [10699,64,196]+[64,1] 5.75 TB/s)
      if ((mode == 2 || mode == 1) && fi >= 0 && p.kind[si] == 2 && p.istride[si] == 0 && no >= 1 && no <= 3 &&
          reinterpret_cast<uintptr_t>(p.in[fi]) % 16 == 0) {
        RaggedScalarParams rp;
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
          bcast_ragged_rowscalar_kernel<T, F, 1, RU><<<static_cast<unsigned>(rb), kBlock, 0, stream>>>(rp, f);
        else
          bcast_ragged_rowscalar_kernel<T, F, 0, RU><<<static_cast<unsigned>(rb), kBlock, 0, stream>>>(rp, f);
        rowscalar_done = true;
      }
    }
    if constexpr (Arity == 2 && !F::kInter) {
      // full-shape operand against one shared row (bias add), whole vectors: the specialised kernel
      const int fi = p.kind[0] == 0 ? 0 : (p.kind[1] == 0 ? 1 : -1);
      const int si = fi < 0 ? -1 : 1 - fi;
      static const bool no_flat_row = getenv("B200PHI_NO_FLAT_ROW") != nullptr;
      if (!rowscalar_done && vec && fi >= 0 && p.kind[si] == 1 && !no_flat_row) {
        FlatRowParams rp;
        rp.full = p.in[fi];
        rp.row = p.in[si];
        rp.out = p.out;
        rp.upr = p.upr;
        rp.total = p.total;
        if (si == 1)
          bcast_flat_row_kernel<T, F, 1, U><<<g, kBlock, 0, stream>>>(rp, f);
        else
          bcast_flat_row_kernel<T, F, 0, U><<<g, kBlock, 0, stream>>>(rp, f);
        rowscalar_done = true;
      }
    }
    if (rowscalar_done) {
    } else if (mode == 2) {
      long long rb = ((static_cast<long long>(p.total) + VE - 1) / VE + kBlock * 2 - 1) / (kBlock * 2);
      if (rb > cap) rb = cap;
      bcast_ragged_kernel<T, F, Arity, 2><<<static_cast<unsigned>(rb), kBlock, 0, stream>>>(p, f);
    } else if (vec) {
      bcast_flat_kernel<T, F, Arity, true, U><<<g, kBlock, 0, stream>>>(p, f);
    } else {
      bcast_flat_kernel<T, F, Arity, false, U><<<g, kBlock, 0, stream>>>(p, f);
    }
  }
  return B200PHI_OK;
}

template <typename T, typename F, int Arity>
static int launch_bcast(const b200phi_elementwise_desc* d, const BcastPlan& bp, F f, void* inter,
                        cudaStream_t stream) {
  constexpr int VE = VecOf<T>::kElems;
  constexpr bool kStaticMask = !F::kInter && Arity == 2;
  // Short rows go to the flat walk: as 16-byte vectors when every row is a whole number of
  // aligned vectors (at most 512 of them: 8 KiB; measured: 1024-vector rows with shared-row operands are faster on the row kernel), as single elements when rows are ragged and short.
  bool whole_vectors = bp.inner % VE == 0 && reinterpret_cast<uintptr_t>(d->out) % 16 == 0 &&
                       (!inter || reinterpret_cast<uintptr_t>(inter) % 16 == 0);
  for (int i = 0; i < Arity; ++i)
    if (bp.stride[i][bp.ndim - 1] != 0)
      whole_vectors = whole_vectors && reinterpret_cast<uintptr_t>(d->in[i].data) % 16 == 0;
  // a full-shape operand against one value per row ([N,C,H*W] + [C,1]) keeps its flat kernel up to
  // 1024 vectors per row (the row kernel's CTAs are short-lived there: 784-vector rows 5.8 TB/s)
  bool row_scalar_pattern = Arity == 2 && !F::kInter && bp.ndim >= 2 && bp.ndim <= 4;
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
  const bool flat_vec = whole_vectors && bp.rows > 1 && bp.inner / VE <= (row_scalar_pattern ? 1024 : 512);
  const bool flat_scalar = !flat_vec && bp.rows > 1 && !whole_vectors && bp.inner <= 1024;
  // ragged rows: the output and the full-shape operands as aligned vectors, the rest gathered
  bool flat_ragged = flat_scalar && bp.inner >= VE && reinterpret_cast<uintptr_t>(d->out) % 16 == 0 &&
                     (!inter || reinterpret_cast<uintptr_t>(inter) % 16 == 0);
  if (flat_ragged) {
    long long run = bp.inner;
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
        flat_ragged = flat_ragged && reinterpret_cast<uintptr_t>(d->in[i].data) % 16 == 0;
      }
    }
    (void)run;
    flat_ragged = flat_ragged && any_full;
  }
  if (bp.use_rows_kernel && !flat_vec && !flat_scalar) {
    RowsParams<Arity> p;
    for (int i = 0; i < Arity; ++i) p.in[i] = d->in[i].data;
    p.out = d->out;
    p.inter = inter;
    p.nouter = bp.ndim - 1;
    for (int k = 0; k < bp.ndim - 1; ++k) {
      p.oext[k] = bp.extent[k];
      for (int i = 0; i < Arity; ++i) p.ostride[i][k] = bp.stride[i][k];
    }
    p.inner = bp.inner;
    for (int i = 0; i < Arity; ++i) p.istride[i] = bp.stride[i][bp.ndim - 1] ? 1 : 0;
    int mask = 0;
    for (int i = 0; i < Arity; ++i) mask |= p.istride[i] << i;
    const long long nvec = (bp.inner + VE - 1) / VE;
    const long long max_chunk = static_cast<long long>(kBlock) * 8;  // 8 vectors per thread
    p.chunks = (nvec + max_chunk - 1) / max_chunk;
    if (p.chunks < 1) p.chunks = 1;
    p.chunk_v = (nvec + p.chunks - 1) / p.chunks;
    const long long blocks = bp.rows * p.chunks;
    if (blocks > 0x7fffffffLL) return set_error(B200PHI_UNSUPPORTED, "elementwise: too many rows");
    const unsigned g = static_cast<unsigned>(blocks);
    if constexpr (kStaticMask) {
      // two-input hot functors: the stream/splat pattern is a template argument
      if (mask == 1)
        bcast_rows_kernel<T, F, Arity, 1, 8><<<g, kBlock, 0, stream>>>(p, f);
      else if (mask == 2)
        bcast_rows_kernel<T, F, Arity, 2, 8><<<g, kBlock, 0, stream>>>(p, f);
      else if (mask == 3)
        bcast_rows_kernel<T, F, Arity, 3, 4><<<g, kBlock, 0, stream>>>(p, f);
      else
        bcast_rows_kernel<T, F, Arity, -1, 4><<<g, kBlock, 0, stream>>>(p, f);
    } else {
      bcast_rows_kernel<T, F, Arity, -1, 4><<<g, kBlock, 0, stream>>>(p, f);
    }
  } else {
    const int rc = launch_flat<T, F, Arity>(d, bp, f, inter, flat_vec ? 1 : (flat_ragged ? 2 : 0), stream);
    if (rc != B200PHI_OK) return rc;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

template <typename T>
static int launch_runtime(const b200phi_elementwise_desc* d, const BcastPlan& bp, void* inter,
                          cudaStream_t s) {
  RuntimeF f{d->binary, d->act, d->compound, d->scale};
  switch (d->arity) {
    case 1:
      return launch_bcast<T, RuntimeF, 1>(d, bp, f, inter, s);
    case 2:
      return launch_bcast<T, RuntimeF, 2>(d, bp, f, inter, s);
    default:
      return launch_bcast<T, RuntimeF, 3>(d, bp, f, inter, s);
  }
}

// hot = binary op without epilogue, or relu(x + y); two inputs; no second output
template <typename T>
static int launch_typed(const b200phi_elementwise_desc* d, const BcastPlan& bp, void* inter,
                        cudaStream_t s) {
  if (d->arity == 2 && !inter && d->compound == 0) {
    if (d->act == B200PHI_ACT_NONE) {
      switch (d->binary) {
        case B200PHI_ADD:
          return launch_bcast<T, StaticF<B200PHI_ADD, 0>, 2>(d, bp, {}, nullptr, s);
        case B200PHI_SUBTRACT:
          return launch_bcast<T, StaticF<B200PHI_SUBTRACT, 0>, 2>(d, bp, {}, nullptr, s);
        case B200PHI_MULTIPLY:
          return launch_bcast<T, StaticF<B200PHI_MULTIPLY, 0>, 2>(d, bp, {}, nullptr, s);
        case B200PHI_DIVIDE:
          return launch_bcast<T, StaticF<B200PHI_DIVIDE, 0>, 2>(d, bp, {}, nullptr, s);
        case B200PHI_MAXIMUM:
          return launch_bcast<T, StaticF<B200PHI_MAXIMUM, 0>, 2>(d, bp, {}, nullptr, s);
        default:
          return launch_bcast<T, StaticF<B200PHI_MINIMUM, 0>, 2>(d, bp, {}, nullptr, s);
      }
    }
    if (d->act == B200PHI_ACT_RELU && d->binary == B200PHI_ADD)
      return launch_bcast<T, StaticF<B200PHI_ADD, B200PHI_ACT_RELU>, 2>(d, bp, {}, nullptr, s);
    // the other unary functors of fused_elemwise_add_activation over x + y (bias + gelu ...): with
    // the functor ids as kernel arguments every element walks two switches (fp32 2.1 TB/s, bf16 0.9)
    if constexpr (std::is_floating_point<T>::value || std::is_same<T, __nv_bfloat16>::value ||
                  std::is_same<T, __half>::value) {
      if (d->binary == B200PHI_ADD) {
        if (d->act == B200PHI_ACT_GELU)
          return launch_bcast<T, StaticF<B200PHI_ADD, B200PHI_ACT_GELU>, 2>(d, bp, {}, nullptr, s);
        if (d->act == B200PHI_ACT_SIGMOID)
          return launch_bcast<T, StaticF<B200PHI_ADD, B200PHI_ACT_SIGMOID>, 2>(d, bp, {}, nullptr, s);
        if (d->act == B200PHI_ACT_TANH)
          return launch_bcast<T, StaticF<B200PHI_ADD, B200PHI_ACT_TANH>, 2>(d, bp, {}, nullptr, s);
      }
    }
  }
  return launch_runtime<T>(d, bp, inter, s);
}

// complex64 / complex128: add and subtract are component-wise, so the call is re-issued on the
// float lanes -- every operand and the output get a trailing dim of 2 (the (re, im) pair), and a
// lower-rank operand placed with the legacy `axis` offset is first padded with trailing 1s so
// that plain right-aligned broadcasting puts it where `axis` would (common_shape.h:35-92).
static int run_elementwise_complex(const b200phi_elementwise_desc* d, cudaStream_t stream) {
  if (!d) return set_error(B200PHI_INVALID_ARGUMENT, "null descriptor");
  if ((d->binary != B200PHI_ADD && d->binary != B200PHI_SUBTRACT) || d->act != B200PHI_ACT_NONE ||
      d->arity < 2 || d->intermediate_out != nullptr)
    return set_error(B200PHI_UNSUPPORTED, "complex dtypes: add and subtract only");
  if (d->out_rank >= B200PHI_MAX_RANK)
    return set_error(B200PHI_UNSUPPORTED, "complex dtypes: rank 9 is not built");
  b200phi_elementwise_desc f = *d;
  f.dtype = d->dtype == B200PHI_COMPLEX64 ? B200PHI_FLOAT32 : B200PHI_FLOAT64;
  f.axis = -1;
  int max_rank = 0;
  for (int i = 0; i < d->arity; ++i) max_rank = d->in[i].rank > max_rank ? d->in[i].rank : max_rank;
  for (int i = 0; i < d->arity; ++i) {
    b200phi_tensor_view& v = f.in[i];
    int rank = d->in[i].rank;
    if (i < 2 && d->axis >= 0 && rank < max_rank) {  // legacy offset: pad with trailing 1s
      const int trailing = max_rank - d->axis - rank;
      if (trailing < 0) return set_error(B200PHI_INVALID_ARGUMENT, "Broadcast dimension mismatch.");
      for (int k = 0; k < trailing; ++k) v.dims[rank++] = 1;
    }
    if (rank >= B200PHI_MAX_RANK) return set_error(B200PHI_UNSUPPORTED, "complex dtypes: rank 9 is not built");
    v.dims[rank++] = 2;
    v.rank = rank;
  }
  f.out_dims[f.out_rank++] = 2;
  return run_elementwise(&f, stream);
}

int run_elementwise(const b200phi_elementwise_desc* d, cudaStream_t stream) {
  if (d && (d->dtype == B200PHI_COMPLEX64 || d->dtype == B200PHI_COMPLEX128))
    return run_elementwise_complex(d, stream);
  BcastPlan bp;
  std::string err;
  int rc = plan_broadcast(d, &bp, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  if (d->binary < B200PHI_ADD || d->binary > B200PHI_MINIMUM)
    return set_error(B200PHI_INVALID_ARGUMENT, "unknown binary functor");
  if (d->act < B200PHI_ACT_NONE || d->act > B200PHI_ACT_TANH)
    return set_error(B200PHI_INVALID_ARGUMENT, "unknown unary functor");
  if (d->compound && (d->arity != 2 || d->act == B200PHI_ACT_NONE))
    return set_error(B200PHI_INVALID_ARGUMENT, "compound == 1 needs two inputs and a unary functor");
  if (bp.out_numel == 0) return B200PHI_OK;
  if (!d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null output pointer");
  for (int i = 0; i < d->arity; ++i)
    if (!d->in[i].data) return set_error(B200PHI_INVALID_ARGUMENT, "null input pointer");

  void* inter = d->intermediate_out;
  if (inter && d->compound) {
    // intermediate = act(in1), shaped like in1: a unary pass over in1 alone
    b200phi_elementwise_desc u = *d;
    u.arity = 1;
    u.compound = 0;
    u.in[0] = d->in[1];
    u.out_rank = d->in[1].rank;
    for (int k = 0; k < d->in[1].rank; ++k) u.out_dims[k] = d->in[1].dims[k];
    u.out = inter;
    u.intermediate_out = nullptr;
    u.axis = -1;
    rc = run_elementwise(&u, stream);
    if (rc != B200PHI_OK) return rc;
    inter = nullptr;
  }
  const bool is_float = d->dtype == B200PHI_FLOAT32 || d->dtype == B200PHI_FLOAT64 ||
                        d->dtype == B200PHI_FLOAT16 || d->dtype == B200PHI_BFLOAT16;
  if (!is_float && d->act != B200PHI_ACT_NONE && d->act != B200PHI_ACT_RELU)
    return set_error(B200PHI_UNSUPPORTED, "this unary functor needs a floating-point dtype");
  switch (d->dtype) {
    case B200PHI_FLOAT32:
      return launch_typed<float>(d, bp, inter, stream);
    case B200PHI_FLOAT64:
      return launch_typed<double>(d, bp, inter, stream);
    case B200PHI_BFLOAT16:
      return launch_typed<__nv_bfloat16>(d, bp, inter, stream);
    case B200PHI_FLOAT16:
      return launch_typed<__half>(d, bp, inter, stream);
    case B200PHI_INT32:
      return launch_typed<int32_t>(d, bp, inter, stream);
    case B200PHI_INT64:
      return launch_typed<long long>(d, bp, inter, stream);
    case B200PHI_INT16:
      return launch_runtime<int16_t>(d, bp, inter, stream);
    case B200PHI_INT8:
      return launch_runtime<int8_t>(d, bp, inter, stream);
    case B200PHI_UINT8:
      return launch_runtime<uint8_t>(d, bp, inter, stream);
    case B200PHI_BOOL:
      return launch_runtime<Bool8>(d, bp, inter, stream);
    default:
      return set_error(B200PHI_UNSUPPORTED, "elementwise: dtype not built");
  }
}

}  // namespace b200phi
