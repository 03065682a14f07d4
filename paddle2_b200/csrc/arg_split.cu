// argmax / argmin kernels that split the reduced extent across thread blocks -- what
// ComputeFullArg (paddle/phi/kernels/gpu/arg_min_max_kernel.cu:85-145) serves with one block per
// (pre, post) slice and scalar strided loads:
//   arg_row_split_kernel  post == 1, few long rows (paddle.argmax's default axis=None flattens the
//                         tensor into ONE row): the row is cut across CTAs
//   arg_col_split_kernel  post > 1 ([pre, n, post], e.g. the leading axis of a row-sharded tensor):
//                         16-byte loads along post, n cut across grid.y
// Both park a (value, index) partial per CTA in caller scratch, take a ticket, and the CTA that
// draws the last ticket combines the partials with the same rule as every other level --
// better value, else lower index -- which is associative and commutative, so the result does not
// depend on the order the CTAs finish in; tickets are reset in-kernel.
#include <string>

#include "arg_common.cuh"
#include "arg_plan.h"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);

template <typename V, typename Op>
__device__ __forceinline__ ArgPair<V> block_arg(ArgPair<V> p, V* s_v, long long* s_i) {
  p = warp_arg<V, Op>(p);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();  // the arrays may still be read by the previous round
  if (lane == 0) {
    s_v[warp] = p.v;
    s_i[warp] = p.i;
  }
  __syncthreads();
  ArgPair<V> r;
  r.v = s_v[0];
  r.i = s_i[0];
#pragma unroll
  for (int w = 1; w < kWarps; ++w) {
    ArgPair<V> q;
    q.v = s_v[w];
    q.i = s_i[w];
    r = Op::combine(r, q);
  }
  return r;
}

// grid = (splits, rows).  An all-identity segment reports (identity, its first index); among
// equal values the lowest index wins, so an all-identity row still answers 0.
template <typename T, typename V, bool kMax, bool kAligned>
__global__ void __launch_bounds__(kBlock)
    arg_row_split_kernel(const T* __restrict__ x, long long n, long long seg, V* __restrict__ pv,
                         long long* __restrict__ pi, unsigned* __restrict__ tickets, void* out,
                         int index_dtype, void* val_out) {
  using Op = ArgOp<V, kMax>;
  __shared__ V s_v[kWarps];
  __shared__ long long s_i[kWarps];
  __shared__ unsigned s_ticket;
  const long long row = blockIdx.y;
  const unsigned splits = gridDim.x;
  const long long c0 = static_cast<long long>(blockIdx.x) * seg;
  long long len = n - c0;
  if (len > seg) len = seg;
  ArgPair<V> p = Op::identity();
  if (len > 0) p = arg_span<T, V, kMax, kBlock, 8, kAligned>(x + row * n + c0, len, threadIdx.x);
  p.i += c0;
  p = block_arg<V, Op>(p, s_v, s_i);
  if (threadIdx.x == 0) {
    pv[row * splits + blockIdx.x] = p.v;
    pi[row * splits + blockIdx.x] = p.i;
    __threadfence();
    s_ticket = atomicAdd(&tickets[row], 1u);
  }
  __syncthreads();
  if (s_ticket != splits - 1) return;
  __threadfence();
  ArgPair<V> r = Op::identity();
  r.i = 0x7fffffffffffffffLL;  // loses every tie: only real partials can win
  for (unsigned s = threadIdx.x; s < splits; s += kBlock) {
    ArgPair<V> q;
    q.v = __ldcg(&pv[row * splits + s]);
    q.i = __ldcg(&pi[row * splits + s]);
    r = Op::combine(r, q);
  }
  r = block_arg<V, Op>(r, s_v, s_i);
  if (threadIdx.x == 0) {
    store_arg<T, V>(out, index_dtype, val_out, row, r);
    tickets[row] = 0;
  }
}

// Batches of UU row loads for the VE columns this thread owns; rows are visited in ascending
// order, so the strict compare keeps each column's first occurrence.
template <typename T, typename V, bool kMax, int VE, int UU>
__device__ __forceinline__ void arg_col_step(const unsigned char* __restrict__& p, size_t stride_bytes,
                                             long long& left, int& row, int TY, V (&bv)[VE], int (&bi)[VE]) {
  using Op = ArgOp<V, kMax>;
  for (; left >= UU; left -= UU) {
    if constexpr (VE > 1) {
      V16 buf[UU];
#pragma unroll
      for (int u = 0; u < UU; ++u) buf[u] = ldg_stream16(p + u * stride_bytes);
      p += UU * stride_bytes;
#pragma unroll
      for (int u = 0; u < UU; ++u) {
        typename VecOf<T>::U w;
        w.raw = buf[u];
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const V v = Cvt<V, T>::to(w.e[e]);
          if (Op::better(v, bv[e])) {
            bv[e] = v;
            bi[e] = row + u * TY;
          }
        }
      }
    } else {
      T buf[UU];
#pragma unroll
      for (int u = 0; u < UU; ++u) buf[u] = ldg_scalar(reinterpret_cast<const T*>(p + u * stride_bytes));
      p += UU * stride_bytes;
#pragma unroll
      for (int u = 0; u < UU; ++u) {
        const V v = Cvt<V, T>::to(buf[u]);
        if (Op::better(v, bv[0])) {
          bv[0] = v;
          bi[0] = row + u * TY;
        }
      }
    }
    row += UU * TY;
  }
}

// [pre, n, post] -> [pre, post].  Block = TX x TY: tx walks post in units of VE elements, ty walks
// n.  grid.x = pre * tiles, grid.y = splits of n.  Shared memory: [TY][TX][VE] values, then indices.
template <typename T, typename V, bool kMax, int VE, int U>
__global__ void __launch_bounds__(kBlock)
    arg_col_split_kernel(const T* __restrict__ x, long long n, long long post, int TX, long long tiles,
                         long long rows_per_split, V* __restrict__ pv, long long* __restrict__ pi,
                         unsigned* __restrict__ tickets, void* out, int index_dtype, void* val_out) {
  using Op = ArgOp<V, kMax>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  V* sv = reinterpret_cast<V*>(smem_raw);
  long long* si = reinterpret_cast<long long*>(smem_raw + static_cast<size_t>(kBlock) * VE * sizeof(V));
  __shared__ unsigned s_ticket;
  const int TY = kBlock / TX;
  const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
  const long long o = blockIdx.x / tiles, tile = blockIdx.x % tiles;
  const long long post_v = post / VE;
  const long long cv = tile * TX + tx;
  const bool live = cv < post_v;
  const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_split;
  long long r1 = r0 + rows_per_split;
  if (r1 > n) r1 = n;
  const unsigned splits = gridDim.y;

  V bv[VE];
  int bi[VE];  // row relative to r0
#pragma unroll
  for (int e = 0; e < VE; ++e) {
    bv[e] = Op::identity().v;
    bi[e] = 0;
  }
  if (live && r0 + ty < r1) {
    const unsigned char* __restrict__ p =
        reinterpret_cast<const unsigned char*>(x + (o * n + r0 + ty) * post + cv * VE);
    const size_t stride_bytes = static_cast<size_t>(TY) * post * sizeof(T);
    long long left = (r1 - r0 - ty + TY - 1) / TY;
    int row = ty;
    arg_col_step<T, V, kMax, VE, U>(p, stride_bytes, left, row, TY, bv, bi);
    if constexpr (U > 4) arg_col_step<T, V, kMax, VE, 4>(p, stride_bytes, left, row, TY, bv, bi);
    if constexpr (U > 2) arg_col_step<T, V, kMax, VE, 2>(p, stride_bytes, left, row, TY, bv, bi);
    if constexpr (U > 1) arg_col_step<T, V, kMax, VE, 1>(p, stride_bytes, left, row, TY, bv, bi);
  }

  // combine across ty (TY is a power of two)
  auto tree = [&](void) {
    __syncthreads();
    for (int half = TY >> 1; half > 0; half >>= 1) {
      if (ty < half) {
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          ArgPair<V> a, b;
          a.v = sv[(ty * TX + tx) * VE + e];
          a.i = si[(ty * TX + tx) * VE + e];
          b.v = sv[((ty + half) * TX + tx) * VE + e];
          b.i = si[((ty + half) * TX + tx) * VE + e];
          a = Op::combine(a, b);
          sv[(ty * TX + tx) * VE + e] = a.v;
          si[(ty * TX + tx) * VE + e] = a.i;
        }
      }
      __syncthreads();
    }
  };
#pragma unroll
  for (int e = 0; e < VE; ++e) {
    sv[(ty * TX + tx) * VE + e] = bv[e];
    si[(ty * TX + tx) * VE + e] = r0 + bi[e];
  }
  tree();

  const long long out_base = o * post + cv * VE;
  if (splits == 1) {
    if (ty == 0 && live) {
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        ArgPair<V> r;
        r.v = sv[tx * VE + e];
        r.i = si[tx * VE + e];
        store_arg<T, V>(out, index_dtype, val_out, out_base + e, r);
      }
    }
    return;
  }

  const long long out_numel = static_cast<long long>(gridDim.x / tiles) * post;
  if (ty == 0 && live) {
#pragma unroll
    for (int e = 0; e < VE; ++e) {
      pv[static_cast<long long>(blockIdx.y) * out_numel + out_base + e] = sv[tx * VE + e];
      pi[static_cast<long long>(blockIdx.y) * out_numel + out_base + e] = si[tx * VE + e];
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(&tickets[blockIdx.x], 1u);
  __syncthreads();
  if (s_ticket != splits - 1) return;
  __threadfence();

  // last CTA of this (pre, tile): combine the `splits` partials (they sit in L2)
  ArgPair<V> acc[VE];
#pragma unroll
  for (int e = 0; e < VE; ++e) {
    acc[e] = Op::identity();
    acc[e].i = 0x7fffffffffffffffLL;
  }
  if (live) {
    for (unsigned s = ty; s < splits; s += TY) {
      ArgPair<V> q[VE];
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        q[e].v = __ldcg(&pv[static_cast<long long>(s) * out_numel + out_base + e]);
        q[e].i = __ldcg(&pi[static_cast<long long>(s) * out_numel + out_base + e]);
      }
#pragma unroll
      for (int e = 0; e < VE; ++e) acc[e] = Op::combine(acc[e], q[e]);
    }
  }
#pragma unroll
  for (int e = 0; e < VE; ++e) {
    sv[(ty * TX + tx) * VE + e] = acc[e].v;
    si[(ty * TX + tx) * VE + e] = acc[e].i;
  }
  tree();
  if (ty == 0 && live) {
#pragma unroll
    for (int e = 0; e < VE; ++e) {
      ArgPair<V> r;
      r.v = sv[tx * VE + e];
      r.i = si[tx * VE + e];
      store_arg<T, V>(out, index_dtype, val_out, out_base + e, r);
    }
  }
  if (threadIdx.x == 0) tickets[blockIdx.x] = 0;
}

template <typename T, typename V, bool kMax>
static int launch_split_typed(const ArgPlan& p, const void* x_, void* out, void* val_out, void* ws,
                              cudaStream_t stream) {
  const T* x = static_cast<const T*>(x_);
  unsigned char* w = static_cast<unsigned char*>(ws);
  V* pv = reinterpret_cast<V*>(w);
  long long* pi = reinterpret_cast<long long*>(w + p.ws_val_bytes);
  unsigned* tickets = reinterpret_cast<unsigned*>(w + p.ws_val_bytes + p.ws_idx_bytes);
  if (p.ws_ticket_bytes) {  // caller scratch arrives uninitialised; tickets must start at zero
    cudaError_t e = cudaMemsetAsync(tickets, 0, p.ws_ticket_bytes, stream);
    if (e != cudaSuccess) return cuda_status(e);
  }
  if (p.kind == kArgRowSplit) {
    const dim3 grid(p.grid_x, p.grid_y);
    if (p.aligned)
      arg_row_split_kernel<T, V, kMax, true><<<grid, kBlock, 0, stream>>>(x, p.n, p.seg, pv, pi, tickets, out,
                                                                          p.index_dtype, val_out);
    else
      arg_row_split_kernel<T, V, kMax, false><<<grid, kBlock, 0, stream>>>(x, p.n, p.seg, pv, pi, tickets, out,
                                                                           p.index_dtype, val_out);
  } else {
    const dim3 grid(p.grid_x, p.grid_y);
    constexpr int VE = VecOf<T>::kElems;
    if (p.vec > 1) {
      const size_t smem = static_cast<size_t>(kBlock) * VE * (sizeof(V) + 8);
      if (smem + 64 > 48 * 1024) {  // 1-byte elements: 16 (value, index) pairs per thread
        cudaError_t ea = cudaFuncSetAttribute(arg_col_split_kernel<T, V, kMax, VE, 8>,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (ea != cudaSuccess) return cuda_status(ea);
      }
      arg_col_split_kernel<T, V, kMax, VE, 8><<<grid, kBlock, smem, stream>>>(
          x, p.n, p.post, p.TX, p.tiles, p.rows_per_split, pv, pi, tickets, out, p.index_dtype, val_out);
    } else {
      const size_t smem = static_cast<size_t>(kBlock) * (sizeof(V) + 8);
      arg_col_split_kernel<T, V, kMax, 1, 8><<<grid, kBlock, smem, stream>>>(
          x, p.n, p.post, p.TX, p.tiles, p.rows_per_split, pv, pi, tickets, out, p.index_dtype, val_out);
    }
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

template <typename T, typename V>
static int launch_split_op(int op, const ArgPlan& p, const void* x, void* out, void* val_out, void* ws,
                           cudaStream_t s) {
  if (op == B200PHI_ARGMAX) return launch_split_typed<T, V, true>(p, x, out, val_out, ws, s);
  return launch_split_typed<T, V, false>(p, x, out, val_out, ws, s);
}

// kArgRowSplit / kArgCol plans; `ws` holds p.ws_total_bytes (may be NULL when that is 0).
int launch_arg_split(int op, const ArgPlan& p, const void* x, void* out, void* val_out, void* ws,
                     cudaStream_t s) {
  switch (p.in_dtype) {
    case B200PHI_FLOAT32:
      return launch_split_op<float, float>(op, p, x, out, val_out, ws, s);
    case B200PHI_FLOAT64:
      return launch_split_op<double, double>(op, p, x, out, val_out, ws, s);
    case B200PHI_BFLOAT16:
      return launch_split_op<__nv_bfloat16, float>(op, p, x, out, val_out, ws, s);
    case B200PHI_FLOAT16:
      return launch_split_op<__half, float>(op, p, x, out, val_out, ws, s);
    case B200PHI_INT64:
      return launch_split_op<long long, long long>(op, p, x, out, val_out, ws, s);
    case B200PHI_INT32:
      return launch_split_op<int32_t, int32_t>(op, p, x, out, val_out, ws, s);
    case B200PHI_INT16:
      return launch_split_op<int16_t, int32_t>(op, p, x, out, val_out, ws, s);
    case B200PHI_UINT8:
      return launch_split_op<uint8_t, int32_t>(op, p, x, out, val_out, ws, s);
    default:
      return set_error(B200PHI_UNSUPPORTED, "argmax/argmin: dtype not built");
  }
}

}  // namespace b200phi
