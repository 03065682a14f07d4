// Mixed-precision add: out(fp32) = x(fp32) + widen(y), y bf16 or fp16 -- the master-gradient
// accumulation that phi::AddKernel routes to MultiPrecisionAddKernelImpl
// (kps/elementwise_kernel.cu:54-99; funcs::MultiPrecisionAddFunctor, elementwise_functor.h:45-50).
// One exactly-rounded fp32 add per element, so results are bit-identical to the reference.
//
// Same [rows, inner] walk as the broadcast engine.  One thread step covers the 8 outputs of one
// 16-byte vector of y: two 16-byte vectors of x in, two out.  An operand that is broadcast
// along the row is read once per row and splatted.
#include <string>

#include "../../include/b200phi.h"
#include "bcast_plan.h"
#include "prim.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();

struct MpAddParams {
  const float* x;
  const void* y;
  float* out;
  int nouter;
  long long oext[B200PHI_MAX_RANK];
  long long ostride[2][B200PHI_MAX_RANK];
  long long inner;
  int istride[2];
  long long chunk_v;  // 8-element groups per chunk
  long long chunks;
};

template <typename TY, int U>
__global__ void __launch_bounds__(kBlock) mpadd_rows_kernel(MpAddParams p) {
  const long long row = blockIdx.x / p.chunks, chunk = blockIdx.x % p.chunks;
  long long off[2] = {0, 0};
  {
    long long idx = row;
    for (int k = p.nouter - 1; k >= 0; --k) {
      const long long e = p.oext[k];
      const long long q = idx / e, rem = idx - q * e;
      off[0] += rem * p.ostride[0][k];
      off[1] += rem * p.ostride[1][k];
      idx = q;
    }
  }
  const float* __restrict__ xs = p.x + off[0];
  const TY* __restrict__ ys = static_cast<const TY*>(p.y) + off[1];
  float* __restrict__ dst = p.out + row * p.inner;
  const bool sx = p.istride[0] != 0, sy = p.istride[1] != 0;
  const float x0 = sx ? 0.f : *xs;
  const float y0 = sy ? 0.f : Cvt<float, TY>::to(ldg_scalar(ys));
  const bool aligned = (reinterpret_cast<uintptr_t>(dst) % 16 == 0) &&
                       (!sx || reinterpret_cast<uintptr_t>(xs) % 16 == 0) &&
                       (!sy || reinterpret_cast<uintptr_t>(ys) % 16 == 0);
  const long long ngrp = p.inner / 8;
  const long long g0 = chunk * p.chunk_v;
  const bool last_chunk = chunk == p.chunks - 1;
  if (aligned) {
    long long g1 = g0 + p.chunk_v;
    if (g1 > ngrp) g1 = ngrp;
    for (long long gb = g0 + threadIdx.x; gb < g1; gb += static_cast<long long>(U) * kBlock) {
      V16 xa[U], xb[U], yv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long gi = gb + static_cast<long long>(u) * kBlock;
        if (gi < g1) {
          if (sx) {
            xa[u] = ldg_stream16(reinterpret_cast<const V16*>(xs) + 2 * gi);
            xb[u] = ldg_stream16(reinterpret_cast<const V16*>(xs) + 2 * gi + 1);
          }
          if (sy) yv[u] = ldg_stream16(reinterpret_cast<const V16*>(ys) + gi);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long gi = gb + static_cast<long long>(u) * kBlock;
        if (gi < g1) {
          float yf[8], xf[8];
          if (sy) {
            unpack_vec<TY, float>(yv[u], yf);
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) yf[e] = y0;
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            xf[e] = sx ? __uint_as_float(xa[u].w[e]) : x0;
            xf[4 + e] = sx ? __uint_as_float(xb[u].w[e]) : x0;
          }
          V16 oa, ob;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            oa.w[e] = __float_as_uint(xf[e] + yf[e]);
            ob.w[e] = __float_as_uint(xf[4 + e] + yf[4 + e]);
          }
          stg_stream16(reinterpret_cast<V16*>(dst) + 2 * gi, oa);
          stg_stream16(reinterpret_cast<V16*>(dst) + 2 * gi + 1, ob);
        }
      }
    }
    if (last_chunk) {
      for (long long i = ngrp * 8 + threadIdx.x; i < p.inner; i += kBlock)
        dst[i] = (sx ? xs[i] : x0) + (sy ? Cvt<float, TY>::to(ldg_scalar(ys + i)) : y0);
    }
  } else {
    const long long e0 = g0 * 8;
    long long e1 = last_chunk ? p.inner : (g0 + p.chunk_v) * 8;
    if (e1 > p.inner) e1 = p.inner;
    for (long long i = e0 + threadIdx.x; i < e1; i += kBlock)
      dst[i] = (sx ? xs[i] : x0) + (sy ? Cvt<float, TY>::to(ldg_scalar(ys + i)) : y0);
  }
}

template <typename TY>
__global__ void __launch_bounds__(kBlock)
    mpadd_generic_kernel(const float* __restrict__ x, const TY* __restrict__ y, float* __restrict__ out,
                         int ndim, MpAddParams p, long long numel) {
  // p.oext / p.ostride hold ALL merged dims here (ndim of them)
  for (long long o = static_cast<long long>(blockIdx.x) * kBlock + threadIdx.x; o < numel;
       o += static_cast<long long>(gridDim.x) * kBlock) {
    long long ox = 0, oy = 0, idx = o;
    for (int k = ndim - 1; k >= 0; --k) {
      const long long e = p.oext[k];
      const long long q = idx / e, rem = idx - q * e;
      ox += rem * p.ostride[0][k];
      oy += rem * p.ostride[1][k];
      idx = q;
    }
    out[o] = x[ox] + Cvt<float, TY>::to(ldg_scalar(y + oy));
  }
}

template <typename TY>
static int launch_mpadd(const b200phi_mixed_add_desc* d, const BcastPlan& bp, cudaStream_t stream) {
  MpAddParams p;
  p.x = static_cast<const float*>(d->x.data);
  p.y = d->y.data;
  p.out = static_cast<float*>(d->out);
  if (bp.use_rows_kernel) {
    p.nouter = bp.ndim - 1;
    for (int k = 0; k < bp.ndim - 1; ++k) {
      p.oext[k] = bp.extent[k];
      p.ostride[0][k] = bp.stride[0][k];
      p.ostride[1][k] = bp.stride[1][k];
    }
    p.inner = bp.inner;
    p.istride[0] = bp.stride[0][bp.ndim - 1] ? 1 : 0;
    p.istride[1] = bp.stride[1][bp.ndim - 1] ? 1 : 0;
    const long long ngrp = (bp.inner + 7) / 8;
    const long long max_chunk = static_cast<long long>(kBlock) * 4;  // 4 groups = 32 outputs per thread
    p.chunks = (ngrp + max_chunk - 1) / max_chunk;
    if (p.chunks < 1) p.chunks = 1;
    p.chunk_v = (ngrp + p.chunks - 1) / p.chunks;
    const long long blocks = bp.rows * p.chunks;
    if (blocks > 0x7fffffffLL) return set_error(B200PHI_UNSUPPORTED, "mixed add: too many rows");
    mpadd_rows_kernel<TY, 4><<<static_cast<unsigned>(blocks), kBlock, 0, stream>>>(p);
  } else {
    for (int k = 0; k < bp.ndim; ++k) {
      p.oext[k] = bp.extent[k];
      p.ostride[0][k] = bp.stride[0][k];
      p.ostride[1][k] = bp.stride[1][k];
    }
    long long blocks = (bp.out_numel + kBlock - 1) / kBlock;
    if (blocks > static_cast<long long>(current_sm_count()) * 8 * 64) blocks = static_cast<long long>(current_sm_count()) * 8 * 64;
    mpadd_generic_kernel<TY><<<static_cast<unsigned>(blocks), kBlock, 0, stream>>>(
        p.x, static_cast<const TY*>(p.y), p.out, bp.ndim, p, bp.out_numel);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

int run_mixed_add(const b200phi_mixed_add_desc* d, cudaStream_t stream) {
  if (!d) return set_error(B200PHI_INVALID_ARGUMENT, "null descriptor");
  if (d->y_dtype != B200PHI_BFLOAT16 && d->y_dtype != B200PHI_FLOAT16)
    return set_error(B200PHI_INVALID_ARGUMENT,
                     "Unsupport x dtype:float32, y dtype for add(x, y) operation: y must be bfloat16 or float16");
  b200phi_elementwise_desc e = {};
  e.dtype = B200PHI_FLOAT32;
  e.arity = 2;
  e.axis = -1;
  e.in[0] = d->x;
  e.in[1] = d->y;
  e.out_rank = d->out_rank;
  for (int k = 0; k < B200PHI_MAX_RANK; ++k) e.out_dims[k] = d->out_dims[k];
  BcastPlan bp;
  std::string err;
  const int rc = plan_broadcast(&e, &bp, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  if (bp.out_numel == 0) return B200PHI_OK;
  if (!d->x.data || !d->y.data || !d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  if (d->y_dtype == B200PHI_BFLOAT16) return launch_mpadd<__nv_bfloat16>(d, bp, stream);
  return launch_mpadd<__half>(d, bp, stream);
}

}  // namespace b200phi

extern "C" int b200phi_mixed_add(const b200phi_mixed_add_desc* d, void* stream) {
  return b200phi::run_mixed_add(d, static_cast<cudaStream_t>(stream));
}
