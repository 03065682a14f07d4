// Row mean + variance in one read of x (config C4: bf16 [16384, 8192]).
// Replaces the six-launch composite behind paddle.var
// (python/paddle/tensor/stat.py:184-197: mean -> subtract -> pow -> sum -> divide,
// each a ReduceAnyKernel / VectorizedBroadcastKernel launch that re-reads a
// tensor-sized buffer) by one kernel: the row is pulled into registers with
// 16-byte streaming loads, the mean is formed, and sum((x-mean)^2) is taken from
// the same registers.  Rows that do not fit the register tile take the same two
// steps with the second pass served from L2.
#include <string>

#include "../../include/b200phi.h"
#include "prim.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();

template <typename Acc, int G>
__device__ __forceinline__ Acc group_sum(Acc v, Acc* smem) {
  return group_combine<Acc, SumOp<Acc>, G>(v, smem);
}

__device__ __forceinline__ void put_stat(void* p, int dtype, long long i, double v) {
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
    case B200PHI_BFLOAT16:
      reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(static_cast<float>(v));
      break;
    default:
      break;
  }
}

// G threads per row (32: one warp per row, 8 rows per CTA; 128: two rows per CTA; 256: one).
// Measured on C4 (16 KiB rows): G = 128 x 8 vectors reads 5.2 TB/s, G = 256 x 4 reads 4.6 TB/s.
// NV = 16-byte vectors each thread keeps in registers; the row must satisfy
// cols <= G*NV*VE, be 16-byte aligned and a multiple of VE (checked on the host).
template <typename T, typename Acc, int G, int NV>
__global__ void __launch_bounds__(kBlock)
    rowstats_reg_kernel(const T* __restrict__ x, long long rows, long long cols, Acc denom,
                        void* mean_out, void* var_out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  __shared__ Acc smem[kWarps];
  const int g = threadIdx.x % G;
  const long long rows_per_cta = kBlock / G;
  const long long nvec = cols / VE;
  for (long long row = static_cast<long long>(blockIdx.x) * rows_per_cta + threadIdx.x / G;
       row < rows; row += static_cast<long long>(gridDim.x) * rows_per_cta) {
    const V16* __restrict__ vp = reinterpret_cast<const V16*>(x + row * cols);
    V16 buf[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const long long v = g + static_cast<long long>(k) * G;
      if (v < nvec) buf[k] = ldg_stream16(vp + v);
    }
    Acc s[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const long long v = g + static_cast<long long>(k) * G;
      if (v < nvec) fold_vec<Acc, SumOp<Acc>, T, 4>(buf[k], s);
    }
    Acc sum = group_sum<Acc, G>((s[0] + s[1]) + (s[2] + s[3]), smem);
    const Acc mean = sum / static_cast<Acc>(cols);
    Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const long long v = g + static_cast<long long>(k) * G;
      if (v < nvec) {
        typename VecOf<T>::U w;
        w.raw = buf[k];
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const Acc dlt = Cvt<Acc, T>::to(w.e[e]) - mean;
          q[e % 4] += dlt * dlt;
        }
      }
    }
    Acc ss = group_sum<Acc, G>((q[0] + q[1]) + (q[2] + q[3]), smem);
    if (g == 0) {
      if (mean_out) put_stat(mean_out, out_dtype, row, static_cast<double>(mean));
      if (var_out) put_stat(var_out, out_dtype, row, static_cast<double>(ss / denom));
    }
  }
}

// Any cols / alignment: one CTA per row, two sweeps (the second is an L2 hit for
// rows up to tens of MB).
template <typename T, typename Acc>
__global__ void __launch_bounds__(kBlock)
    rowstats_sweep_kernel(const T* __restrict__ x, long long rows, long long cols, Acc denom,
                          void* mean_out, void* var_out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  __shared__ Acc smem[kWarps];
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* __restrict__ r = x + row * cols;
    const uintptr_t addr = reinterpret_cast<uintptr_t>(r);
    long long head = static_cast<long long>(((16 - (addr & 15)) & 15) / sizeof(T));
    if (head > cols) head = cols;
    const long long nvec = (cols - head) / VE;
    const long long tail0 = head + nvec * VE;
    const V16* __restrict__ vp = reinterpret_cast<const V16*>(r + head);
    Acc s[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
    if (threadIdx.x < head) s[0] += Cvt<Acc, T>::to(ldg_scalar(r + threadIdx.x));
    long long v = threadIdx.x;
    for (; v + 3LL * kBlock < nvec; v += 4LL * kBlock) {
      V16 b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) b[u] = ldg_stream16(vp + v + static_cast<long long>(u) * kBlock);
#pragma unroll
      for (int u = 0; u < 4; ++u) fold_vec<Acc, SumOp<Acc>, T, 4>(b[u], s);
    }
    for (; v < nvec; v += kBlock) {
      V16 b = ldg_stream16(vp + v);
      fold_vec<Acc, SumOp<Acc>, T, 4>(b, s);
    }
    for (long long i = tail0 + threadIdx.x; i < cols; i += kBlock) s[1] += Cvt<Acc, T>::to(ldg_scalar(r + i));
    const Acc sum = block_combine<Acc, SumOp<Acc>>((s[0] + s[1]) + (s[2] + s[3]), smem);
    const Acc mean = sum / static_cast<Acc>(cols);

    Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
    if (threadIdx.x < head) {
      const Acc dlt = Cvt<Acc, T>::to(ldg_scalar(r + threadIdx.x)) - mean;
      q[0] += dlt * dlt;
    }
    for (v = threadIdx.x; v < nvec; v += kBlock) {
      typename VecOf<T>::U w;
      w.raw = *reinterpret_cast<const V16*>(vp + v);  // plain load: served by L2
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        const Acc dlt = Cvt<Acc, T>::to(w.e[e]) - mean;
        q[e % 4] += dlt * dlt;
      }
    }
    for (long long i = tail0 + threadIdx.x; i < cols; i += kBlock) {
      const Acc dlt = Cvt<Acc, T>::to(ldg_scalar(r + i)) - mean;
      q[1] += dlt * dlt;
    }
    const Acc ss = block_combine<Acc, SumOp<Acc>>((q[0] + q[1]) + (q[2] + q[3]), smem);
    if (threadIdx.x == 0) {
      if (mean_out) put_stat(mean_out, out_dtype, row, static_cast<double>(mean));
      if (var_out) put_stat(var_out, out_dtype, row, static_cast<double>(ss / denom));
    }
  }
}

template <typename T, typename Acc>
static int launch_rowstats(const b200phi_rowstats_desc* d, int out_dtype, cudaStream_t stream) {
  constexpr int VE = VecOf<T>::kElems;
  const T* x = static_cast<const T*>(d->x);
  // paddle.var: n-1 when unbiased and n > 1, else n (stat.py:193-196)
  const long long n = d->cols;
  const Acc denom = static_cast<Acc>((d->unbiased && n > 1) ? n - 1 : n);
  const bool aligned = (reinterpret_cast<uintptr_t>(x) % 16 == 0) && (n % VE == 0);
  const long long nvec = n / VE;
  const int sms = current_sm_count();
  if (aligned && nvec <= 32 * 8 && d->rows >= static_cast<long long>(sms) * 16) {
    long long ctas = (d->rows + kWarps - 1) / kWarps;
    if (ctas > 0x7fffffffLL) ctas = 0x7fffffffLL;
    rowstats_reg_kernel<T, Acc, 32, 8><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, d->mean, d->var, out_dtype);
  } else if (aligned && nvec <= 128 * 8) {
    long long ctas = (d->rows + 1) / 2;
    if (ctas > 0x7fffffffLL) ctas = 0x7fffffffLL;
    rowstats_reg_kernel<T, Acc, 128, 8><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, d->mean, d->var, out_dtype);
  } else {
    long long ctas = d->rows > 0x7fffffffLL ? 0x7fffffffLL : d->rows;
    rowstats_sweep_kernel<T, Acc><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, d->mean, d->var, out_dtype);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

int run_rowstats(const b200phi_rowstats_desc* d, cudaStream_t stream) {
  if (!d) return set_error(B200PHI_INVALID_ARGUMENT, "null descriptor");
  if (d->rows < 0 || d->cols < 0) return set_error(B200PHI_INVALID_ARGUMENT, "negative extent");
  if (d->rows == 0 || d->cols == 0)
    return set_error(B200PHI_INVALID_ARGUMENT, "Tensor need be reduced must not empty.");
  if (!d->x || (!d->mean && !d->var)) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  int out_dtype = d->out_dtype == B200PHI_UNDEFINED ? d->in_dtype : d->out_dtype;
  const bool ok_out = out_dtype == d->in_dtype ||
                      (out_dtype == B200PHI_FLOAT32 &&
                       (d->in_dtype == B200PHI_BFLOAT16 || d->in_dtype == B200PHI_FLOAT16));
  if (!ok_out) return set_error(B200PHI_UNSUPPORTED, "rowstats: output dtype not built");
  switch (d->in_dtype) {
    case B200PHI_FLOAT32:
      return launch_rowstats<float, float>(d, out_dtype, stream);
    case B200PHI_FLOAT64:
      return launch_rowstats<double, double>(d, out_dtype, stream);
    case B200PHI_BFLOAT16:
      return launch_rowstats<__nv_bfloat16, float>(d, out_dtype, stream);
    case B200PHI_FLOAT16:
      return launch_rowstats<__half, float>(d, out_dtype, stream);
    default:
      return set_error(B200PHI_UNSUPPORTED, "rowstats: floating-point input only");
  }
}

}  // namespace b200phi
