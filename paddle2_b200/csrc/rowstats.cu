// Row mean + variance in one read of x (config C4: bf16 [16384, 8192]).
// Replaces the six-launch composite behind paddle.var
// (python/paddle/tensor/stat.py:184-197: mean -> subtract -> pow -> sum -> divide,
// each a ReduceAnyKernel / VectorizedBroadcastKernel launch that re-reads a
// tensor-sized buffer) by one kernel: the row is pulled into registers with
// 16-byte streaming loads, the mean is formed, and sum((x-mean)^2) is taken from
// the same registers.  Rows that do not fit the register tile take the same two
// steps with the second pass served from L2.
#include <stdlib.h>

#include <string>

#include "../../include/b200phi.h"
#include "prim.cuh"
#include "rows_staged.cuh"
#include "tma.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();

template <typename Acc, int G>
__device__ __forceinline__ Acc group_sum(Acc v, Acc* smem) {
  return group_combine<Acc, SumOp<Acc>, G>(v, smem);
}

__device__ __forceinline__ void put_stat(void* p, int dtype, long long i, float v) {
  switch (dtype) {
    case B200PHI_FLOAT32:
      reinterpret_cast<float*>(p)[i] = v;
      break;
    case B200PHI_FLOAT16:
      reinterpret_cast<__half*>(p)[i] = __float2half_rn(v);
      break;
    case B200PHI_BFLOAT16:
      reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
      break;
    default:
      break;
  }
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
// kPeel: rows of any alignment -- up to VE - 1 head elements before the first 16-byte boundary
// and up to VE - 1 tail elements after the last whole vector, one per thread (g < head, g < ntail),
// loaded together with the vectors and carried through both register sweeps.
// kFull: slots 0 .. kFull - 1 are held by every thread of the row (the host guarantees
// nvec >= kFull * G), so only NV - kFull slots carry a predicate.  With eight per-slot predicates
// next to the head / tail ones the compiler ran out of predicate registers and issued the last
// loads after the first loaded value had been consumed: two memory round trips per row.
template <typename T, typename Acc, int G, int NV, bool kPeel = false, int kFull = 0>
__global__ void __launch_bounds__(kBlock, kFull > 0 ? 4 : 1)  // (kFull: 64 registers, four CTAs per SM)
    rowstats_reg_kernel(const T* __restrict__ x, long long rows, long long cols64, Acc denom,
                        Acc cols_acc, void* mean_out, void* var_out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  constexpr int kW = G / 32;
  // [combine 0|1][parity][warp]: alternating parity per row leaves one barrier per combine
  __shared__ Acc smem[2][2][kWarps];
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long rows_per_cta = kBlock / G;
  // A row is at most G * NV vectors (+ head and tail): everything inside a row is 32-bit, and a
  // slot's liveness is one compare against nl (64-bit slot indices recomputed in the load loop and
  // in both sweeps: ~100 instructions per vector for ~40 of arithmetic on bf16).
  const int cols = static_cast<int>(cols64);
  int nvec = cols / VE;
  int parity = 0;
  for (long long row = static_cast<long long>(blockIdx.x) * rows_per_cta + grp; row < rows;
       row += static_cast<long long>(gridDim.x) * rows_per_cta) {
    const T* __restrict__ r = x + row * cols64;
    int head = 0;
    T hraw, traw;
    bool has_h = false, has_t = false;
    if constexpr (kPeel) {
      head = static_cast<int>(((16u - (static_cast<unsigned>(reinterpret_cast<uintptr_t>(r)) & 15u)) & 15u) / sizeof(T));
      if (head > cols) head = cols;
      nvec = (cols - head) / VE;
      const int tail0 = head + nvec * VE;
      has_h = g < head;
      has_t = tail0 + g < cols;
      if (has_h) hraw = ldg_scalar(r + g);
      if (has_t) traw = ldg_scalar(r + tail0 + g);
    }
    const V16* __restrict__ vp = reinterpret_cast<const V16*>(r + head) + g;
    const int nl = nvec > g ? (nvec - g + G - 1) / G : 0;  // this thread's live slots: 0 .. nl - 1
    V16 buf[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      if (k < kFull || k < nl) buf[k] = ldg_stream16(vp + k * G);
    }
    // (the head / tail elements are widened only now: their conversion ahead of the vector loads
    // made every warp wait for a 2-byte load before it issued the 16-byte ones)
    Acc hval = Acc(0), tval = Acc(0);
    if constexpr (kPeel) {
      if (has_h) hval = Cvt<Acc, T>::to(hraw);
      if (has_t) tval = Cvt<Acc, T>::to(traw);
    }
    Acc s[4] = {hval, tval, Acc(0), Acc(0)};
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      if (k < kFull || k < nl) {
        Acc val[VE];
        unpack_vec<T, Acc>(buf[k], val);
#pragma unroll
        for (int e = 0; e < VE; ++e) s[e % 4] += val[e];
      }
    }
    Acc sum = warp_combine<Acc, SumOp<Acc>>((s[0] + s[1]) + (s[2] + s[3]));
    if constexpr (G > 32) {
      if (lane == 0) smem[0][parity][warp] = sum;
      group_barrier<G>();
      sum = smem[0][parity][grp * kW];
#pragma unroll
      for (int w = 1; w < kW; ++w) sum += smem[0][parity][grp * kW + w];
    }
    const Acc mean = sum / cols_acc;
    Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
    if constexpr (kPeel) {
      if (has_h) q[0] = (hval - mean) * (hval - mean);
      if (has_t) q[1] = (tval - mean) * (tval - mean);
    }
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      if (k < kFull || k < nl) {
        Acc val[VE];
        unpack_vec<T, Acc>(buf[k], val);
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const Acc dlt = val[e] - mean;
          q[e % 4] += dlt * dlt;
        }
      }
    }
    Acc ss = warp_combine<Acc, SumOp<Acc>>((q[0] + q[1]) + (q[2] + q[3]));
    if constexpr (G > 32) {
      if (lane == 0) smem[1][parity][warp] = ss;
      group_barrier<G>();
      if (g == 0) {
        ss = smem[1][parity][grp * kW];
#pragma unroll
        for (int w = 1; w < kW; ++w) ss += smem[1][parity][grp * kW + w];
      }
    }
    parity ^= 1;
    if (g == 0) {
      if (mean_out) put_stat(mean_out, out_dtype, row, mean);
      if (var_out) put_stat(var_out, out_dtype, row, ss / denom);
    }
  }
}

// The same two sweeps for R rows at a time: rows shorter than G*8 vectors would leave most of the
// eight load slots of rowstats_reg_kernel empty (a 256-column bf16 row is one vector per lane:
// 1.2 TB/s; 4096 columns on G = 128 half fill it: 4.0 TB/s), so a group takes R rows with NV
// vectors per thread each, R * NV = 8 loads in flight, and the R sums share each shuffle step
// and each barrier.
template <typename T, typename Acc, int G, int NV, int R>
__global__ void __launch_bounds__(kBlock, 4)  // 64 registers: four CTAs per SM
    rowstats_multi_kernel(const T* __restrict__ x, long long rows, long long cols, Acc denom,
                          Acc cols_acc, void* mean_out, void* var_out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  constexpr int kW = G >= 32 ? G / 32 : 1;
  constexpr int kTree = G < 32 ? G / 2 : 16;  // shuffle tree inside the row's lanes
  constexpr int kGroups = kBlock / G;
  __shared__ Acc smem[2][2][R][kWarps];
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long nvec = cols / VE;
  int parity = 0;
  // Loop bound.  R == 1 forms: it depends on blockIdx alone (every group of the CTA runs the same
  // passes, dead rows predicated off), so the body is provably convergent and the shuffles compile
  // to bare SHFL; with a per-warp bound each one carried a convergence check and a collective
  // slow path (bf16 768 columns mean+var 5.47 -> 6.11 TB/s, 1024 columns 6.40 -> 6.95).
  // R == 2 forms keep the per-warp bound (the warp's FIRST group's row, so its lanes leave
  // together): measured faster there (bf16 256 columns: max 7.05 vs 6.66, mean+var 5.87 vs 5.10).
  constexpr bool kWarpLoop = (R == 2);
  const int first_grp = kWarpLoop ? (G < 32 ? (threadIdx.x >> 5) * (32 / (G < 32 ? G : 32)) : grp) : 0;
  const long long lag = static_cast<long long>(grp - first_grp) * R;
  for (long long w0 = (static_cast<long long>(blockIdx.x) * kGroups + first_grp) * R; w0 < rows;
       w0 += static_cast<long long>(gridDim.x) * kGroups * R) {
    const long long r0 = w0 + lag;
    V16 buf[R][NV];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const V16* __restrict__ vp = reinterpret_cast<const V16*>(x + (r0 + r) * cols);
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long v = g + static_cast<long long>(k) * G;
        if (r0 + r < rows && v < nvec) buf[r][k] = ldg_stream16(vp + v);
      }
    }
    Acc sum[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      Acc s[2] = {Acc(0), Acc(0)};
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long v = g + static_cast<long long>(k) * G;
        if (r0 + r < rows && v < nvec) {
          Acc val[VE];
          unpack_vec<T, Acc>(buf[r][k], val);
#pragma unroll
          for (int e = 0; e < VE; ++e) s[e & 1] += val[e];
        }
      }
      sum[r] = s[0] + s[1];
    }
#pragma unroll
    for (int m = kTree; m > 0; m >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) sum[r] += shfl_xor(sum[r], m);
    }
    if constexpr (G > 32) {
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) smem[0][parity][r][warp] = sum[r];
      }
      group_barrier<G>();
#pragma unroll
      for (int r = 0; r < R; ++r) {
        sum[r] = smem[0][parity][r][grp * kW];
#pragma unroll
        for (int w = 1; w < kW; ++w) sum[r] += smem[0][parity][r][grp * kW + w];
      }
    }
    Acc ss[R], mean[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      mean[r] = sum[r] / cols_acc;
      Acc q[2] = {Acc(0), Acc(0)};
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long v = g + static_cast<long long>(k) * G;
        if (r0 + r < rows && v < nvec) {
          Acc val[VE];
          unpack_vec<T, Acc>(buf[r][k], val);
#pragma unroll
          for (int e = 0; e < VE; ++e) {
            const Acc dlt = val[e] - mean[r];
            q[e & 1] += dlt * dlt;
          }
        }
      }
      ss[r] = q[0] + q[1];
    }
#pragma unroll
    for (int m = kTree; m > 0; m >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) ss[r] += shfl_xor(ss[r], m);
    }
    if constexpr (G > 32) {
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) smem[1][parity][r][warp] = ss[r];
      }
      group_barrier<G>();
      if (g == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          ss[r] = smem[1][parity][r][grp * kW];
#pragma unroll
          for (int w = 1; w < kW; ++w) ss[r] += smem[1][parity][r][grp * kW + w];
        }
      }
    }
    parity ^= 1;
    if (g == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        if (r0 + r < rows) {
          if (mean_out) put_stat(mean_out, out_dtype, r0 + r, mean[r]);
          if (var_out) put_stat(var_out, out_dtype, r0 + r, ss[r] / denom);
        }
      }
    }
  }
}

// Asynchronous-copy staging (LDGSTS): a persistent CTA owns rows blockIdx.x, +gridDim.x, ...;
// every thread copies ITS OWN NV vectors of the rows ahead into a private slot of a shared-memory
// ring with cp.async and later reads back only what it copied itself, so no barrier or mbarrier
// guards the data: cp.async.wait_group is enough.  kStages-1 rows per CTA are always in flight
// while the current row is unpacked ONCE into registers and both passes run from there.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <typename T, typename Acc, int NV, int kStages>
__global__ void __launch_bounds__(kBlock)
    rowstats_async_kernel(const T* __restrict__ x, long long rows, long long cols, Acc denom,
                          Acc cols_acc, void* mean_out, void* var_out, int out_dtype) {
  extern __shared__ __align__(128) unsigned char ring[];  // [kStages][NV][kBlock] V16
  __shared__ Acc part[2][2][kWarps];
  constexpr int VE = VecOf<T>::kElems;
  V16* stage = reinterpret_cast<V16*>(ring);
  const int nvec = static_cast<int>(cols / VE);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long my_rows = (rows > blockIdx.x) ? (rows - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  auto issue = [&](long long k) {
    if (k < my_rows) {
      const V16* src = reinterpret_cast<const V16*>(x + (blockIdx.x + k * static_cast<long long>(gridDim.x)) * cols);
      V16* dst = stage + static_cast<size_t>(k % kStages) * (NV * kBlock);
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int v = threadIdx.x + j * kBlock;
        if (v < nvec) cp_async16(dst + v, src + v);
      }
    }
    cp_async_commit();  // always: keeps the group count uniform
  };

#pragma unroll
  for (int s = 0; s < kStages - 1; ++s) issue(s);
  int parity = 0;
  for (long long k = 0; k < my_rows; ++k) {
    issue(k + kStages - 1);
    cp_async_wait<kStages - 1>();  // row k's copies (this thread's own) have landed
    const V16* cur = stage + static_cast<size_t>(k % kStages) * (NV * kBlock);
    Acc val[NV * VE];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int v = threadIdx.x + j * kBlock;
      if (v < nvec) {
        unpack_vec<T, Acc>(cur[v], &val[j * VE]);
      } else {
#pragma unroll
        for (int e = 0; e < VE; ++e) val[j * VE + e] = Acc(0);
      }
    }
    Acc a[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
#pragma unroll
    for (int i = 0; i < NV * VE; ++i) a[i % 4] += val[i];
    Acc sum = warp_combine<Acc, SumOp<Acc>>((a[0] + a[1]) + (a[2] + a[3]));
    if (lane == 0) part[0][parity][warp] = sum;
    __syncthreads();
    sum = part[0][parity][0];
#pragma unroll
    for (int w = 1; w < kWarps; ++w) sum += part[0][parity][w];
    const Acc mean = sum / cols_acc;
    Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      if (threadIdx.x + j * kBlock < nvec) {
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const Acc dlt = val[j * VE + e] - mean;
          q[e % 4] += dlt * dlt;
        }
      }
    }
    Acc ss = warp_combine<Acc, SumOp<Acc>>((q[0] + q[1]) + (q[2] + q[3]));
    if (lane == 0) part[1][parity][warp] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
      ss = part[1][parity][0];
#pragma unroll
      for (int w = 1; w < kWarps; ++w) ss += part[1][parity][w];
      const long long row = blockIdx.x + k * static_cast<long long>(gridDim.x);
      if (mean_out) put_stat(mean_out, out_dtype, row, mean);
      if (var_out) put_stat(var_out, out_dtype, row, ss / denom);
    }
    parity ^= 1;
  }
  cp_async_wait<0>();
}

template <typename T, typename Acc, int NV, int kStages>
static int launch_rowstats_async(const T* x, const b200phi_rowstats_desc* d, int out_dtype, Acc denom,
                                 int sms, cudaStream_t stream) {
  const size_t smem = static_cast<size_t>(kStages) * NV * kBlock * sizeof(V16);
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    cudaError_t ea = cudaFuncSetAttribute(rowstats_async_kernel<T, Acc, NV, kStages>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (ea != cudaSuccess) return cuda_status(ea);
    attr_set[dev] = true;
  }
  int per_sm = static_cast<int>((200 * 1024) / (smem + 1024));
  if (per_sm > 6) per_sm = 6;
  if (per_sm < 1) per_sm = 1;
  long long ctas = static_cast<long long>(sms) * per_sm;
  if (ctas > d->rows) ctas = d->rows;
  rowstats_async_kernel<T, Acc, NV, kStages><<<static_cast<unsigned>(ctas), kBlock, smem, stream>>>(
      x, d->rows, d->cols, denom, static_cast<Acc>(d->cols), d->mean, d->var, out_dtype);
  return B200PHI_OK;
}

// Short rows of any alignment (101, 1001 columns): the sweep kernel below spends a CTA per row
// (a 101-column row: 0.07-0.18 TB/s).  Here G = 8, 16 or 32 lanes take one row with NE scalar
// loads per lane (consecutive lanes read consecutive elements, consecutive groups consecutive
// rows: a warp still reads one contiguous span), keep the row in registers as Acc, and do both
// sweeps from there; the shuffle tree stays inside the G lanes.  cols <= G * NE.
template <typename T, typename Acc, int G, int NE>
__global__ void __launch_bounds__(kBlock)
    rowstats_scalar_kernel(const T* __restrict__ x, long long rows, int cols, Acc denom, Acc cols_acc,
                           void* mean_out, void* var_out, int out_dtype) {
  constexpr int kPerWarp = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  // warp-uniform loop bound: the shuffles use the full mask
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
    Acc val[NE];
#pragma unroll
    for (int k = 0; k < NE; ++k) {
      const int i = short_row_index<T, G, NE>(k, g, odd);
      val[k] = (live && i >= 0 && i < cols) ? Cvt<Acc, T>::to(raw[k]) : Acc(0);
    }
    Acc s0 = Acc(0), s1 = Acc(0);
#pragma unroll
    for (int k = 0; k < NE; k += 2) {
      s0 += val[k];
      if (k + 1 < NE) s1 += val[k + 1];
    }
    Acc sum = s0 + s1;
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) sum += shfl_xor(sum, m);
    const Acc mean = sum / cols_acc;
    Acc q0 = Acc(0), q1 = Acc(0);
#pragma unroll
    for (int k = 0; k < NE; ++k) {
      const int i = short_row_index<T, G, NE>(k, g, odd);
      const Acc dlt = (i >= 0 && i < cols) ? val[k] - mean : Acc(0);
      if (k & 1)
        q1 += dlt * dlt;
      else
        q0 += dlt * dlt;
    }
    Acc ss = q0 + q1;
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) ss += shfl_xor(ss, m);
    if (live && g == 0) {
      if (mean_out) put_stat(mean_out, out_dtype, row, mean);
      if (var_out) put_stat(var_out, out_dtype, row, ss / denom);
    }
  }
}

// Long rows in ONE sweep of global memory, any alignment.  Every thread keeps (n, mean, M2) and
// merges one 16-byte vector at a time with the pairwise update of Chan, Golub & LeVeque: the
// vector's own mean and sum of squared deviations are exact two-step quantities over its VE
// registers, and the merge only ever subtracts two means, so nothing cancels catastrophically
// (the accuracy is that of the two-sweep form, not of sum / sum-of-squares).  Threads and warps
// merge their triples the same way.  Replaces the two-sweep kernel below wherever a row does not
// fit the register kernels: its second sweep came from HBM again once the resident CTAs' rows
// outgrew L2 (fp32 [21365,50257]: 3.6 TB/s).
template <typename Acc>
struct Moments {
  Acc n, mean, m2;
};
template <typename Acc>
__device__ __forceinline__ void moments_merge(Moments<Acc>& a, const Moments<Acc>& b) {
  const Acc n = a.n + b.n;
  if (n > Acc(0)) {
    const Acc delta = b.mean - a.mean;
    const Acc fb = b.n / n;
    a.mean += delta * fb;
    a.m2 += b.m2 + delta * delta * a.n * fb;
    a.n = n;
  }
}
template <typename T, typename Acc>
__device__ __forceinline__ void moments_add_vec(Moments<Acc>& a, const V16& raw) {
  constexpr int VE = VecOf<T>::kElems;
  Acc val[VE];
  unpack_vec<T, Acc>(raw, val);
  Acc s0 = Acc(0), s1 = Acc(0);
#pragma unroll
  for (int e = 0; e < VE; e += 2) {
    s0 += val[e];
    s1 += val[e + 1];
  }
  const Acc vmean = (s0 + s1) * (Acc(1) / VE);
  Acc q0 = Acc(0), q1 = Acc(0);
#pragma unroll
  for (int e = 0; e < VE; e += 2) {
    const Acc d0 = val[e] - vmean, d1 = val[e + 1] - vmean;
    q0 += d0 * d0;
    q1 += d1 * d1;
  }
  const Acc n = a.n + Acc(VE);
  const Acc delta = vmean - a.mean;
  const Acc fb = Acc(VE) / n;
  a.mean += delta * fb;
  a.m2 += (q0 + q1) + delta * delta * a.n * fb;
  a.n = n;
}
template <typename Acc>
__device__ __forceinline__ void moments_add_one(Moments<Acc>& a, Acc v) {
  const Acc n = a.n + Acc(1);
  const Acc delta = v - a.mean;
  a.mean += delta / n;
  a.m2 += delta * (v - a.mean);
  a.n = n;
}

template <typename T, typename Acc, int U>
__global__ void __launch_bounds__(kBlock)
    rowstats_stream_kernel(const T* __restrict__ x, long long rows, long long cols, Acc denom,
                           void* mean_out, void* var_out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  __shared__ Acc sm[3][kWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* __restrict__ r = x + row * cols;
    const uintptr_t addr = reinterpret_cast<uintptr_t>(r);
    long long head = static_cast<long long>(((16 - (addr & 15)) & 15) / sizeof(T));
    if (head > cols) head = cols;
    const long long nvec = (cols - head) / VE;
    const long long tail0 = head + nvec * VE;
    const V16* __restrict__ vp = reinterpret_cast<const V16*>(r + head);
    Moments<Acc> a;
    a.n = a.mean = a.m2 = Acc(0);
    if (threadIdx.x < head) moments_add_one<Acc>(a, Cvt<Acc, T>::to(ldg_scalar(r + threadIdx.x)));
    for (long long vb = threadIdx.x; vb < nvec; vb += static_cast<long long>(U) * kBlock) {
      V16 buf[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long v = vb + static_cast<long long>(u) * kBlock;
        if (v < nvec) buf[u] = ldg_stream16(vp + v);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long v = vb + static_cast<long long>(u) * kBlock;
        if (v < nvec) moments_add_vec<T, Acc>(a, buf[u]);
      }
    }
    for (long long i = tail0 + threadIdx.x; i < cols; i += kBlock)
      moments_add_one<Acc>(a, Cvt<Acc, T>::to(ldg_scalar(r + i)));
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) {
      Moments<Acc> b;
      b.n = shfl_xor(a.n, m);
      b.mean = shfl_xor(a.mean, m);
      b.m2 = shfl_xor(a.m2, m);
      moments_merge<Acc>(a, b);
    }
    if (lane == 0) {
      sm[0][warp] = a.n;
      sm[1][warp] = a.mean;
      sm[2][warp] = a.m2;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
      for (int w = 1; w < kWarps; ++w) {
        Moments<Acc> b;
        b.n = sm[0][w];
        b.mean = sm[1][w];
        b.m2 = sm[2][w];
        moments_merge<Acc>(a, b);
      }
      if (mean_out) put_stat(mean_out, out_dtype, row, a.mean);
      if (var_out) put_stat(var_out, out_dtype, row, a.m2 / denom);
    }
    __syncthreads();
  }
}

// Full rows of exactly G * NV vectors (config C4: 8192 bf16 = 128 x 8) in one sweep of the
// registers and ONE barrier: the pairwise-moments merge with every count known at compile time
// (after k vectors a thread holds k*VE elements; every lane and warp holds the same count), so
// the merges are a few multiply-adds with constant weights and only (mean, M2) travel.
template <typename T, typename Acc, int G, int NV>
__global__ void __launch_bounds__(kBlock, 4)
    rowstats_moments_kernel(const T* __restrict__ x, long long rows, long long cols, Acc denom,
                            void* mean_out, void* var_out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  constexpr int kW = G / 32;
  __shared__ Acc sm[2][2][kWarps];  // [mean | m2][parity][warp]
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long rows_per_cta = kBlock / G;
  int parity = 0;
  for (long long row = static_cast<long long>(blockIdx.x) * rows_per_cta + grp; row < rows;
       row += static_cast<long long>(gridDim.x) * rows_per_cta) {
    const V16* __restrict__ vp = reinterpret_cast<const V16*>(x + row * cols);
    V16 buf[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) buf[k] = ldg_stream16(vp + g + k * G);
    Acc mean = Acc(0), m2 = Acc(0);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      Acc val[VE];
      unpack_vec<T, Acc>(buf[k], val);
      Acc s0 = Acc(0), s1 = Acc(0);
#pragma unroll
      for (int e = 0; e < VE; e += 2) {
        s0 += val[e];
        s1 += val[e + 1];
      }
      const Acc vmean = (s0 + s1) * (Acc(1) / VE);
      Acc q0 = Acc(0), q1 = Acc(0);
#pragma unroll
      for (int e = 0; e < VE; e += 2) {
        const Acc d0 = val[e] - vmean, d1 = val[e + 1] - vmean;
        q0 += d0 * d0;
        q1 += d1 * d1;
      }
      // merge VE new elements into k*VE held ones: weights 1/(k+1) and k*VE/(k+1)
      const Acc delta = vmean - mean;
      mean += delta * (Acc(1) / (k + 1));
      m2 += (q0 + q1) + delta * delta * (Acc(k) * VE / (k + 1));
    }
    // lanes: equal counts c, 2c, 4c, ... : mean = average, M2 += delta^2 * c / 2
    Acc cnt = Acc(NV * VE);
#pragma unroll
    for (int m = 1; m < 32; m <<= 1) {
      const Acc om = shfl_xor(mean, m), oq = shfl_xor(m2, m);
      const Acc delta = om - mean;
      mean = (mean + om) * Acc(0.5);
      m2 += oq + delta * delta * (cnt * Acc(0.5));
      cnt += cnt;
    }
    if constexpr (G > 32) {
      if (lane == 0) {
        sm[0][parity][warp] = mean;
        sm[1][parity][warp] = m2;
      }
      group_barrier<G>();
    }
    if (g == 0) {
      Acc c = cnt;
#pragma unroll
      for (int w = 1; w < kW; ++w) {  // merge a warp (cnt elements) into w warps' worth
        const Acc om = sm[0][parity][grp * kW + w], oq = sm[1][parity][grp * kW + w];
        const Acc delta = om - mean;
        mean += delta * (Acc(1) / (w + 1));
        m2 += oq + delta * delta * (c * Acc(w) / (w + 1));
      }
      if (mean_out) put_stat(mean_out, out_dtype, row, mean);
      if (var_out) put_stat(var_out, out_dtype, row, m2 / denom);
    }
    parity ^= 1;
  }
}

// Any cols / alignment: one CTA per row, two sweeps (the second is an L2 hit for
// rows up to tens of MB).
template <typename T, typename Acc>
__global__ void __launch_bounds__(kBlock)
    rowstats_sweep_kernel(const T* __restrict__ x, long long rows, long long cols, Acc denom,
                          Acc cols_acc, void* mean_out, void* var_out, int out_dtype) {
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
    const Acc mean = sum / cols_acc;

    Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
    if (threadIdx.x < head) {
      const Acc dlt = Cvt<Acc, T>::to(ldg_scalar(r + threadIdx.x)) - mean;
      q[0] += dlt * dlt;
    }
    // second sweep: served by L2; four loads in flight like the first
    v = threadIdx.x;
    for (; v + 3LL * kBlock < nvec; v += 4LL * kBlock) {
      V16 b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) b[u] = ldg_cached16(vp + v + static_cast<long long>(u) * kBlock);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        Acc val[VE];
        unpack_vec<T, Acc>(b[u], val);
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const Acc dlt = val[e] - mean;
          q[e % 4] += dlt * dlt;
        }
      }
    }
    for (; v < nvec; v += kBlock) {
      Acc val[VE];
      unpack_vec<T, Acc>(ldg_cached16(vp + v), val);
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        const Acc dlt = val[e] - mean;
        q[e % 4] += dlt * dlt;
      }
    }
    for (long long i = tail0 + threadIdx.x; i < cols; i += kBlock) {
      const Acc dlt = Cvt<Acc, T>::to(ldg_scalar(r + i)) - mean;
      q[1] += dlt * dlt;
    }
    const Acc ss = block_combine<Acc, SumOp<Acc>>((q[0] + q[1]) + (q[2] + q[3]), smem);
    if (threadIdx.x == 0) {
      if (mean_out) put_stat(mean_out, out_dtype, row, mean);
      if (var_out) put_stat(var_out, out_dtype, row, ss / denom);
    }
  }
}

// Persistent CTAs with a ring of whole rows in shared memory, filled by one producer thread
// with bulk asynchronous copies (TMA) while the 256 consumer threads take mean and
// sum((x-mean)^2) of earlier rows out of the ring: the next rows are always in flight, so the
// two block combines per row no longer leave the memory system idle (the register kernel above
// is latency-bound between rows: 5.2 TB/s on config C4).
// Requirements (host-checked): row_bytes % 16 == 0, x 16-byte aligned, stages * row_bytes fits.
// NV > 0: the row has at most kBlock*NV vectors and each consumer keeps its share as unpacked
// accumulation-type values in registers (one conversion, one shared-memory read per element;
// the stage is handed back to the producer before any arithmetic).  NV == 0: any row that fits
// a stage; both passes read shared memory.
template <typename T, typename Acc, int NV, bool kExact>
__global__ void __launch_bounds__(kBlock + 32)
    rowstats_tma_kernel(const T* __restrict__ x, long long rows, long long cols, int stages, Acc denom,
                        Acc cols_acc, void* mean_out, void* var_out, int out_dtype) {
  extern __shared__ __align__(128) unsigned char ring[];
  constexpr int VE = VecOf<T>::kElems;
  const uint32_t row_bytes = static_cast<uint32_t>(cols * sizeof(T));
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + static_cast<size_t>(stages) * row_bytes);
  uint64_t* empty = full + stages;
  __shared__ Acc s_part[2][kWarps];
  const long long my_rows = (rows > blockIdx.x) ? (rows - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kWarps);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (threadIdx.x >= kBlock) {  // producer warp
    if (threadIdx.x == kBlock) {
      for (long long k = 0; k < my_rows; ++k) {
        const int s = static_cast<int>(k % stages);
        const uint32_t phase = static_cast<uint32_t>((k / stages) & 1);
        if (k >= stages) mbar_wait(&empty[s], phase ^ 1u);
        const long long row = blockIdx.x + k * static_cast<long long>(gridDim.x);
        mbar_arrive_expect_tx(&full[s], row_bytes);
        bulk_load(ring + static_cast<size_t>(s) * row_bytes, x + row * cols, row_bytes, &full[s]);
      }
    }
    return;
  }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = static_cast<int>(cols / VE);
  for (long long k = 0; k < my_rows; ++k) {
    const int s = static_cast<int>(k % stages);
    const uint32_t phase = static_cast<uint32_t>((k / stages) & 1);
    mbar_wait(&full[s], phase);
    const V16* __restrict__ sv = reinterpret_cast<const V16*>(ring + static_cast<size_t>(s) * row_bytes);
    Acc sum, ss;
    Acc mean;
    if constexpr (NV > 0) {
      Acc val[NV * VE];
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int v = threadIdx.x + j * kBlock;
        if (kExact || v < nvec) {  // kExact: the row is exactly kBlock*NV vectors, no predicates
          unpack_vec<T, Acc>(sv[v], &val[j * VE]);
        } else {
#pragma unroll
          for (int e = 0; e < VE; ++e) val[j * VE + e] = Acc(0);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);  // the row now lives in registers
      Acc a[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
#pragma unroll
      for (int i = 0; i < NV * VE; ++i) a[i % 4] += val[i];
      sum = warp_combine<Acc, SumOp<Acc>>((a[0] + a[1]) + (a[2] + a[3]));
      if (lane == 0) s_part[0][warp] = sum;
      consumer_barrier<kBlock>();
      sum = s_part[0][0];
#pragma unroll
      for (int w = 1; w < kWarps; ++w) sum += s_part[0][w];
      mean = sum / cols_acc;  // cols converted once on the host: I2F.S64 per row was the hottest line
      Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        if (kExact || threadIdx.x + j * kBlock < nvec) {
#pragma unroll
          for (int e = 0; e < VE; ++e) {
            const Acc dlt = val[j * VE + e] - mean;
            q[e % 4] += dlt * dlt;
          }
        }
      }
      ss = warp_combine<Acc, SumOp<Acc>>((q[0] + q[1]) + (q[2] + q[3]));
    } else {
      Acc a[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
      for (int v = threadIdx.x; v < nvec; v += kBlock) fold_vec<Acc, SumOp<Acc>, T, 4>(sv[v], a);
      sum = warp_combine<Acc, SumOp<Acc>>((a[0] + a[1]) + (a[2] + a[3]));
      if (lane == 0) s_part[0][warp] = sum;
      consumer_barrier<kBlock>();
      sum = s_part[0][0];
#pragma unroll
      for (int w = 1; w < kWarps; ++w) sum += s_part[0][w];
      mean = sum / cols_acc;
      Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
      for (int v = threadIdx.x; v < nvec; v += kBlock) {
        Acc val[VE];
        unpack_vec<T, Acc>(sv[v], val);
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const Acc dlt = val[e] - mean;
          q[e % 4] += dlt * dlt;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
      ss = warp_combine<Acc, SumOp<Acc>>((q[0] + q[1]) + (q[2] + q[3]));
    }
    if (lane == 0) s_part[1][warp] = ss;
    consumer_barrier<kBlock>();  // also orders this row's reads of s_part[0] before the next row's writes
    if (threadIdx.x == 0) {
      ss = s_part[1][0];
#pragma unroll
      for (int w = 1; w < kWarps; ++w) ss += s_part[1][w];
      const long long row = blockIdx.x + k * static_cast<long long>(gridDim.x);
      if (mean_out) put_stat(mean_out, out_dtype, row, mean);
      if (var_out) put_stat(var_out, out_dtype, row, ss / denom);
    }
  }
}

template <typename T, typename Acc, int NV, bool kExact>
static int launch_rowstats_tma(const T* x, const b200phi_rowstats_desc* d, int out_dtype, Acc denom,
                               int stages, size_t smem, long long ctas, cudaStream_t stream) {
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    cudaError_t ea = cudaFuncSetAttribute(rowstats_tma_kernel<T, Acc, NV, kExact>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (ea != cudaSuccess) return cuda_status(ea);
    attr_set[dev] = true;
  }
  rowstats_tma_kernel<T, Acc, NV, kExact><<<static_cast<unsigned>(ctas), kBlock + 32, smem, stream>>>(
      x, d->rows, d->cols, stages, denom, static_cast<Acc>(d->cols), d->mean, d->var, out_dtype);
  return B200PHI_OK;
}


// Rows of 0.6 .. 16 KiB that are not whole aligned vectors (301- .. 4097-column rows): one WARP
// per row, the row kept in registers -- up to VE-1 head elements (one per lane), NV 16-byte
// vectors per lane from the first aligned address, up to VE-1 tail elements -- so both passes of
// the two-pass variance read registers, no barrier, no second sweep of L2.  The CTA-per-row
// sweep kernel above spends two block reductions per 4 KiB row (fp32 1025 columns: 1.6 TB/s).
// G < 32: 32 / G rows side by side in a warp (16-bit rows of 100..1000 columns are only 12..125
// vectors: a whole warp per row left most lanes without a load); needs G >= VE - 1 so that every
// possible head / tail element has a lane.
template <typename T, typename Acc, int NV, int G = 32>
__global__ void __launch_bounds__(kBlock, NV <= 8 ? 4 : (NV <= 16 ? 2 : 1))
    rowstats_warp_peel_kernel(const T* __restrict__ x, long long rows, long long cols64, Acc denom,
                              Acc cols_acc, void* mean_out, void* var_out, int out_dtype) {
  constexpr int VE = VecOf<T>::kElems;
  static_assert(G >= VE - 1, "a lane per possible head / tail element");
  constexpr int kPerWarp = 32 / G;
  constexpr int kPerCta = kWarps * kPerWarp;
  const int wl = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lane = wl % G, sub = wl / G;
  const int cols = static_cast<int>(cols64);  // (rows of at most 16 KiB)
  // (loop bound on blockIdx alone: the shuffles compile to bare SHFL; 32-bit row arithmetic on a
  // CTA-uniform base: on short 16-bit rows the kernel is bound by instruction issue)
  for (long long c0 = static_cast<long long>(blockIdx.x) * kPerCta; c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * kPerCta) {
    const T* __restrict__ xb = x + c0 * cols64;
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
    const int ntail = cols - tail0;
    const V16* __restrict__ vp = reinterpret_cast<const V16*>(r + head);
    const bool has_h = live && lane < head, has_t = live && lane < ntail;
    Acc hv = Acc(0), tv = Acc(0);
    if (has_h) hv = Cvt<Acc, T>::to(ldg_scalar(r + lane));
    if (has_t) tv = Cvt<Acc, T>::to(ldg_scalar(r + tail0 + lane));
    V16 buf[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = lane + G * k;
      if (live && v < nvec) buf[k] = ldg_stream16(vp + v);
    }
    Acc s[4] = {hv, tv, Acc(0), Acc(0)};
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = lane + G * k;
      if (live && v < nvec) fold_vec<Acc, SumOp<Acc>, T, 4>(buf[k], s);
    }
    Acc sum = (s[0] + s[1]) + (s[2] + s[3]);
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) sum += shfl_xor(sum, m);
    const Acc mean = sum / cols_acc;
    Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
    if (has_h) q[0] = (hv - mean) * (hv - mean);
    if (has_t) q[1] = (tv - mean) * (tv - mean);
    // The second sweep widens the raw vectors again.  Left to itself the compiler keeps the first
    // sweep's widened values alive across the shuffle tree instead (NV = 16: 147 registers, one CTA
    // per SM, 3.7 TB/s on 4097 bf16 columns where the 64 raw registers + accumulators give 5.4).
    if constexpr (sizeof(T) == 2) {
#pragma unroll
      for (int k = 0; k < NV; ++k) {
#pragma unroll
        for (int j = 0; j < 4; ++j) asm volatile("" : "+r"(buf[k].w[j]));
      }
    }
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int v = lane + G * k;
      if (live && v < nvec) {
        Acc val[VE];
        unpack_vec<T, Acc>(buf[k], val);
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          const Acc dlt = val[e] - mean;
          q[e % 4] = fma(dlt, dlt, q[e % 4]);
        }
      }
    }
    Acc ss = (q[0] + q[1]) + (q[2] + q[3]);
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) ss += shfl_xor(ss, m);
    if (live && lane == 0) {
      if (mean_out) put_stat(mean_out, out_dtype, c0 + lr, mean);
      if (var_out) put_stat(var_out, out_dtype, c0 + lr, ss / denom);
    }
  }
}

// mean / variance consumer of rows_staged_kernel (rows_staged.cuh): the two sweeps of the two-pass
// form over the thread's own row in shared memory.
template <typename T, typename Acc>
struct StagedStats {
  struct Params {
    Acc denom, cols_acc;
    void* mean_out;
    void* var_out;
    int out_dtype;
  };
  template <int L>
  static __device__ __forceinline__ void run(const T* __restrict__ sp, int n, int i0, int i1, int rot,
                                             bool live, long long row, int h, const Params& p) {
    Acc s[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
    smem_visit<T>(sp, n, i0, i1, rot, [&](int, auto x, int slot) { s[slot % 4] += Cvt<Acc, decltype(x)>::to(x); });
    const Acc sum = group_all<L>((s[0] + s[1]) + (s[2] + s[3]), [](Acc a, Acc b) { return a + b; });
    const Acc mean = sum / p.cols_acc;
    Acc q[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
    smem_visit<T>(sp, n, i0, i1, rot, [&](int, auto x, int slot) {
      const Acc dlt = Cvt<Acc, decltype(x)>::to(x) - mean;
      q[slot % 4] += dlt * dlt;
    });
    const Acc ss = group_all<L>((q[0] + q[1]) + (q[2] + q[3]), [](Acc a, Acc b) { return a + b; });
    if (live && h == 0) {
      if (p.mean_out) put_stat(p.mean_out, p.out_dtype, row, mean);
      if (p.var_out) put_stat(p.var_out, p.out_dtype, row, ss / p.denom);
    }
  }
};

// Smallest row, in bytes, that takes the warp-per-row peel kernel rather than the sub-warp scalar
// one (B200PHI_ROWSTATS_PEEL_MIN overrides, for measurement).
// Measured (4 GiB tensors): fp32 301 columns (1204 B) 4.64 vs 3.79 TB/s for the scalar kernel,
// 201 columns (804 B) 3.22 vs 4.91; bf16 301 columns (602 B) 1.97 vs 1.59, 201 columns 1.48 vs 2.11.
static long long peel_min_bytes(size_t esize) {
  static const long long env = [] {
    const char* e = getenv("B200PHI_ROWSTATS_PEEL_MIN");
    return e ? atoll(e) : -1LL;
  }();
  return env >= 0 ? env : (esize == 2 ? 600 : 1024);
}

// Smallest row, in bytes, that takes the one-sweep moments kernel instead of the two-sweep one
// (B200PHI_ROWSTATS_STREAM_MIN overrides, for measurement).
static long long stream_min_bytes(bool aligned) {
  static const long long env = [] {
    const char* e = getenv("B200PHI_ROWSTATS_STREAM_MIN");
    return e ? atoll(e) : -1LL;
  }();
  if (env >= 0) return env;
  (void)aligned;
  return 24 * 1024;
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
  const long long row_bytes = n * static_cast<long long>(sizeof(T));
  // The TMA ring was measured at 4.1 TB/s on C4 against 5.2 TB/s for the register kernel
  // (instruction-issue bound: profiles/r01_rowstats_tma.md); it stays available for study.
  static const bool no_tma = getenv("B200PHI_ROWSTATS_TMA") == nullptr;
  // the asynchronous-copy ring was measured at 4.9 TB/s on C4 against 5.8 for the register
  // kernel: opt-in for study only
  static const bool no_async = getenv("B200PHI_ROWSTATS_ASYNC") == nullptr;
  static const int force_g = getenv("B200PHI_ROWSTATS_G") ? atoi(getenv("B200PHI_ROWSTATS_G")) : 0;
  static const bool no_multi = getenv("B200PHI_ROWSTATS_NO_MULTI") != nullptr;
  static const bool rs_warp_peel_aligned =
      !(getenv("B200PHI_ROWSTATS_WARP_PEEL_ALIGNED") && atoi(getenv("B200PHI_ROWSTATS_WARP_PEEL_ALIGNED")) == 0);
  // aligned 16-bit rows of more than this many vectors (and at most 128) take the sub-warp peel
  // kernel instead of rowstats_multi_kernel (B200PHI_ROWSTATS_PEEL_ALIGNED_MIN overrides)
  static const long long rs_peel_aligned_min =
      getenv("B200PHI_ROWSTATS_PEEL_ALIGNED_MIN") ? atoll(getenv("B200PHI_ROWSTATS_PEEL_ALIGNED_MIN")) : 32;  // (bf16 384 / 512 columns 4.8 / 5.7 -> 6.1 / 6.9 TB/s)
  static const bool no_staged = getenv("B200PHI_NO_ROW_STAGED") != nullptr;
  if (!no_staged && !no_multi && force_g == 0 && staged_rows_ok(d->rows, row_bytes, static_cast<int>(sizeof(T)), kStagedStats, reinterpret_cast<uintptr_t>(x) % 16 == 0)) {
    // short rows that are not whole aligned vectors: a thread per row out of shared memory
    typename StagedStats<T, Acc>::Params sp{denom, static_cast<Acc>(n), d->mean, d->var, out_dtype};
    return launch_rows_staged<T, StagedStats<T, Acc>>(x, d->rows, static_cast<int>(n), sp, stream);
  }
  if (aligned && (nvec == 128 * 8 || nvec == 64 * 8 || nvec == 32 * 8) && force_g == 0 && d->rows >= 64 &&
      getenv("B200PHI_ROWSTATS_TWO_SWEEP") == nullptr) {
    // full power-of-two rows (the common hidden sizes): one sweep of the registers, one barrier
    // (C4: 43.6 us against 46.4 for the two-sweep register kernel)
    const int GG = static_cast<int>(nvec / 8);
    long long ctas = (d->rows + (kBlock / GG) - 1) / (kBlock / GG);
    if (ctas > 0x7fffffffLL) ctas = 0x7fffffffLL;
    const unsigned gr = static_cast<unsigned>(ctas);
    if (GG == 128)
      rowstats_moments_kernel<T, Acc, 128, 8><<<gr, kBlock, 0, stream>>>(x, d->rows, n, denom, d->mean, d->var, out_dtype);
    else if (GG == 64)
      rowstats_moments_kernel<T, Acc, 64, 8><<<gr, kBlock, 0, stream>>>(x, d->rows, n, denom, d->mean, d->var, out_dtype);
    else
      rowstats_moments_kernel<T, Acc, 32, 8><<<gr, kBlock, 0, stream>>>(x, d->rows, n, denom, d->mean, d->var, out_dtype);
  } else if (sizeof(T) == 2 && (!aligned || nvec > rs_peel_aligned_min) && n / VE >= 4 && n / VE <= 128 && d->rows >= 64 &&
             getenv("B200PHI_ROWSTATS_NO_PEEL") == nullptr && getenv("B200PHI_ROWSTATS_PEEL_MIN") == nullptr) {
    // 16-bit rows of 32..1031 columns that are not whole aligned vectors: 8 or 16 lanes per row
    // (the scalar-load kernel read these at 2.0-3.2 TB/s, a whole warp per row at 2.0 on 301 columns)
    const long long nv_row = n / VE;
    const int lanes = nv_row <= 64 ? 8 : 16;
    long long ctas = (d->rows + kWarps * (32 / lanes) - 1) / (kWarps * (32 / lanes));
    const long long cap = static_cast<long long>(sms) * 8 * 8;
    if (ctas > cap) ctas = cap;
    const unsigned gsz = static_cast<unsigned>(ctas);
    if constexpr (sizeof(T) == 2) {
      if (nv_row <= 16)
        rowstats_warp_peel_kernel<T, Acc, 2, 8><<<gsz, kBlock, 0, stream>>>(x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
      else if (nv_row <= 32)
        rowstats_warp_peel_kernel<T, Acc, 4, 8><<<gsz, kBlock, 0, stream>>>(x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
      else if (nv_row <= 64)
        rowstats_warp_peel_kernel<T, Acc, 8, 8><<<gsz, kBlock, 0, stream>>>(x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
      else
        rowstats_warp_peel_kernel<T, Acc, 8, 16><<<gsz, kBlock, 0, stream>>>(x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
    }
  } else if (aligned && !no_multi && force_g == 0 && (nvec <= 128 || (nvec > 256 && nvec <= 512)) && d->rows >= 64 &&
             !(sizeof(T) == 2 && nvec > rs_peel_aligned_min && nvec <= 128)) {
    // R rows per group so that all eight load slots are used (see rowstats_multi_kernel)
    const Acc na = static_cast<Acc>(n);
#define B200PHI_RS_MULTI(GG, NN, RR)                                                          \
  {                                                                                           \
    long long ctas = (d->rows + (kBlock / GG) * RR - 1) / ((kBlock / GG) * RR);                \
    const long long cap = static_cast<long long>(sms) * 8 * 8;                                 \
    if (ctas > cap) ctas = cap;                                                                \
    rowstats_multi_kernel<T, Acc, GG, NN, RR><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>( \
        x, d->rows, n, denom, na, d->mean, d->var, out_dtype);                                 \
  }
    if (nvec <= 32)  // sub-warp groups: the shuffle tree spans 8 or 16 lanes, not 32
      B200PHI_RS_MULTI(8, 4, 2)
    else if (nvec <= 64)
      B200PHI_RS_MULTI(16, 4, 2)
    else if (nvec <= 128)
      B200PHI_RS_MULTI(16, 8, 1)
    else
      B200PHI_RS_MULTI(64, 8, 1)
#undef B200PHI_RS_MULTI
  } else if (sizeof(T) == 2 && aligned && nvec > 128 && nvec < 256 && d->rows >= 64 && rs_warp_peel_aligned) {
    // aligned 16-bit rows of 2..4 KiB that are not the moments kernel's power of two: the warp-per-row
    // peel kernel (head and tail empty) instead of the two-barrier register kernel
    long long ctas = (d->rows + kWarps - 1) / kWarps;
    const long long cap = static_cast<long long>(sms) * 8 * 8;
    if (ctas > cap) ctas = cap;
    rowstats_warp_peel_kernel<T, Acc, 8><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
  } else if (aligned && nvec <= 32 * 8 && d->rows >= static_cast<long long>(sms) * 16) {
    long long ctas = (d->rows + kWarps - 1) / kWarps;
    if (ctas > 0x7fffffffLL) ctas = 0x7fffffffLL;
    rowstats_reg_kernel<T, Acc, 32, 8><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
  } else if (aligned && !no_tma && row_bytes >= 4096 && row_bytes <= 98304) {
    // ring of whole rows: 2..6 stages, as many CTAs per SM as 200 KiB of shared memory allow
    int stages = static_cast<int>((192 * 1024) / row_bytes);
    stages = stages > 12 ? 4 : (stages >= 8 ? 4 : (stages >= 6 ? 3 : 2));
    const size_t smem = static_cast<size_t>(stages) * row_bytes + 2 * stages * sizeof(uint64_t);
    int per_sm = static_cast<int>((200 * 1024) / (smem + 1024));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 4) per_sm = 4;
    long long ctas = static_cast<long long>(sms) * per_sm;
    if (ctas > d->rows) ctas = d->rows;
    int rc;
    if (nvec == kBlock * 4)
      rc = launch_rowstats_tma<T, Acc, 4, true>(x, d, out_dtype, denom, stages, smem, ctas, stream);
    else if (nvec == kBlock * 2)
      rc = launch_rowstats_tma<T, Acc, 2, true>(x, d, out_dtype, denom, stages, smem, ctas, stream);
    else if (nvec <= kBlock * 2)
      rc = launch_rowstats_tma<T, Acc, 2, false>(x, d, out_dtype, denom, stages, smem, ctas, stream);
    else if (nvec <= kBlock * 4)
      rc = launch_rowstats_tma<T, Acc, 4, false>(x, d, out_dtype, denom, stages, smem, ctas, stream);
    else
      rc = launch_rowstats_tma<T, Acc, 0, false>(x, d, out_dtype, denom, stages, smem, ctas, stream);
    if (rc != B200PHI_OK) return rc;
  } else if (aligned && force_g == 0 && !no_async && sizeof(T) <= 4 && nvec > 128 * 2 && nvec <= 256 * 4) {
    // 4..16 KiB rows: asynchronous-copy ring, 2 rows per CTA in flight behind the one in registers
    int rc = (nvec <= 256 * 2) ? launch_rowstats_async<T, Acc, 2, 3>(x, d, out_dtype, denom, sms, stream)
                               : launch_rowstats_async<T, Acc, 4, 3>(x, d, out_dtype, denom, sms, stream);
    if (rc != B200PHI_OK) return rc;
  } else if (aligned && nvec <= 256 * 4 && force_g == 256) {
    long long ctas = d->rows > 0x7fffffffLL ? 0x7fffffffLL : d->rows;
    rowstats_reg_kernel<T, Acc, 256, 4><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
  } else if (aligned && nvec <= 128 * 8) {
    long long ctas = (d->rows + 1) / 2;
    if (ctas > 0x7fffffffLL) ctas = 0x7fffffffLL;
    rowstats_reg_kernel<T, Acc, 128, 8><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
  } else if (!aligned && row_bytes >= peel_min_bytes(sizeof(T)) && row_bytes < 16400 && d->rows >= 64 && getenv("B200PHI_ROWSTATS_NO_PEEL") == nullptr) {
    // 1..16 KiB rows that are not whole aligned vectors: a warp per row, row in registers
    long long ctas = (d->rows + kWarps - 1) / kWarps;
    const long long cap = static_cast<long long>(sms) * 8 * 8;
    if (ctas > cap) ctas = cap;
    if (row_bytes < 2064)
      rowstats_warp_peel_kernel<T, Acc, 4><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
          x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
    else if (row_bytes < 4112)
      rowstats_warp_peel_kernel<T, Acc, 8><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
          x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
    else if (row_bytes < 8208)
      rowstats_warp_peel_kernel<T, Acc, 16><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
          x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
    else if (row_bytes < 12304)
      rowstats_warp_peel_kernel<T, Acc, 24><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
          x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
    else
      rowstats_warp_peel_kernel<T, Acc, 32><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
          x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
  } else if (!aligned && row_bytes >= 16400 && row_bytes <= 32768 && d->rows >= 64 &&
             getenv("B200PHI_ROWSTATS_NO_PEEL") == nullptr) {
    // 16..32 KiB rows that are not whole aligned vectors: the whole CTA on one row, row in
    // registers (8 vectors per thread + a head / tail element), two register sweeps.  The two-sweep
    // global kernel read these at 3.0-3.9 TB/s (its second sweep misses L2), the one-sweep stream
    // kernel at 3.0-4.1.
    long long ctas = d->rows > 0x7fffffffLL ? 0x7fffffffLL : d->rows;
    // every row holds at least (n - (VE - 1)) / VE whole vectors: 4 full slots per thread from
    // 1024 of them (16 KiB), 6 from 1536 (24 KiB)
    const long long min_nvec = (n - (VE - 1)) / VE;
    if (min_nvec >= 6 * 256)
      rowstats_reg_kernel<T, Acc, 256, 8, true, 6><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
          x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
    else if (min_nvec >= 4 * 256)
      rowstats_reg_kernel<T, Acc, 256, 8, true, 4><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
          x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
    else
      rowstats_reg_kernel<T, Acc, 256, 8, true><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
          x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
  } else if (n <= 1023 && d->rows >= 64) {  // (one slot spare: a 16-bit row may start mid-word)
    // short rows that are ragged or unaligned: sub-warp groups, scalar loads, row in registers
    const int c = static_cast<int>(n);
    const Acc na = static_cast<Acc>(n);
#define B200PHI_RS_SCALAR(GG, NN)                                                              \
  {                                                                                            \
    long long ctas = (d->rows + kWarps * (32 / GG) - 1) / (kWarps * (32 / GG));                \
    const long long cap = static_cast<long long>(sms) * 8 * 8;                                 \
    if (ctas > cap) ctas = cap;                                                                \
    rowstats_scalar_kernel<T, Acc, GG, NN><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>( \
        x, d->rows, c, denom, na, d->mean, d->var, out_dtype);                                 \
  }
    if (n < 128)
      B200PHI_RS_SCALAR(8, 16)
    else if (n < 256)
      B200PHI_RS_SCALAR(16, 16)
    else if (n < 512)
      B200PHI_RS_SCALAR(32, 16)
    else
      B200PHI_RS_SCALAR(32, 32)
#undef B200PHI_RS_SCALAR
  } else if (getenv("B200PHI_ROWSTATS_SWEEP") == nullptr && row_bytes >= stream_min_bytes(aligned)) {
    // (measured: 16 KiB rows are faster on the two-sweep kernel, 4.2 vs 3.5 TB/s -- the triple
    // merges and two barriers per row weigh more on short rows; 32 KiB rows 5.2 vs 3.7)
    // long rows (and medium rows of any alignment): one sweep, pairwise moment merges
    long long ctas = d->rows > 0x7fffffffLL ? 0x7fffffffLL : d->rows;
    const long long cap = static_cast<long long>(sms) * 8 * 4;
    if (ctas > cap) ctas = cap;
    rowstats_stream_kernel<T, Acc, 4><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, d->mean, d->var, out_dtype);
  } else {
    long long ctas = d->rows > 0x7fffffffLL ? 0x7fffffffLL : d->rows;
    rowstats_sweep_kernel<T, Acc><<<static_cast<unsigned>(ctas), kBlock, 0, stream>>>(
        x, d->rows, n, denom, static_cast<Acc>(n), d->mean, d->var, out_dtype);
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
