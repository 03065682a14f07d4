// Reduction kernels for sm_100a.  Four shapes of work after the host has merged
// the dims (reduce_plan.h):
//   row_reduce_kernel    [rows, cols] -> [rows]       contiguous axis reduced   (replaces ReduceAnyKernel
//                                                     in its kReduceLastDim mode, reduce_function.h:573-699)
//   row_split_kernel     same, few very long rows: the row is cut across CTAs and the last CTA to
//                        finish combines the partials (replaces CubTensorReduce :938-1039 and the
//                        Eigen detour for numel > INT32_MAX, kps/reduce_kernel.cu:196-243)
//   col_reduce_kernel    [outer, R, inner] -> [outer, inner]  (replaces ReduceHigherDimKernel :706-788
//                        and its second pass; one launch, last-block finalise)
//   generic_reduce_kernel  any alternating keep/reduce pattern up to rank 9 (kReduceAny mode)
// All index math is 64-bit: a 2^31-element shard is an ordinary input.
#pragma once

#include <type_traits>

#include "prim.cuh"

namespace b200phi {

constexpr int kFoldLanes = 4;  // independent accumulators per thread

template <typename Acc, typename Op>
__device__ __forceinline__ Acc fold_lanes(Acc (&a)[kFoldLanes]) {
  return Op::combine(Op::combine(a[0], a[1]), Op::combine(a[2], a[3]));
}

// Consume as many batches of UU strided 16-byte vectors as fit: all UU loads are
// issued before the first use so UU requests per thread are in flight.
template <typename T, typename Acc, typename Op, int G, int UU>
__device__ __forceinline__ void span_step(const V16* __restrict__ vp, long long& v, long long nvec,
                                          Acc (&acc)[kFoldLanes]) {
  for (; v + static_cast<long long>(UU - 1) * G < nvec; v += static_cast<long long>(UU) * G) {
    V16 buf[UU];
#pragma unroll
    for (int u = 0; u < UU; ++u) buf[u] = ldg_stream16(vp + v + static_cast<long long>(u) * G);
#pragma unroll
    for (int u = 0; u < UU; ++u) fold_vec<Acc, Op, T, kFoldLanes>(buf[u], acc);
  }
}

// Reduce `cols` contiguous elements starting at `row` with a group of G threads
// (g = index in group). Handles any element alignment: scalar head up to the
// first 16-byte boundary, 16-byte vectors, scalar tail.
template <typename T, typename Acc, typename Op, int G, int U>
__device__ __forceinline__ Acc reduce_span(const T* __restrict__ row, long long cols, int g) {
  constexpr int VE = VecOf<T>::kElems;
  Acc acc[kFoldLanes];
#pragma unroll
  for (int i = 0; i < kFoldLanes; ++i) acc[i] = Op::identity();

  const uintptr_t addr = reinterpret_cast<uintptr_t>(row);
  long long head = static_cast<long long>(((16 - (addr & 15)) & 15) / sizeof(T));
  if (head > cols) head = cols;
  const long long nvec = (cols - head) / VE;
  const long long tail0 = head + nvec * VE;

  if (g < head) acc[0] = Op::combine(acc[0], load_as<Acc, Op>(ldg_scalar(row + g)));

  const V16* __restrict__ vp = reinterpret_cast<const V16*>(row + head);
  long long v = g;
  span_step<T, Acc, Op, G, U>(vp, v, nvec, acc);
  if constexpr (U > 4) span_step<T, Acc, Op, G, 4>(vp, v, nvec, acc);
  if constexpr (U > 2) span_step<T, Acc, Op, G, 2>(vp, v, nvec, acc);
  if constexpr (U > 1) span_step<T, Acc, Op, G, 1>(vp, v, nvec, acc);
  for (long long i = tail0 + g; i < cols; i += G)
    acc[1] = Op::combine(acc[1], load_as<Acc, Op>(ldg_scalar(row + i)));
  return fold_lanes<Acc, Op>(acc);
}

// Every row starts on a 16-byte boundary and holds a whole number of vectors (checked on the
// host): no head/tail peel, no per-row address inspection.
template <typename T, typename Acc, typename Op, int G, int U>
__device__ __forceinline__ Acc reduce_span_aligned(const T* __restrict__ row, long long nvec, int g) {
  Acc acc[kFoldLanes];
#pragma unroll
  for (int i = 0; i < kFoldLanes; ++i) acc[i] = Op::identity();
  const V16* __restrict__ vp = reinterpret_cast<const V16*>(row);
  long long v = g;
  span_step<T, Acc, Op, G, U>(vp, v, nvec, acc);
  if constexpr (U > 4) span_step<T, Acc, Op, G, 4>(vp, v, nvec, acc);
  if constexpr (U > 2) span_step<T, Acc, Op, G, 2>(vp, v, nvec, acc);
  if constexpr (U > 1) span_step<T, Acc, Op, G, 1>(vp, v, nvec, acc);
  return fold_lanes<Acc, Op>(acc);
}

// G threads per row (G = 32: a warp per row, no barrier at all; 64/128: sub-CTA groups on
// named barriers; 256: the CTA works on one row).  Measured on C1 (16 KiB rows): G = 256 with
// one row per CTA reads 7.09 TB/s, G = 32 reads 6.74 TB/s (profiles/r01_microbench.md).
template <typename T, typename Acc, typename Op, int G, int U, bool kAligned>
__global__ void __launch_bounds__(kBlock)
    row_reduce_kernel(const T* __restrict__ x, long long rows, long long cols, Epilogue ep) {
  __shared__ Acc smem[2 * kWarps];
  constexpr int kRowsPerCta = kBlock / G;
  constexpr int VE = VecOf<T>::kElems;
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  int parity = 0;
  for (long long row = static_cast<long long>(blockIdx.x) * kRowsPerCta + grp; row < rows;
       row += static_cast<long long>(gridDim.x) * kRowsPerCta) {
    Acc v;
    if constexpr (kAligned) {
      v = reduce_span_aligned<T, Acc, Op, G, U>(x + row * cols, cols / VE, g);
    } else {
      v = reduce_span<T, Acc, Op, G, U>(x + row * cols, cols, g);
    }
    v = group_combine_leader<Acc, Op, G>(v, smem, parity);
    parity ^= 1;
    if (g == 0) store_final_16<Op, T>(ep, row, v);
  }
}

// Rows of up to ~16 KiB (hidden sizes, feature maps: the LayerNorm / softmax / per-channel
// widths).  The cascade in reduce_span_aligned serves such a row in two or three dependent rounds
// of one to four loads (a 1568-byte bf16 row: 3 vectors per lane, read at 3.4 TB/s; a 6 KiB row
// 2.6 TB/s).  Here a group of G threads takes R rows at a time and issues ALL of their loads --
// NV predicated 16-byte loads per row per thread, R * NV = 8 in flight -- in one round; the R
// partial results then share one shuffle tree pass (interleaved chains) and one barrier.
template <typename T, typename Acc, typename Op, int G, int NV, int R>
__global__ void __launch_bounds__(kBlock)
    row_batch_kernel(const T* __restrict__ x, long long rows, long long nvec, Epilogue ep) {
  constexpr int VE = VecOf<T>::kElems;
  constexpr int kGroups = kBlock / G;
  constexpr int kW = G >= 32 ? G / 32 : 1;
  __shared__ Acc smem[2][R][kWarps];
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long cols = nvec * VE;
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
    Acc a[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      Acc acc[kFoldLanes];
#pragma unroll
      for (int i = 0; i < kFoldLanes; ++i) acc[i] = Op::identity();
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long v = g + static_cast<long long>(k) * G;
        if (r0 + r < rows && v < nvec) fold_vec<Acc, Op, T, kFoldLanes>(buf[r][k], acc);
      }
      a[r] = fold_lanes<Acc, Op>(acc);
    }
    // G < 32: several rows share a warp, the tree stays inside the row's G lanes
#pragma unroll
    for (int m = (G < 32 ? G / 2 : 16); m > 0; m >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) a[r] = Op::combine(a[r], shfl_xor(a[r], m));
    }
    if constexpr (G > 32) {
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) smem[parity][r][warp] = a[r];
      }
      group_barrier<G>();
      if (g == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          a[r] = smem[parity][r][grp * kW];
#pragma unroll
          for (int w = 1; w < kW; ++w) a[r] = Op::combine(a[r], smem[parity][r][grp * kW + w]);
        }
      }
      parity ^= 1;
    }
    if (g == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r)
        if (r0 + r < rows) store_final_16<Op, T>(ep, r0 + r, a[r]);
    }
  }
}

// Aligned rows of 1..16 vectors (8..64 fp32, 8..128 bf16 elements: class counts, heads, small
// feature maps).  G = 1, 2, 4, 8 or 16 lanes per row, one 16-byte load per lane per row, R rows
// per group with all R loads issued together.  Rows are dealt round-robin over the warp's groups,
// so one warp instruction reads 32/G consecutive rows: a contiguous 512 bytes when the row is
// exactly G vectors.  row_small_kernel reads such a row element by element in a loop with one
// load in flight (fp32 [., 32]: 4.1 TB/s).
template <typename T, typename Acc, typename Op, int G, int R>
__global__ void __launch_bounds__(kBlock)
    row_tiny_kernel(const T* __restrict__ x, long long rows, int nvec, Epilogue ep) {
  constexpr int kPerWarp = 32 / G;
  constexpr int kPerCta = kWarps * kPerWarp * R;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  // The loop bound depends on blockIdx alone: the compiler then knows the body is not divergent
  // and emits bare shuffles.  With a per-warp bound every shuffle carried a convergence check
  // and a collective slow path (24 of them per pass at G = 8), and the kernel was issue-bound.
  for (long long c0 = static_cast<long long>(blockIdx.x) * kPerCta; c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * kPerCta) {
    const long long left = rows - c0;
    const int live = left > kPerCta ? kPerCta : static_cast<int>(left);  // rows of this pass
    const int first = warp * (kPerWarp * R) + sub;                       // this lane's first row
    const V16* __restrict__ xv = reinterpret_cast<const V16*>(x) + c0 * nvec;
    V16 buf[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = first + r * kPerWarp;
      if (row < live && g < nvec) buf[r] = ldg_stream16(xv + (row * nvec + g));
    }
    Acc a[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = first + r * kPerWarp;
      Acc acc[kFoldLanes];
#pragma unroll
      for (int i = 0; i < kFoldLanes; ++i) acc[i] = Op::identity();
      if (row < live && g < nvec) fold_vec<Acc, Op, T, kFoldLanes>(buf[r], acc);
      a[r] = fold_lanes<Acc, Op>(acc);
    }
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) a[r] = Op::combine(a[r], shfl_xor(a[r], m));
    }
    if (g == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int row = first + r * kPerWarp;
        if (row < live) store_final_16<Op, T>(ep, c0 + row, a[r]);
      }
    }
  }
}

// Short rows of any alignment, 129..1023 elements (odd hidden sizes): G = 16 or 32 lanes per row,
// NE scalar loads per lane issued together (4-byte pair loads for 16-bit types in the 32-per-lane
// form), shuffle tree inside the G lanes.  The peel in reduce_span costs such a row a scalar
// head, a few vectors and a scalar tail in sequence (fp32 1001 columns: 4.5 TB/s).
template <typename T, typename Acc, typename Op, int G, int NE>
__global__ void __launch_bounds__(kBlock)
    row_scalar_kernel(const T* __restrict__ x, long long rows, int cols, Epilogue ep) {
  constexpr int kPerWarp = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  // (loop bound on blockIdx alone: bare shuffles, see row_tiny_kernel)
  for (long long c0 = static_cast<long long>(blockIdx.x) * (kWarps * kPerWarp); c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * (kWarps * kPerWarp)) {
    const long long row = c0 + warp * kPerWarp + sub;
    const bool live = row < rows;
    T raw[NE];
    int odd;
    short_row_load<T, G, NE>(x + row * cols, cols, g, live, raw, &odd);
    Acc acc[2] = {Op::identity(), Op::identity()};
#pragma unroll
    for (int k = 0; k < NE; ++k) {
      const int i = short_row_index<T, G, NE>(k, g, odd);
      if (live && i >= 0 && i < cols) acc[k & 1] = Op::combine(acc[k & 1], load_as<Acc, Op>(raw[k]));
    }
    Acc v = Op::combine(acc[0], acc[1]);
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) v = Op::combine(v, shfl_xor(v, m));
    if (live && g == 0) store_final_16<Op, T>(ep, row, v);
  }
}

// Rows of 1- and 2-byte elements that are NOT whole aligned 16-byte vectors (odd hidden sizes: 101 ..
// 2047 bf16 columns): G = 8, 16 or 32 lanes per row, 32 / G rows side by side in a warp.  A row
// starts up to VE - 1 elements before a 16-byte boundary and ends up to VE - 1 elements after one, so
// a lane issues, all together, one head element (lane g < head), NV 16-byte vectors from the first
// aligned address and one tail element, folds them and joins a shuffle tree inside its G lanes.
// row_scalar_kernel reads such rows with 2-byte (or paired 4-byte) loads: bf16 201 columns 3.7 TB/s,
// max on 1001 columns 3.5 TB/s.
// RB rows per lane group are loaded before any is folded.  The kernel is bound by instruction
// issue (ncu: 85 % issue-slot use on 101 bf16 columns), so the per-row address arithmetic is 32-bit:
// a CTA-uniform 64-bit base per iteration plus (row within the iteration) * cols.
template <typename T, typename Acc, typename Op, int G, int NV, int RB = 1>
__global__ void __launch_bounds__(kBlock, NV * RB >= 16 ? 2 : 4)  // (64 registers; 16 vectors per lane: 128, two CTAs per SM)
    row_peel_kernel(const T* __restrict__ x, long long rows, int cols, Epilogue ep) {
  constexpr int VE = VecOf<T>::kElems;
  static_assert(G >= VE - 1, "a lane per possible head / tail element");
  constexpr int kPerWarp = 32 / G;
  constexpr int kPerCta = kWarps * kPerWarp * RB;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  // (loop bound on blockIdx alone: bare shuffles, see row_tiny_kernel)
  for (long long c0 = static_cast<long long>(blockIdx.x) * kPerCta; c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * kPerCta) {
    const T* __restrict__ xb = x + c0 * cols;  // CTA-uniform
    const unsigned xlow = static_cast<unsigned>(reinterpret_cast<uintptr_t>(xb)) & 15u;
    const long long left = rows - c0;
    const int nlive = left < kPerCta ? static_cast<int>(left) : kPerCta;
    bool live[RB], has_h[RB], has_t[RB];
    int nvec[RB];
    T hv[RB], tv[RB];
    V16 buf[RB][NV];
#pragma unroll
    for (int b = 0; b < RB; ++b) {
      const int lr = (warp * RB + b) * kPerWarp + sub;
      live[b] = lr < nlive;
      const unsigned off = static_cast<unsigned>(lr) * static_cast<unsigned>(cols);  // elements
      const T* __restrict__ r = xb + off;
      int head = static_cast<int>(((16u - ((xlow + off * static_cast<unsigned>(sizeof(T))) & 15u)) & 15u) /
                                  sizeof(T));
      if (head > cols) head = cols;
      nvec[b] = (cols - head) / VE;
      const int tail0 = head + nvec[b] * VE;
      const V16* __restrict__ vp = reinterpret_cast<const V16*>(r + head);
      has_h[b] = live[b] && g < head;
      has_t[b] = live[b] && tail0 + g < cols;
      if (has_h[b]) hv[b] = ldg_scalar(r + g);
      if (has_t[b]) tv[b] = ldg_scalar(r + tail0 + g);
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int v = g + k * G;
        if (live[b] && v < nvec[b]) buf[b][k] = ldg_stream16(vp + v);
      }
    }
#pragma unroll
    for (int b = 0; b < RB; ++b) {
      Acc acc[kFoldLanes];
#pragma unroll
      for (int i = 0; i < kFoldLanes; ++i) acc[i] = Op::identity();
      if (has_h[b]) acc[2] = Op::combine(acc[2], load_as<Acc, Op>(hv[b]));
      if (has_t[b]) acc[3] = Op::combine(acc[3], load_as<Acc, Op>(tv[b]));
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int v = g + k * G;
        if (live[b] && v < nvec[b]) fold_vec<Acc, Op, T, kFoldLanes>(buf[b][k], acc);
      }
      Acc a = fold_lanes<Acc, Op>(acc);
#pragma unroll
      for (int m = G / 2; m > 0; m >>= 1) a = Op::combine(a, shfl_xor(a, m));
      if (live[b] && g == 0) store_final_16<Op, T>(ep, c0 + (warp * RB + b) * kPerWarp + sub, a);
    }
  }
}

// The reduce family's consumer for rows_staged_kernel: one thread folds its whole row out of shared
// memory into four accumulators (the element order inside a row is the thread's own business: sums
// are compared at the family's tolerance, every other operator is order-independent).
template <typename T, typename Acc, typename Op>
struct StagedReduce {
  using Params = Epilogue;
  template <int L>
  static __device__ __forceinline__ void run(const T* __restrict__ sp, int cols, int i0, int i1, int rot,
                                             bool live, long long row, int h, const Epilogue& ep) {
    Acc a[kFoldLanes];
#pragma unroll
    for (int i = 0; i < kFoldLanes; ++i) a[i] = Op::identity();
    smem_visit<T>(sp, cols, i0, i1, rot, [&](int, auto v, int slot) {
      Acc w;
      if constexpr (Is16Float<T>::value) {  // v is the element widened to float
        if constexpr (Op::kLogical) {
          w = static_cast<Acc>(v != 0.f ? 1 : 0);
        } else {
          w = pre_as<Acc, Op>(static_cast<Acc>(v));
        }
      } else {
        w = load_as<Acc, Op>(v);
      }
      a[slot % kFoldLanes] = Op::combine(a[slot % kFoldLanes], w);
    });
    const Acc v = group_all<L>(fold_lanes<Acc, Op>(a), [](Acc p, Acc q) { return Op::combine(p, q); });
    if (live && h == 0) store_final_16<Op, T>(ep, row, v);
  }
};

// Short rows (softmax / classifier widths: a few to ~128 elements): G = 1, 2, 4, 8 or 16 lanes
// per row, 32/G rows per warp, scalar loads (consecutive rows are contiguous, so a warp still
// reads one contiguous span and L1 serves the strided revisits), xor-shuffle combine inside the
// G-lane group.  A full warp per 40-byte row would leave 22 of 32 lanes idle.
template <typename T, typename Acc, typename Op, int G>
__global__ void __launch_bounds__(kBlock)
    row_small_kernel(const T* __restrict__ x, long long rows, int cols, Epilogue ep) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane % G;
  constexpr int kRowsPerWarp = 32 / G;
  // the loop condition is CTA-uniform: every lane takes part in the shuffles
  for (long long c0 = static_cast<long long>(blockIdx.x) * (kWarps * kRowsPerWarp); c0 < rows;
       c0 += static_cast<long long>(gridDim.x) * (kWarps * kRowsPerWarp)) {
    const long long row = c0 + warp * kRowsPerWarp + lane / G;
    const bool valid = row < rows;
    Acc acc = Op::identity();
    if (valid) {
      const T* __restrict__ r = x + row * cols;
      for (int i = g; i < cols; i += G) acc = Op::combine(acc, load_as<Acc, Op>(r[i]));
    }
#pragma unroll
    for (int m = G >> 1; m > 0; m >>= 1) acc = Op::combine(acc, shfl_xor(acc, m));
    if (valid && g == 0) store_final_16<Op, T>(ep, row, acc);
  }
}

// Few, very long rows (a full reduction is rows == 1).  grid = (splits, rows).
// Each CTA reduces one 16-byte-aligned-friendly segment, parks its partial in
// `partial[row*splits + s]`, takes a ticket; the CTA that draws the last ticket
// combines the partials in index order (deterministic, no float atomics) and
// resets the ticket so the workspace can be reused / graph-replayed.
template <typename T, typename Acc, typename Op, int U>
__global__ void __launch_bounds__(kBlock)
    row_split_kernel(const T* __restrict__ x, long long rows, long long cols, long long seg,
                     Acc* __restrict__ partial, unsigned* __restrict__ tickets, Epilogue ep) {
  __shared__ Acc smem[kWarps];
  __shared__ unsigned s_ticket;
  const long long row = blockIdx.y;
  const unsigned splits = gridDim.x;
  const long long c0 = static_cast<long long>(blockIdx.x) * seg;
  long long len = cols - c0;
  if (len > seg) len = seg;
  Acc v = Op::identity();
  if (len > 0) v = reduce_span<T, Acc, Op, kBlock, U>(x + row * cols + c0, len, threadIdx.x);
  v = block_combine<Acc, Op>(v, smem);
  if (threadIdx.x == 0) {
    partial[row * splits + blockIdx.x] = v;
    __threadfence();
    s_ticket = atomicAdd(&tickets[row], 1u);
  }
  __syncthreads();
  if (s_ticket != splits - 1) return;
  __threadfence();
  Acc r = Op::identity();
  for (unsigned s = threadIdx.x; s < splits; s += kBlock)
    r = Op::combine(r, __ldcg(&partial[row * splits + s]));
  r = block_combine<Acc, Op>(r, smem);
  if (threadIdx.x == 0) {
    store_final_16<Op, T>(ep, row, r);
    tickets[row] = 0;
  }
}

// Batches of UU row loads for the column this thread owns.  The address advances by
// pointer increments with a precomputed byte stride: a fresh 64-bit multiply per load makes
// ptxas interleave address math with the loads and only ~2 stay in flight (measured:
// 6.3 TB/s vs 7.3 TB/s on config C5, profiles/r01_col_reduce.md).
template <typename T, typename Acc, typename Op, int VE, int UU>
__device__ __forceinline__ void col_step(const unsigned char* __restrict__& p, size_t stride_bytes,
                                         long long& left, Acc (&acc)[VE]) {
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
        for (int e = 0; e < VE; ++e) acc[e] = Op::combine(acc[e], load_as<Acc, Op>(w.e[e]));
      }
    } else {
      T buf[UU];
#pragma unroll
      for (int u = 0; u < UU; ++u) buf[u] = ldg_scalar(reinterpret_cast<const T*>(p + u * stride_bytes));
      p += UU * stride_bytes;
#pragma unroll
      for (int u = 0; u < UU; ++u) acc[0] = Op::combine(acc[0], load_as<Acc, Op>(buf[u]));
    }
  }
}

// [outer, R, inner] -> [outer, inner].  Block = TX x TY (TX*TY == kBlock): tx walks
// the kept contiguous axis in units of VE elements (VE = 16/sizeof(T) when inner
// allows 16-byte loads, else 1), ty walks the reduced axis.  grid.x = outer * tiles,
// grid.y = splits of R.  U row-loads are in flight per thread.
template <typename T, typename Acc, typename Op, int VE, int U>
__global__ void __launch_bounds__(kBlock)
    col_reduce_kernel(const T* __restrict__ x, long long R, long long inner, int TX,
                      long long tiles, long long rows_per_split, Acc* __restrict__ partial,
                      unsigned* __restrict__ tickets, Epilogue ep) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Acc* smem = reinterpret_cast<Acc*>(smem_raw);  // [TY][TX][VE]
  __shared__ unsigned s_ticket;
  // A kernel queued behind this one with programmatic dependent launch (the peer exchange of
  // b200phi_sharded_reduce_p2p) may be scheduled now; it waits for this grid to finish before
  // it reads anything.  No effect on an ordinary launch.
  asm volatile("griddepcontrol.launch_dependents;");
  const int TY = kBlock / TX;
  const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
  const long long o = blockIdx.x / tiles, tile = blockIdx.x % tiles;
  const long long inner_v = inner / VE;  // VE divides inner on the vector path
  const long long cv = tile * TX + tx;   // column (in vectors) this thread owns
  const bool live = cv < inner_v;
  const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_split;
  long long r1 = r0 + rows_per_split;
  if (r1 > R) r1 = R;
  const unsigned splits = gridDim.y;

  Acc acc[VE];
#pragma unroll
  for (int e = 0; e < VE; ++e) acc[e] = Op::identity();

  if (live && r0 + ty < r1) {
    const unsigned char* __restrict__ p =
        reinterpret_cast<const unsigned char*>(x + (o * R + r0 + ty) * inner + cv * VE);
    const size_t stride_bytes = static_cast<size_t>(TY) * inner * sizeof(T);
    long long left = (r1 - r0 - ty + TY - 1) / TY;  // rows this thread visits
    col_step<T, Acc, Op, VE, U>(p, stride_bytes, left, acc);
    if constexpr (U > 4) col_step<T, Acc, Op, VE, 4>(p, stride_bytes, left, acc);
    if constexpr (U > 2) col_step<T, Acc, Op, VE, 2>(p, stride_bytes, left, acc);
    if constexpr (U > 1) col_step<T, Acc, Op, VE, 1>(p, stride_bytes, left, acc);
  }

  // combine across ty through shared memory (tree over TY, TY is a power of two)
#pragma unroll
  for (int e = 0; e < VE; ++e) smem[(ty * TX + tx) * VE + e] = acc[e];
  __syncthreads();
  for (int half = TY >> 1; half > 0; half >>= 1) {
    if (ty < half) {
#pragma unroll
      for (int e = 0; e < VE; ++e)
        smem[(ty * TX + tx) * VE + e] = Op::combine(smem[(ty * TX + tx) * VE + e],
                                                    smem[((ty + half) * TX + tx) * VE + e]);
    }
    __syncthreads();
  }

  const long long out_base = o * inner + cv * VE;
  if (splits == 1) {
    if (ty == 0 && live) {
#pragma unroll
      for (int e = 0; e < VE; ++e) store_final_16<Op, T>(ep, out_base + e, smem[tx * VE + e]);
    }
    return;
  }

  const long long out_numel = static_cast<long long>(gridDim.x / tiles) * inner;
  if (ty == 0 && live) {
#pragma unroll
    for (int e = 0; e < VE; ++e)
      partial[static_cast<long long>(blockIdx.y) * out_numel + out_base + e] = smem[tx * VE + e];
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(&tickets[blockIdx.x], 1u);
  __syncthreads();
  if (s_ticket != splits - 1) return;
  __threadfence();

  // last CTA for this (outer, tile): combine the `splits` partials in index order.
  // Loads are issued in batches of 8 before the first use: the partials sit in L2 and a
  // dependent chain of `splits` loads would put their latencies end to end.
#pragma unroll
  for (int e = 0; e < VE; ++e) acc[e] = Op::identity();
  if (live) {
    unsigned s = ty;
    for (; s + 7u * TY < splits; s += 8u * TY) {
      Acc b[8][VE];
#pragma unroll
      for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int e = 0; e < VE; ++e)
          b[u][e] = __ldcg(&partial[static_cast<long long>(s + u * TY) * out_numel + out_base + e]);
#pragma unroll
      for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int e = 0; e < VE; ++e) acc[e] = Op::combine(acc[e], b[u][e]);
    }
    for (; s < splits; s += TY) {
#pragma unroll
      for (int e = 0; e < VE; ++e)
        acc[e] = Op::combine(acc[e], __ldcg(&partial[static_cast<long long>(s) * out_numel + out_base + e]));
    }
  }
#pragma unroll
  for (int e = 0; e < VE; ++e) smem[(ty * TX + tx) * VE + e] = acc[e];
  __syncthreads();
  for (int half = TY >> 1; half > 0; half >>= 1) {
    if (ty < half) {
#pragma unroll
      for (int e = 0; e < VE; ++e)
        smem[(ty * TX + tx) * VE + e] = Op::combine(smem[(ty * TX + tx) * VE + e],
                                                    smem[((ty + half) * TX + tx) * VE + e]);
    }
    __syncthreads();
  }
  if (ty == 0 && live) {
#pragma unroll
    for (int e = 0; e < VE; ++e) store_final_16<Op, T>(ep, out_base + e, smem[tx * VE + e]);
  }
  if (threadIdx.x == 0) tickets[blockIdx.x] = 0;
}

// Arbitrary keep/reduce interleaving (up to 9 merged dims).  The host hands over
// the kept dims and the reduced dims separately, each with extents and input
// strides, innermost last (the job of IndexCalculator, index_calculator.h:53-107).
struct GenericDims {
  int nkeep, nred;
  long long keep_extent[kMaxRank], keep_stride[kMaxRank];
  long long red_extent[kMaxRank], red_stride[kMaxRank];
  long long out_numel, red_numel;
};

__device__ __forceinline__ long long unravel(long long idx, int n, const long long* extent,
                                             const long long* stride) {
  long long off = 0;
  for (int d = n - 1; d >= 0; --d) {
    const long long e = extent[d];
    long long q, rem;
    if ((static_cast<unsigned long long>(idx) | static_cast<unsigned long long>(e)) >> 32) {
      q = idx / e;
      rem = idx - q * e;
    } else {
      const unsigned qi = static_cast<unsigned>(idx) / static_cast<unsigned>(e);
      q = qi;
      rem = static_cast<unsigned>(idx) - qi * static_cast<unsigned>(e);
    }
    off += rem * stride[d];
    idx = q;
  }
  return off;
}

// Block = TX x TY.  tx walks outputs (coalesced when the innermost merged dim is
// kept), ty walks the flattened reduce space (coalesced when it is reduced; the
// host then picks TX == 1).
template <typename T, typename Acc, typename Op>
__global__ void __launch_bounds__(kBlock)
    generic_reduce_kernel(const T* __restrict__ x, GenericDims gd, int TX, Epilogue ep) {
  __shared__ Acc smem[kBlock];
  const int TY = kBlock / TX;
  const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
  const long long o = static_cast<long long>(blockIdx.x) * TX + tx;
  Acc acc = Op::identity();
  if (o < gd.out_numel) {
    const long long base = unravel(o, gd.nkeep, gd.keep_extent, gd.keep_stride);
    // walk the innermost reduced dim directly, unravel only the outer part
    const long long in_e = gd.nred > 0 ? gd.red_extent[gd.nred - 1] : 1;
    const long long in_s = gd.nred > 0 ? gd.red_stride[gd.nred - 1] : 0;
    const long long outer_n = gd.red_numel / in_e;
    if (in_e >= TY || outer_n == 1) {
      for (long long jo = 0; jo < outer_n; ++jo) {
        const long long ob = base + unravel(jo, gd.nred - 1, gd.red_extent, gd.red_stride);
        for (long long ji = ty; ji < in_e; ji += TY)
          acc = Op::combine(acc, load_as<Acc, Op>(ldg_scalar(x + ob + ji * in_s)));
      }
    } else {
      for (long long j = ty; j < gd.red_numel; j += TY) {
        const long long off = unravel(j, gd.nred, gd.red_extent, gd.red_stride);
        acc = Op::combine(acc, load_as<Acc, Op>(ldg_scalar(x + base + off)));
      }
    }
  }
  smem[ty * TX + tx] = acc;
  __syncthreads();
  for (int half = TY >> 1; half > 0; half >>= 1) {
    if (ty < half)
      smem[ty * TX + tx] = Op::combine(smem[ty * TX + tx], smem[(ty + half) * TX + tx]);
    __syncthreads();
  }
  if (ty == 0 && o < gd.out_numel) store_final_16<Op, T>(ep, o, smem[tx]);
}

}  // namespace b200phi
