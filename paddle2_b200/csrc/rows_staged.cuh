// Short rows of any alignment through shared memory: a THREAD per row.
//
// Rows of a few hundred bytes that are not whole aligned 16-byte vectors (a 101-column bf16 row
// is 202 bytes and starts on a different byte offset in each of eight consecutive rows) cost the
// sub-warp kernels ~45 warp instructions per row -- head / tail element loads, per-slot
// predicates, two shuffle trees for every four rows -- and those kernels are bound by instruction
// issue at 2.5-4.8 TB/s (ncu: 74-87 % issue-slot use).  Here a CTA copies kStThreads consecutive
// rows -- one contiguous byte span, whatever the rows' alignment -- into shared memory with ONE
// bulk asynchronous copy (TMA, 1-D form: cp.async.bulk + mbarrier transaction bytes; no thread
// spends issue slots on the copy), and each thread then walks its own row in shared memory: no
// idle lanes, no shuffles, no predicates per slot, coalesced result stores.  A ring of `stages`
// spans keeps the next copies in flight while a span is consumed.
//
// The bulk copy needs 16-byte-aligned ends, so a span is widened to the enclosing aligned window;
// the window only leaves the tensor at the tensor's own two ends, where up to 15 bytes are copied
// with byte loads instead.
//
// Rows longer than ~64 bytes get L = 2 or 4 lanes each (contiguous L-ths of the row, joined by
// one or two xor shuffles): what bounds the kernel is the shared memory a resident thread pins
// (stages * row_bytes / L), i.e. how many warps an SM holds to hide the shared-memory latency of
// the walk (measured on 101 bf16 columns: 2 CTAs of 128 threads per SM 4.7 TB/s, 4 CTAs 6.1).
//
// A consumer C provides
//   using Params = ...;                                   (passed by value to the kernel)
//   template <int L> static __device__ void run(const T* row_in_smem, int cols, int i0, int i1, int rot,
//                                               bool live, long long row, int h, const Params&);
// for lane h of the row's L lanes and its element range [i0, i1); every thread of the CTA calls
// it (dead rows with i0 == i1 and live == false), so that it may use group_all<L>.
//
// Rows that ARE whole 16-byte vectors (and start on 16-byte boundaries) are read with 16-byte
// shared loads, which are served 8 threads at a time from 8 groups of 4 banks: with a pitch of nv
// vectors, threads that all start at vector 0 share 8 / gcd(nv, 8) groups (a 128-byte row: one
// group, an 8-way conflict).  Such rows are visited from a staggered start and wrapped around
// (smem_visit): thread t starts floor((t mod 8) * gcd(nv, 8) / 8) vectors into its row, which
// puts the 8 threads of a phase on 8 different groups.  Consumers whose result depends on the
// visiting order (argmax ties) compare indices explicitly then.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "prim.cuh"
#include "reduce_plan.h"
#include "tma.cuh"

namespace b200phi {

template <typename T, typename C, int L>
__global__ void __launch_bounds__(kStThreads * L)
    rows_staged_kernel(const T* __restrict__ x, long long rows, int cols, int stages, int stage_bytes,
                       int rot_share, int rpt, typename C::Params prm) {
  extern __shared__ __align__(128) unsigned char st_smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(st_smem + static_cast<size_t>(stages) * stage_bytes);
  const int tid = threadIdx.x;
  const int r = tid / L, h = tid % L;
  const long long row_bytes = static_cast<long long>(cols) * static_cast<long long>(sizeof(T));
  // rpt rows per thread and span (very short rows: a span's fixed costs -- barrier, mbarrier wait,
  // window arithmetic -- would otherwise be paid once per 16-byte row); thread t takes rows
  // t, t + 128, ... of the span, so the bank pattern and the coalesced result stores stay those of
  // one row per thread
  const int span_rows = kStThreads * rpt;
  const long long ntiles = (rows + span_rows - 1) / span_rows;
  const long long my = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  const uintptr_t xb = reinterpret_cast<uintptr_t>(x);
  const uintptr_t xe = xb + static_cast<uintptr_t>(rows * row_bytes);
  const int chunk = (cols + L - 1) / L;
  const int i0 = h * chunk < cols ? h * chunk : cols;
  const int i1 = i0 + chunk < cols ? i0 + chunk : cols;
  // staggered start, in 16-byte vectors (smem_visit), for rows that are whole aligned vectors:
  // rot_share = gcd(vectors per row, 8) threads of a load phase share a bank group when they all
  // start at vector 0, so thread t starts floor((t mod 8) * rot_share / 8) vectors in.
  // rot_share = 0 for every other row (rot = -1: plain walk).
  int rot = -1;
  if (L == 1 && rot_share > 0) rot = (((tid & 7) * rot_share) >> 3) % (static_cast<int>(row_bytes) >> 4);

  // Start the copy of this CTA's i-th span into stage i % stages (every thread calls).
  auto issue = [&](long long i) {
    const int s = static_cast<int>(i % stages);
    const long long row0 = (blockIdx.x + i * static_cast<long long>(gridDim.x)) * span_rows;
    const long long left = rows - row0;
    const int n = left < span_rows ? static_cast<int>(left) : span_rows;
    const uintptr_t b0 = xb + static_cast<uintptr_t>(row0 * row_bytes);
    const uintptr_t b1 = b0 + static_cast<uintptr_t>(n * row_bytes);
    const uintptr_t a0 = b0 & ~static_cast<uintptr_t>(15);
    const uintptr_t a1 = (b1 + 15) & ~static_cast<uintptr_t>(15);
    uintptr_t lo = a0 < xb ? ((b0 + 15) & ~static_cast<uintptr_t>(15)) : a0;
    uintptr_t hi = a1 > xe ? (b1 & ~static_cast<uintptr_t>(15)) : a1;
    if (hi < lo) hi = lo;
    unsigned char* dst = st_smem + static_cast<size_t>(s) * stage_bytes;
    if (tid == 0) {
      mbar_arrive_expect_tx(&full[s], static_cast<uint32_t>(hi - lo));
      if (hi > lo) bulk_load(dst + (lo - a0), reinterpret_cast<const void*>(lo), static_cast<uint32_t>(hi - lo), &full[s]);
    }
    // the tensor's own first / last bytes when they are not 16-byte aligned (at most 15 each)
    const uintptr_t e0 = lo < b1 ? lo : b1;
    if (b0 + tid < e0) dst[b0 + tid - a0] = *reinterpret_cast<const unsigned char*>(b0 + tid);
    const uintptr_t t0 = hi > e0 ? hi : e0;
    if (t0 + tid < b1) dst[t0 + tid - a0] = *reinterpret_cast<const unsigned char*>(t0 + tid);
  };

  for (long long i = 0; i < stages - 1 && i < my; ++i) issue(i);
  for (long long i = 0; i < my; ++i) {
    __syncthreads();  // span i - 1 has been consumed by every thread: its stage is free again
    if (i + stages - 1 < my) issue(i + stages - 1);
    if (stages == 1) __syncthreads();  // (the byte-copied edges of THIS span were stored just now)
    const int s = static_cast<int>(i % stages);
    mbar_wait(&full[s], static_cast<uint32_t>((i / stages) & 1));
    const long long row0 = (blockIdx.x + i * static_cast<long long>(gridDim.x)) * span_rows;
    const uintptr_t b0 = xb + static_cast<uintptr_t>(row0 * row_bytes);
    const unsigned char* base = st_smem + static_cast<size_t>(s) * stage_bytes + (b0 & 15);
    for (int j = 0; j < rpt; ++j) {
      const int rs = j * kStThreads + r;  // row of the span
      const bool live = row0 + rs < rows;
      const unsigned char* sp = base + static_cast<size_t>(rs) * static_cast<size_t>(row_bytes);
      C::template run<L>(reinterpret_cast<const T*>(sp), cols, live ? i0 : 0, live ? i1 : 0, rot, live, row0 + rs, h, prm);
    }
  }
}

int current_sm_count();
int cuda_status(cudaError_t e);

// stages: B200PHI_STAGED_STAGES (2..4, default 2), read once; lanes per row: B200PHI_STAGED_LANES
// (1, 2 or 4; default by row length).
int staged_stages();
int staged_lanes();
int staged_rows_per_thread();  // B200PHI_STAGED_RPT (1..8; 0 / unset: about 256 bytes of rows per thread)

template <typename T, typename C, int L>
int launch_rows_staged_l(const T* x, long long rows, int cols, const typename C::Params& prm, cudaStream_t stream) {
  const long long row_bytes = static_cast<long long>(cols) * static_cast<long long>(sizeof(T));
  // rows per thread: about 256 bytes of rows per thread and span (one lane per row only)
  int rpt = 1;
  if (L == 1) {
    const int want = staged_rows_per_thread();
    rpt = want > 0 ? want : static_cast<int>(256 / row_bytes);
    rpt = rpt < 1 ? 1 : (rpt > 8 ? 8 : rpt);
    while (rpt > 1 && row_bytes * rpt * kStThreads > 48 * 1024) --rpt;  // (a span stays under 48 KiB)
  }
  const int stage_bytes = staged_stage_bytes(row_bytes * rpt);
  int stages = staged_stages();
  while (stages > 1 && static_cast<long long>(stages) * stage_bytes > 100 * 1024) --stages;
  const size_t smem = static_cast<size_t>(stages) * stage_bytes + 64;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    cudaError_t ea = cudaFuncSetAttribute(rows_staged_kernel<T, C, L>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (ea != cudaSuccess) return cuda_status(ea);
    attr_set[dev] = true;
  }
  const long long ntiles = (rows + static_cast<long long>(kStThreads) * rpt - 1) / (static_cast<long long>(kStThreads) * rpt);
  long long per_sm = static_cast<long long>(220 * 1024) / static_cast<long long>(smem + 1024);
  const long long by_threads = 2048 / (kStThreads * L);
  if (per_sm > by_threads) per_sm = by_threads;
  if (per_sm < 1) per_sm = 1;
  long long ctas = static_cast<long long>(current_sm_count()) * per_sm;
  if (ctas > ntiles) ctas = ntiles;
  int rot_share = 0;  // rows that are whole vectors AND start on 16-byte boundaries: staggered vector walk
  if (row_bytes % 16 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0) {
    const long long nv = row_bytes / 16;
    rot_share = (nv % 8 == 0) ? 8 : ((nv % 4 == 0) ? 4 : ((nv % 2 == 0) ? 2 : 1));
  }
  rows_staged_kernel<T, C, L><<<static_cast<unsigned>(ctas), kStThreads * L, smem, stream>>>(x, rows, cols, stages, stage_bytes, rot_share, rpt, prm);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return 0;
}

template <typename T, typename C>
int launch_rows_staged(const T* x, long long rows, int cols, const typename C::Params& prm, cudaStream_t stream) {
  const long long row_bytes = static_cast<long long>(cols) * static_cast<long long>(sizeof(T));
  int lanes = staged_lanes();
  if (lanes == 0 || row_bytes % 16 == 0) lanes = 1;  // (the staggered start is per row)
  if (lanes == 1) return launch_rows_staged_l<T, C, 1>(x, rows, cols, prm, stream);
  if (lanes == 2) return launch_rows_staged_l<T, C, 2>(x, rows, cols, prm, stream);
  return launch_rows_staged_l<T, C, 4>(x, rows, cols, prm, stream);
}

}  // namespace b200phi
