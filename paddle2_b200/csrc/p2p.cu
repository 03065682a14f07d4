// Sharded reduce without a collective library: the local reduce, then a small exchange kernel
// over NVLink peer memory (CUDA IPC mappings).
//
//   1. the ordinary reduce kernels (byte for byte the ones b200phi_reduce launches) write this
//      rank's partial -- in the accumulation type, so a bf16 / fp16 result is rounded once, at the
//      very end -- straight into the rank's own slot of the exchange buffer;
//   2. p2p_exchange_kernel, a handful of CTAs launched with programmatic dependent launch so that
//      its launch latency hides under the reduce's tail (it parks on griddepcontrol.wait): each CTA
//      pushes its slice of the partial into every peer's buffer, publishes an epoch flag with a
//      system-scope release store, waits for the peers' flags of the same slice and combines the
//      nranks slots in rank order, so every rank ends with identical bits.
//
// A peer that never publishes (not launched, failed its own argument checks, died) is a
// TIMEOUT: the CTA writes nothing to `out`, records the epoch in a host-mapped status word,
// and the next call on this group -- or b200phi_p2p_status -- reports B200PHI_PEER_TIMEOUT.
// Nothing is ever combined from slots that were not published.
//
// Replaces SumKernel + ProcessGroupNCCL::AllReduce (process_group_nccl.cc:267-298) like
// b200phi_sharded_reduce does.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <map>
#include <mutex>
#include <string>

#include "../../include/b200phi.h"
#include "prim.cuh"
#include "reduce_plan.h"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();
int run_reduce(const b200phi_reduce_desc* d, cudaStream_t stream);
Tunables tunables();

namespace {

constexpr int kMaxExchangeCtas = 64;  // all of them co-resident on any device, whatever else runs

// Every rank's buffer (all ranks can address all buffers):
//   slots [2 parities][nranks][cells] 8-byte cells   rank r's partial, written BY rank r
//   flags [nranks][max_tiles] uint32                 epoch of rank r's last publication, per CTA
long long slot_cells(long long max_out) { return (max_out + 1) & ~1LL; }  // 16-byte multiple
size_t slots_bytes(int nranks, long long max_out) {
  return static_cast<size_t>(2) * nranks * slot_cells(max_out) * 8;
}

struct ExchangeArgs {
  int nranks, rank, parity;
  unsigned epoch;
  long long n;           // elements of the partial
  long long slot_bytes;  // bytes between two ranks' slots
  long long max_tiles;
  unsigned long long flags_offset;
  unsigned long long timeout_ns;
  unsigned char* base[8];
  unsigned* status;  // host-mapped: 0, or the epoch that timed out
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// slots are written by peers through this GPU's L2: never serve them from L1
__device__ __forceinline__ V16 ld_volatile16(const void* p) {
  V16 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3])
               : "l"(p)
               : "memory");
  return r;
}

// P: storage type of the partial; Acc / Op: how partials combine.  One 16-byte unit per thread
// per round.  Indexing of a.base[] is by unrolled constant index only (a run-time index would
// put the argument struct in local memory).
template <typename P, typename Acc, typename Op>
__global__ void __launch_bounds__(kBlock)
    p2p_exchange_kernel(const __grid_constant__ ExchangeArgs a, const __grid_constant__ Epilogue ep) {
  constexpr int VE = 16 / sizeof(P);
  __shared__ int s_timeout;
  // the reduce in front of this kernel must have finished and its stores be visible
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const long long units = (a.n * static_cast<long long>(sizeof(P)) + 15) / 16;
  const long long per_cta = (units + gridDim.x - 1) / gridDim.x;
  const long long u0 = static_cast<long long>(blockIdx.x) * per_cta;
  long long u1 = u0 + per_cta;
  if (u1 > units) u1 = units;
  const long long mine = (static_cast<long long>(a.parity) * a.nranks + a.rank) * a.slot_bytes;
  unsigned char* own = nullptr;
#pragma unroll
  for (int p = 0; p < 8; ++p)
    if (p == a.rank) own = a.base[p];
  if (threadIdx.x == 0) s_timeout = 0;

  // push this CTA's slice of the partial into every peer's copy of this rank's slot
  for (long long u = u0 + threadIdx.x; u < u1; u += kBlock) {
    const V16 v = ld_volatile16(own + mine + u * 16);
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      if (p < a.nranks && p != a.rank) *reinterpret_cast<V16*>(a.base[p] + mine + u * 16) = v;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < a.nranks) {
    unsigned char* peer = nullptr;
#pragma unroll
    for (int p = 0; p < 8; ++p)
      if (p == static_cast<int>(threadIdx.x)) peer = a.base[p];
    unsigned* f = reinterpret_cast<unsigned*>(peer + a.flags_offset) +
                  static_cast<long long>(a.rank) * a.max_tiles + blockIdx.x;
    st_release_sys(f, a.epoch);
    const unsigned* wait_on = reinterpret_cast<const unsigned*>(own + a.flags_offset) +
                              static_cast<long long>(threadIdx.x) * a.max_tiles + blockIdx.x;
    const unsigned long long t0 = global_timer_ns();
    while (static_cast<int>(ld_acquire_sys(wait_on) - a.epoch) < 0) {
      if (global_timer_ns() - t0 > a.timeout_ns) {
        s_timeout = 1;
        break;
      }
    }
  }
  __syncthreads();
  if (s_timeout) {  // a peer never published: report, and combine NOTHING
    if (threadIdx.x == 0) {
      *reinterpret_cast<volatile unsigned*>(a.status) = a.epoch;
      __threadfence_system();
    }
    return;
  }
  for (long long u = u0 + threadIdx.x; u < u1; u += kBlock) {
    Acc r[VE];
#pragma unroll
    for (int e = 0; e < VE; ++e) r[e] = Op::identity();
    for (int k = 0; k < a.nranks; ++k) {  // rank order: the same on every rank
      typename VecOf<P>::U w;
      w.raw = ld_volatile16(own + (static_cast<long long>(a.parity) * a.nranks + k) * a.slot_bytes + u * 16);
#pragma unroll
      for (int e = 0; e < VE; ++e) r[e] = Op::combine(r[e], Cvt<Acc, P>::to(w.e[e]));
    }
#pragma unroll
    for (int e = 0; e < VE; ++e) {
      const long long i = u * VE + e;
      if (i < a.n) store_result<Acc>(ep, i, r[e]);
    }
  }
}

template <typename P, typename Acc, typename Op>
int launch_exchange(const ExchangeArgs& a, const Epilogue& ep, unsigned grid, cudaStream_t stream) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kBlock);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, p2p_exchange_kernel<P, Acc, Op>, a, ep);
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

template <typename P, typename Acc>
int dispatch_exchange_op(int op, const ExchangeArgs& a, const Epilogue& ep, unsigned grid, cudaStream_t stream) {
  switch (op) {
    case B200PHI_SUM:
    case B200PHI_MEAN:
      return launch_exchange<P, Acc, SumOp<Acc>>(a, ep, grid, stream);
    case B200PHI_PROD:
      return launch_exchange<P, Acc, ProdOp<Acc>>(a, ep, grid, stream);
    case B200PHI_MAX:
    case B200PHI_AMAX:
    case B200PHI_ANY:  // on 0/1 bytes
      return launch_exchange<P, Acc, MaxOp<Acc>>(a, ep, grid, stream);
    default:  // MIN, AMIN, ALL
      return launch_exchange<P, Acc, MinOp<Acc>>(a, ep, grid, stream);
  }
}

int dispatch_exchange(int partial_dtype, int op, const ExchangeArgs& a, const Epilogue& ep, unsigned grid,
                      cudaStream_t stream) {
  switch (partial_dtype) {
    case B200PHI_FLOAT32:
      return dispatch_exchange_op<float, float>(op, a, ep, grid, stream);
    case B200PHI_FLOAT64:
      return dispatch_exchange_op<double, double>(op, a, ep, grid, stream);
    case B200PHI_INT64:
      return dispatch_exchange_op<long long, long long>(op, a, ep, grid, stream);
    case B200PHI_INT32:
      return dispatch_exchange_op<int32_t, long long>(op, a, ep, grid, stream);
    case B200PHI_INT16:
      return dispatch_exchange_op<int16_t, long long>(op, a, ep, grid, stream);
    case B200PHI_INT8:
      return dispatch_exchange_op<int8_t, long long>(op, a, ep, grid, stream);
    case B200PHI_UINT8:
    case B200PHI_BOOL:
      return dispatch_exchange_op<uint8_t, long long>(op, a, ep, grid, stream);
    case B200PHI_BFLOAT16:  // max / min only: sums travel as FLOAT32
      return dispatch_exchange_op<__nv_bfloat16, float>(op, a, ep, grid, stream);
    case B200PHI_FLOAT16:
      return dispatch_exchange_op<__half, float>(op, a, ep, grid, stream);
    default:
      return set_error(B200PHI_UNSUPPORTED, "p2p exchange: partial dtype not built");
  }
}

// ---- buffers allocated by b200phi_p2p_alloc and their host-mapped status words ----------------
struct OwnBuffer {
  unsigned* status_host;
  unsigned* status_dev;
};
std::mutex g_mu;
std::map<void*, OwnBuffer> g_own;

bool find_own(void* p, OwnBuffer* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_own.find(p);
  if (it == g_own.end()) return false;
  *out = it->second;
  return true;
}

int timeout_error(unsigned epoch) {
  char buf[200];
  snprintf(buf, sizeof(buf),
           "p2p exchange: a peer did not publish its partial for epoch %u before the timeout; no "
           "result was written for that call and the group is out of step (tear it down)",
           epoch);
  return set_error(B200PHI_PEER_TIMEOUT, buf);
}

}  // namespace

// The dtype a rank's partial travels in: the accumulation type for sums into a 16-bit float
// (one rounding, after the ranks are combined), the output dtype otherwise.
int sharded_partial_dtype(int op, int out_dtype) {
  const bool sum_like = op == B200PHI_SUM || op == B200PHI_MEAN;
  if (sum_like && (out_dtype == B200PHI_FLOAT16 || out_dtype == B200PHI_BFLOAT16)) return B200PHI_FLOAT32;
  return out_dtype;
}

int run_sharded_reduce_p2p(const b200phi_reduce_desc* d, long long global_count,
                           const b200phi_p2p_group* g, unsigned epoch, cudaStream_t stream) {
  if (!g || g->nranks < 1 || g->nranks > 8 || g->rank < 0 || g->rank >= g->nranks)
    return set_error(B200PHI_INVALID_ARGUMENT, "bad p2p group");
  if (epoch == 0) return set_error(B200PHI_INVALID_ARGUMENT, "epoch starts at 1 and grows by 1 per call");
  if (g->max_out_numel < 1 || g->max_tiles < 1) return set_error(B200PHI_INVALID_ARGUMENT, "bad p2p group sizes");
  if (d && d->op > B200PHI_AMIN)
    return set_error(B200PHI_UNSUPPORTED, "p2p sharded reduce: the norm reductions are not built");
  ReducePlan plan;
  std::string err;
  int rc = plan_reduce(d, tunables(), &plan, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  if (!d->x || !d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  if (plan.out_numel > g->max_out_numel)
    return set_error(B200PHI_WORKSPACE_TOO_SMALL, "p2p group buffers are too small for this reduce");
  for (int r = 0; r < g->nranks; ++r)
    if (!g->base[r]) return set_error(B200PHI_INVALID_ARGUMENT, "p2p group has an unmapped peer");
  const bool mean = d->op == B200PHI_MEAN;
  if (mean && global_count <= 0) return set_error(B200PHI_INVALID_ARGUMENT, "global_reduce_count must be > 0");
  OwnBuffer own;
  if (!find_own(g->base[g->rank], &own))
    return set_error(B200PHI_INVALID_ARGUMENT, "p2p group: base[rank] was not allocated by b200phi_p2p_alloc");
  const unsigned bad = *reinterpret_cast<volatile unsigned*>(own.status_host);
  if (bad != 0) return timeout_error(bad);

  const int partial_dtype = sharded_partial_dtype(d->op, plan.out_dtype);
  const long long cells = slot_cells(g->max_out_numel);
  ExchangeArgs a;
  a.nranks = g->nranks;
  a.rank = g->rank;
  a.parity = static_cast<int>(epoch & 1u);
  a.epoch = epoch;
  a.n = plan.out_numel;
  a.slot_bytes = cells * 8;
  a.max_tiles = g->max_tiles;
  a.flags_offset = slots_bytes(g->nranks, g->max_out_numel);
  a.timeout_ns = static_cast<unsigned long long>(tunables().p2p_timeout_ms) * 1000000ull;
  for (int r = 0; r < 8; ++r) a.base[r] = r < g->nranks ? static_cast<unsigned char*>(g->base[r]) : nullptr;
  a.status = own.status_dev;

  // 1. local partial into this rank's own slot (sum for a mean: the division is by the GLOBAL count)
  b200phi_reduce_desc local = *d;
  if (mean) local.op = B200PHI_SUM;
  local.out_dtype = partial_dtype;
  local.out = a.base[a.rank] + (static_cast<long long>(a.parity) * a.nranks + a.rank) * a.slot_bytes;
  rc = run_reduce(&local, stream);
  if (rc != B200PHI_OK) return rc;

  // 2. exchange + combine; the same CTA count on every rank (it depends on the shape and the
  //    group's sizes only), one flag per CTA
  const long long units = (a.n * static_cast<long long>(dtype_size(partial_dtype)) + 15) / 16;
  long long grid = (units + kBlock - 1) / kBlock;
  if (grid > kMaxExchangeCtas) grid = kMaxExchangeCtas;
  if (grid > g->max_tiles) grid = g->max_tiles;
  if (grid < 1) grid = 1;
  const Epilogue ep = make_epilogue(d->out, plan.out_dtype, mean, mean ? global_count : 1);
  return dispatch_exchange(partial_dtype, d->op, a, ep, static_cast<unsigned>(grid), stream);
}

}  // namespace b200phi

using namespace b200phi;

extern "C" {

size_t b200phi_p2p_buffer_bytes(int nranks, int64_t max_out_numel, int64_t max_tiles) {
  if (nranks < 1 || nranks > 8 || max_out_numel < 1 || max_tiles < 1) return 0;
  return slots_bytes(nranks, max_out_numel) + static_cast<size_t>(nranks) * max_tiles * 4 + 256;
}

int b200phi_p2p_alloc(size_t bytes, void** ptr, unsigned char handle[64]) {
  if (!ptr || !handle || bytes == 0) return set_error(B200PHI_INVALID_ARGUMENT, "bad p2p alloc arguments");
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) return cuda_status(e);
  e = cudaMemset(p, 0, bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  void* host = nullptr;
  void* dev = nullptr;
  if (e == cudaSuccess) e = cudaHostAlloc(&host, 64, cudaHostAllocMapped);
  if (e == cudaSuccess) {
    memset(host, 0, 64);
    e = cudaHostGetDevicePointer(&dev, host, 0);
  }
  if (e != cudaSuccess) {
    if (host) cudaFreeHost(host);
    cudaFree(p);
    return cuda_status(e);
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  memcpy(handle, &h, 64);
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_own[p] = OwnBuffer{static_cast<unsigned*>(host), static_cast<unsigned*>(dev)};
  }
  *ptr = p;
  return B200PHI_OK;
}

int b200phi_p2p_open(const unsigned char handle[64], void** ptr) {
  if (!ptr || !handle) return set_error(B200PHI_INVALID_ARGUMENT, "bad p2p open arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return cuda_status(e);
  *ptr = p;
  return B200PHI_OK;
}

int b200phi_p2p_close(void* ptr) {
  if (!ptr) return B200PHI_OK;
  cudaError_t e = cudaIpcCloseMemHandle(ptr);
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

int b200phi_p2p_free(void* ptr) {
  if (!ptr) return B200PHI_OK;
  OwnBuffer own = {nullptr, nullptr};
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_own.find(ptr);
    if (it != g_own.end()) {
      own = it->second;
      g_own.erase(it);
    }
  }
  if (own.status_host) cudaFreeHost(own.status_host);
  cudaError_t e = cudaFree(ptr);
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

int b200phi_p2p_status(void* own_ptr) {
  OwnBuffer own;
  if (!find_own(own_ptr, &own))
    return set_error(B200PHI_INVALID_ARGUMENT, "b200phi_p2p_status: not a buffer from b200phi_p2p_alloc");
  const unsigned bad = *reinterpret_cast<volatile unsigned*>(own.status_host);
  if (bad != 0) return timeout_error(bad);
  return B200PHI_OK;
}

int b200phi_sharded_reduce_p2p(const b200phi_reduce_desc* d, int64_t global_reduce_count,
                               const b200phi_p2p_group* g, uint32_t epoch, void* stream) {
  return run_sharded_reduce_p2p(d, global_reduce_count, g, epoch, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
