// Fused sharded reduce: the local column reduce and the cross-GPU exchange in ONE kernel.
// Each rank reduces its row shard; the thread block that finishes an output tile stores it
// straight into every peer's memory over NVLink (CUDA IPC mappings), publishes an epoch flag
// and combines the ranks' slots in rank order (reduce_kernels.cuh, p2p_finish).  No NCCL call,
// no second launch: the collective latency left on config C5 is one NVLink round trip.
// Replaces SumKernel + ProcessGroupNCCL::AllReduce (process_group_nccl.cc:267-298) like
// b200phi_sharded_reduce does, which stays as the general (any dtype / any shape) route.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <string>

#include "../../include/b200phi.h"
#include "reduce_launch.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int current_sm_count();

static size_t p2p_slots_bytes(int nranks, long long max_out) {
  return static_cast<size_t>(2) * nranks * max_out * 8;
}

template <typename T, typename Acc, typename Op>
static int launch_col_p2p(const ReducePlan& p, const b200phi_reduce_desc* d, const Epilogue& ep,
                          const P2PArgs& px, cudaStream_t stream) {
  Acc* partial = static_cast<Acc*>(d->workspace);
  unsigned* tickets =
      reinterpret_cast<unsigned*>(static_cast<unsigned char*>(d->workspace) + p.ws_partial_bytes);
  if (p.ws_ticket_bytes) {
    cudaError_t e = cudaMemsetAsync(tickets, 0, p.ws_ticket_bytes, stream);
    if (e != cudaSuccess) return cuda_status(e);
  }
  const size_t smem = static_cast<size_t>(kBlock) * p.vec * sizeof(Acc);
  col_reduce_kernel<T, Acc, Op, VecOf<T>::kElems, 8, P2PArgs>
      <<<dim3(p.grid_x, p.grid_y), kBlock, smem, stream>>>(static_cast<const T*>(d->x), p.R, p.inner, p.TX,
                                                           p.tiles, p.rows_per_split, partial, tickets, ep, px);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

int run_sharded_reduce_p2p(const b200phi_reduce_desc* d, long long global_count,
                           const b200phi_p2p_group* g, unsigned epoch, cudaStream_t stream) {
  if (!g || g->nranks < 1 || g->nranks > 8 || g->rank < 0 || g->rank >= g->nranks)
    return set_error(B200PHI_INVALID_ARGUMENT, "bad p2p group");
  if (epoch == 0) return set_error(B200PHI_INVALID_ARGUMENT, "epoch starts at 1 and grows by 1 per call");
  ReducePlan plan;
  std::string err;
  int rc = plan_reduce(d, default_tunables(current_sm_count()), &plan, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  if (!d->x || !d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  if (plan.kind != kCol || plan.vec <= 1)
    return set_error(B200PHI_UNSUPPORTED,
                     "fused p2p reduce covers the vectorised column reduce; use b200phi_sharded_reduce");
  if (plan.out_numel > g->max_out_numel || static_cast<long long>(plan.grid_x) > g->max_tiles)
    return set_error(B200PHI_WORKSPACE_TOO_SMALL, "p2p group buffers are too small for this reduce");
  if (plan.ws_total_bytes > 0 && (!d->workspace || d->workspace_bytes < plan.ws_total_bytes))
    return set_error(B200PHI_WORKSPACE_TOO_SMALL, "workspace too small");
  P2PArgs px;
  px.nranks = g->nranks;
  px.rank = g->rank;
  px.parity = static_cast<int>(epoch & 1u);
  px.epoch = epoch;
  px.max_out = g->max_out_numel;
  px.max_tiles = g->max_tiles;
  for (int r = 0; r < 8; ++r) px.base[r] = r < g->nranks ? static_cast<unsigned char*>(g->base[r]) : nullptr;
  px.flags_offset = p2p_slots_bytes(g->nranks, g->max_out_numel);
  for (int r = 0; r < g->nranks; ++r)
    if (!px.base[r]) return set_error(B200PHI_INVALID_ARGUMENT, "p2p group has an unmapped peer");
  const bool mean = d->op == B200PHI_MEAN;
  if (mean && global_count <= 0) return set_error(B200PHI_INVALID_ARGUMENT, "global_reduce_count must be > 0");
  const Epilogue ep = make_epilogue(d->out, plan.out_dtype, mean, mean ? global_count : plan.red_numel);
  const int op = d->op;
  const bool sum = op == B200PHI_SUM || mean;
  const bool mx = op == B200PHI_MAX || op == B200PHI_AMAX;
  const bool mn = op == B200PHI_MIN || op == B200PHI_AMIN;
  if (plan.acc == 0) {
    if (d->in_dtype == B200PHI_FLOAT32) {
      if (sum) return launch_col_p2p<float, float, SumOp<float>>(plan, d, ep, px, stream);
      if (mx) return launch_col_p2p<float, float, MaxOp<float>>(plan, d, ep, px, stream);
      if (mn) return launch_col_p2p<float, float, MinOp<float>>(plan, d, ep, px, stream);
    } else if (d->in_dtype == B200PHI_BFLOAT16 && sum) {
      return launch_col_p2p<__nv_bfloat16, float, SumOp<float>>(plan, d, ep, px, stream);
    }
  } else if (plan.acc == 1 && d->in_dtype == B200PHI_FLOAT64 && sum) {
    return launch_col_p2p<double, double, SumOp<double>>(plan, d, ep, px, stream);
  } else if (plan.acc == 2 && d->in_dtype == B200PHI_INT64 && sum) {
    return launch_col_p2p<long long, long long, SumOp<long long>>(plan, d, ep, px, stream);
  }
  return set_error(B200PHI_UNSUPPORTED, "fused p2p reduce: dtype/op not built; use b200phi_sharded_reduce");
}

}  // namespace b200phi

using namespace b200phi;

extern "C" {

size_t b200phi_p2p_buffer_bytes(int nranks, int64_t max_out_numel, int64_t max_tiles) {
  if (nranks < 1 || nranks > 8 || max_out_numel < 1 || max_tiles < 1) return 0;
  return p2p_slots_bytes(nranks, max_out_numel) + static_cast<size_t>(nranks) * max_tiles * 4 + 256;
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
  if (e != cudaSuccess) {
    cudaFree(p);
    return cuda_status(e);
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  memcpy(handle, &h, 64);
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
  cudaError_t e = cudaFree(ptr);
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

int b200phi_sharded_reduce_p2p(const b200phi_reduce_desc* d, int64_t global_reduce_count,
                               const b200phi_p2p_group* g, uint32_t epoch, void* stream) {
  return run_sharded_reduce_p2p(d, global_reduce_count, g, epoch, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
