// Sharded global reduce (config C5): local reduce on each GPU + allreduce of the
// small partial on the SAME stream, no host synchronisation in between.
// Replaces SumKernel + ProcessGroupNCCL::AllReduce
// (paddle/fluid/distributed/collective/process_group_nccl.cc:267-298 ->
//  paddle/phi/core/distributed/nccl_comm_context.cc:184-206) and the
// Partial->Replicate reshard (p_to_r_reshard_function.cc:67).  NCCL is resolved
// with dlopen at first use, the way phi::dynload does, so the library itself
// has no link-time NCCL dependency.
#include <dlfcn.h>
#include <nccl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

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
size_t reduce_workspace_bytes(const b200phi_reduce_desc* d);
int sharded_partial_dtype(int op, int out_dtype);

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t);
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  const char* (*GetErrorString)(ncclResult_t);
  bool ok;
  std::string why;
};

static NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    api.ok = false;
    const char* names[] = {getenv("B200PHI_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names) {
      if (!n || !*n) continue;
      h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (h) break;
    }
    if (!h) {
      api.why = std::string("cannot dlopen libnccl.so.2: ") + (dlerror() ? dlerror() : "");
      return;
    }
#define B200PHI_SYM(field, name)                                           \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, name));      \
  if (!api.field) {                                                        \
    api.why = std::string("libnccl lacks ") + name;                        \
    return;                                                                \
  }
    B200PHI_SYM(GetUniqueId, "ncclGetUniqueId")
    B200PHI_SYM(CommInitRank, "ncclCommInitRank")
    B200PHI_SYM(CommDestroy, "ncclCommDestroy")
    B200PHI_SYM(AllReduce, "ncclAllReduce")
    B200PHI_SYM(AllGather, "ncclAllGather")
    B200PHI_SYM(GetErrorString, "ncclGetErrorString")
#undef B200PHI_SYM
    api.ok = true;
  });
  return &api;
}

static int nccl_status(NcclApi* a, ncclResult_t r) {
  return set_error(B200PHI_NCCL_ERROR, std::string("NCCL: ") + a->GetErrorString(r));
}

// After the allreduce: out[i] = convert(partial[i] / count) (count == 1 unless MEAN).  `partial`
// may alias `out` when the two dtypes are the same.
template <typename P>
__global__ void __launch_bounds__(kBlock)
    finish_partial_kernel(const P* partial, long long n, Epilogue ep) {
  for (long long i = static_cast<long long>(blockIdx.x) * kBlock + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * kBlock)
    store_result<P>(ep, i, partial[i]);
}

static bool nccl_dtype(int dtype, ncclDataType_t* ty) {
  switch (dtype) {
    case B200PHI_FLOAT32: *ty = ncclFloat32; return true;
    case B200PHI_FLOAT64: *ty = ncclFloat64; return true;
    case B200PHI_FLOAT16: *ty = ncclFloat16; return true;
    case B200PHI_BFLOAT16: *ty = ncclBfloat16; return true;
    case B200PHI_INT32: *ty = ncclInt32; return true;
    case B200PHI_INT64: *ty = ncclInt64; return true;
    case B200PHI_INT8: *ty = ncclInt8; return true;
    case B200PHI_UINT8:
    case B200PHI_BOOL: *ty = ncclUint8; return true;
    default: return false;
  }
}

static size_t round256(size_t v) { return (v + 255) / 256 * 256; }

// Scratch of the sharded reduce: [wide partial, if the result is a 16-bit float sum][local reduce's own]
size_t sharded_reduce_workspace_bytes(const b200phi_reduce_desc* d) {
  ReducePlan plan;
  std::string err;
  if (plan_reduce(d, tunables(), &plan, &err) != B200PHI_OK) return 0;
  size_t need = reduce_workspace_bytes(d);
  if (sharded_partial_dtype(d->op, plan.out_dtype) != plan.out_dtype)
    need = round256(need) + round256(static_cast<size_t>(plan.out_numel) * 4);
  return need;
}

int run_sharded_reduce(const b200phi_reduce_desc* d, long long global_count, void* comm,
                       cudaStream_t stream) {
  if (!comm) return run_reduce(d, stream);
  if (d && d->op > B200PHI_AMIN)
    return set_error(B200PHI_UNSUPPORTED,
                     "sharded reduce: the norm reductions are not built (reduce locally, then "
                     "allreduce the partial with SUM / MAX / MIN)");
  NcclApi* a = nccl_api();
  if (!a->ok) return set_error(B200PHI_NCCL_ERROR, a->why);

  // ---- every argument is checked before anything is enqueued: a rank that returned an error
  //      has launched nothing, so it cannot leave a half-done result or a one-sided collective
  ReducePlan plan;
  std::string err;
  int rc = plan_reduce(d, tunables(), &plan, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  if (!d->x || !d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  const bool mean = d->op == B200PHI_MEAN;
  if (mean && global_count <= 0)
    return set_error(B200PHI_INVALID_ARGUMENT, "global_reduce_count must be > 0");
  // Sums into bf16 / fp16 travel as fp32 partials (MPType) and are rounded ONCE, after the ranks
  // are combined; an allreduce in the 16-bit type would round per rank and per NCCL step.
  const int partial_dtype = sharded_partial_dtype(d->op, plan.out_dtype);
  const bool wide = partial_dtype != plan.out_dtype;
  ncclDataType_t ty;
  if (!nccl_dtype(partial_dtype, &ty))
    return set_error(B200PHI_UNSUPPORTED, "sharded reduce: output dtype has no NCCL type");
  if (mean && partial_dtype != B200PHI_FLOAT32 && partial_dtype != B200PHI_FLOAT64 &&
      partial_dtype != B200PHI_INT64 && partial_dtype != B200PHI_INT32)
    return set_error(B200PHI_UNSUPPORTED, "sharded mean: dtype not built");
  ncclRedOp_t op;
  switch (d->op) {  // red-op mapping of gpu/all_reduce_kernel.cu:46-65
    case B200PHI_SUM:
    case B200PHI_MEAN: op = ncclSum; break;
    case B200PHI_PROD: op = ncclProd; break;
    case B200PHI_MAX:
    case B200PHI_AMAX:
    case B200PHI_ANY: op = ncclMax; break;
    default: op = ncclMin; break;  // MIN, AMIN, ALL
  }
  b200phi_reduce_desc local = *d;
  if (mean) local.op = B200PHI_SUM;  // divide by the GLOBAL count after the exchange
  void* partial = d->out;
  if (wide) {
    const size_t wide_bytes = round256(static_cast<size_t>(plan.out_numel) * 4);
    const size_t inner = reduce_workspace_bytes(d);
    if (!d->workspace || d->workspace_bytes < wide_bytes + inner ||
        reinterpret_cast<uintptr_t>(d->workspace) % 256 != 0) {
      char buf[200];
      snprintf(buf, sizeof(buf),
               "sharded reduce into a 16-bit float needs %zu bytes of 256-byte aligned workspace for "
               "the fp32 partial (b200phi_sharded_reduce_workspace), got %zu",
               wide_bytes + inner, d->workspace ? d->workspace_bytes : (size_t)0);
      return set_error(B200PHI_WORKSPACE_TOO_SMALL, buf);
    }
    partial = d->workspace;
    local.out = partial;
    local.out_dtype = partial_dtype;
    local.workspace = static_cast<unsigned char*>(d->workspace) + wide_bytes;
    local.workspace_bytes = d->workspace_bytes - wide_bytes;
  }

  // A shard with nothing to reduce locally (one row: the Partial -> Replicate reshard of an
  // already-reduced tensor) is all-reduced straight out of x: no copy in front of the collective.
  const bool plain_op = d->op == B200PHI_SUM || d->op == B200PHI_MEAN || d->op == B200PHI_MAX ||
                        d->op == B200PHI_MIN || d->op == B200PHI_PROD || d->op == B200PHI_AMAX ||
                        d->op == B200PHI_AMIN;
  const void* send = partial;
  if (plan.red_numel == 1 && plain_op && plan.in_dtype == partial_dtype) {
    send = d->x;
  } else {
    rc = run_reduce(&local, stream);
    if (rc != B200PHI_OK) return rc;
  }
  ncclResult_t r = a->AllReduce(send, partial, static_cast<size_t>(plan.out_numel), ty, op,
                                static_cast<ncclComm_t>(comm), stream);
  if (r != ncclSuccess) return nccl_status(a, r);

  if (mean || wide) {
    const long long n = plan.out_numel;
    long long blocks = (n + kBlock - 1) / kBlock;
    const long long cap = static_cast<long long>(current_sm_count()) * 8;
    if (blocks > cap) blocks = cap;
    const unsigned g = static_cast<unsigned>(blocks);
    const Epilogue ep = make_epilogue(d->out, plan.out_dtype, mean, mean ? global_count : 1);
    switch (partial_dtype) {
      case B200PHI_FLOAT32:
        finish_partial_kernel<float><<<g, kBlock, 0, stream>>>(static_cast<const float*>(partial), n, ep);
        break;
      case B200PHI_FLOAT64:
        finish_partial_kernel<double><<<g, kBlock, 0, stream>>>(static_cast<const double*>(partial), n, ep);
        break;
      case B200PHI_INT64:
        finish_partial_kernel<long long><<<g, kBlock, 0, stream>>>(static_cast<const long long*>(partial), n, ep);
        break;
      case B200PHI_INT32:
        finish_partial_kernel<int><<<g, kBlock, 0, stream>>>(static_cast<const int*>(partial), n, ep);
        break;
      default:
        break;  // unreachable: rejected before the first launch
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_status(e);
  }
  return B200PHI_OK;
}

// byte-wise all-gather on `stream` (used by the sharded arg-reduce)
int comm_all_gather_bytes(const void* send, void* recv, size_t bytes_per_rank, void* comm,
                          cudaStream_t stream) {
  NcclApi* a = nccl_api();
  if (!a->ok) return set_error(B200PHI_NCCL_ERROR, a->why);
  ncclResult_t r = a->AllGather(send, recv, bytes_per_rank, ncclChar, static_cast<ncclComm_t>(comm), stream);
  if (r != ncclSuccess) return nccl_status(a, r);
  return B200PHI_OK;
}

int comm_unique_id(unsigned char id[128]) {
  NcclApi* a = nccl_api();
  if (!a->ok) return set_error(B200PHI_NCCL_ERROR, a->why);
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId u;
  ncclResult_t r = a->GetUniqueId(&u);
  if (r != ncclSuccess) return nccl_status(a, r);
  memcpy(id, &u, 128);
  return B200PHI_OK;
}

int comm_init_rank(const unsigned char id[128], int nranks, int rank, void** comm) {
  NcclApi* a = nccl_api();
  if (!a->ok) return set_error(B200PHI_NCCL_ERROR, a->why);
  if (!comm || nranks < 1 || rank < 0 || rank >= nranks)
    return set_error(B200PHI_INVALID_ARGUMENT, "bad communicator arguments");
  ncclUniqueId u;
  memcpy(&u, id, 128);
  ncclComm_t c = nullptr;
  ncclResult_t r = a->CommInitRank(&c, nranks, u, rank);
  if (r != ncclSuccess) return nccl_status(a, r);
  *comm = c;
  return B200PHI_OK;
}

int comm_destroy(void* comm) {
  if (!comm) return B200PHI_OK;
  NcclApi* a = nccl_api();
  if (!a->ok) return set_error(B200PHI_NCCL_ERROR, a->why);
  ncclResult_t r = a->CommDestroy(static_cast<ncclComm_t>(comm));
  if (r != ncclSuccess) return nccl_status(a, r);
  return B200PHI_OK;
}

}  // namespace b200phi
