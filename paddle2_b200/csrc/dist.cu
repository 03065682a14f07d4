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

template <typename T>
__global__ void __launch_bounds__(kBlock) divide_inplace_kernel(T* p, long long n, double count) {
  for (long long i = static_cast<long long>(blockIdx.x) * kBlock + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * kBlock) {
    if constexpr (std::numeric_limits<T>::is_integer) {
      p[i] = static_cast<T>(p[i] / static_cast<long long>(count));
    } else {
      p[i] = p[i] / static_cast<T>(count);
    }
  }
}
__global__ void __launch_bounds__(kBlock)
    divide_inplace_bf16(__nv_bfloat16* p, long long n, float count) {
  for (long long i = static_cast<long long>(blockIdx.x) * kBlock + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * kBlock)
    p[i] = __float2bfloat16_rn(__bfloat162float(p[i]) / count);
}
__global__ void __launch_bounds__(kBlock) divide_inplace_f16(__half* p, long long n, float count) {
  for (long long i = static_cast<long long>(blockIdx.x) * kBlock + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * kBlock)
    p[i] = __float2half_rn(__half2float(p[i]) / count);
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

  ReducePlan plan;
  std::string err;
  int rc = plan_reduce(d, default_tunables(current_sm_count()), &plan, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);

  b200phi_reduce_desc local = *d;
  if (d->op == B200PHI_MEAN) local.op = B200PHI_SUM;  // divide by the GLOBAL count after the exchange
  rc = run_reduce(&local, stream);
  if (rc != B200PHI_OK) return rc;

  ncclDataType_t ty;
  switch (plan.out_dtype) {
    case B200PHI_FLOAT32: ty = ncclFloat32; break;
    case B200PHI_FLOAT64: ty = ncclFloat64; break;
    case B200PHI_FLOAT16: ty = ncclFloat16; break;
    case B200PHI_BFLOAT16: ty = ncclBfloat16; break;
    case B200PHI_INT32: ty = ncclInt32; break;
    case B200PHI_INT64: ty = ncclInt64; break;
    case B200PHI_INT8: ty = ncclInt8; break;
    case B200PHI_UINT8:
    case B200PHI_BOOL: ty = ncclUint8; break;
    default:
      return set_error(B200PHI_UNSUPPORTED, "sharded reduce: output dtype has no NCCL type");
  }
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
  ncclResult_t r = a->AllReduce(d->out, d->out, static_cast<size_t>(plan.out_numel), ty, op,
                                static_cast<ncclComm_t>(comm), stream);
  if (r != ncclSuccess) return nccl_status(a, r);

  if (d->op == B200PHI_MEAN) {
    if (global_count <= 0) return set_error(B200PHI_INVALID_ARGUMENT, "global_reduce_count must be > 0");
    const long long n = plan.out_numel;
    long long blocks = (n + kBlock - 1) / kBlock;
    if (blocks > 148 * 8) blocks = 148 * 8;
    const unsigned g = static_cast<unsigned>(blocks);
    switch (plan.out_dtype) {
      case B200PHI_FLOAT32:
        divide_inplace_kernel<float><<<g, kBlock, 0, stream>>>(static_cast<float*>(d->out), n, (double)global_count);
        break;
      case B200PHI_FLOAT64:
        divide_inplace_kernel<double><<<g, kBlock, 0, stream>>>(static_cast<double*>(d->out), n, (double)global_count);
        break;
      case B200PHI_INT32:
        divide_inplace_kernel<int><<<g, kBlock, 0, stream>>>(static_cast<int*>(d->out), n, (double)global_count);
        break;
      case B200PHI_INT64:
        divide_inplace_kernel<long long><<<g, kBlock, 0, stream>>>(static_cast<long long*>(d->out), n, (double)global_count);
        break;
      case B200PHI_BFLOAT16:
        divide_inplace_bf16<<<g, kBlock, 0, stream>>>(static_cast<__nv_bfloat16*>(d->out), n, (float)global_count);
        break;
      case B200PHI_FLOAT16:
        divide_inplace_f16<<<g, kBlock, 0, stream>>>(static_cast<__half*>(d->out), n, (float)global_count);
        break;
      default:
        return set_error(B200PHI_UNSUPPORTED, "sharded mean: dtype not built");
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
