// extern "C" surface of libb200phi.so (include/b200phi.h).  Nothing here touches
// phi or torch types; the host logic is reduce_plan.cc / bcast_plan.cc.
#include <cuda_runtime.h>
#include <stdio.h>

#include <mutex>
#include <string>

#include "../../include/b200phi.h"
#include "reduce_launch.cuh"

namespace b200phi {

static thread_local std::string g_last_error;

int set_error(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}

int cuda_status(cudaError_t e) {
  return set_error(B200PHI_CUDA_ERROR, std::string("CUDA: ") + cudaGetErrorString(e));
}

// Device properties are immutable: cache per device ordinal.
int current_sm_count() {
  static std::mutex mu;
  static int cache[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    cudaGetLastError();
    return 148;
  }
  if (dev < 0 || dev >= 64) return 148;
  std::lock_guard<std::mutex> lk(mu);
  if (cache[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
      cudaGetLastError();
      n = 148;
    }
    cache[dev] = n;
  }
  return cache[dev];
}

// Tunables of the current device (environment part cached, see reduce_plan.cc).
Tunables tunables() { return default_tunables(current_sm_count()); }

// Ring depth of the staged short-row kernels (rows_staged.cuh): B200PHI_STAGED_STAGES, read once.
int staged_stages() {
  static const int v = [] {
    const char* e = getenv("B200PHI_STAGED_STAGES");
    int n = e ? atoi(e) : 1;
    return n < 1 ? 1 : (n > 4 ? 4 : n);
  }();
  return v;
}
// Rows per thread and span of the same kernels: B200PHI_STAGED_RPT (1..8; 0: by row length).
int staged_rows_per_thread() {
  static const int v = [] {
    const char* e = getenv("B200PHI_STAGED_RPT");
    const int n = e ? atoi(e) : 0;
    return n < 0 ? 0 : (n > 8 ? 8 : n);
  }();
  return v;
}
// Lanes per row of the same kernels: B200PHI_STAGED_LANES (1, 2 or 4).  One lane measured best
// at every row length tried (101 bf16 columns: sum 6.2 / 6.0 / 4.1 TB/s, mean+var 5.2 / 4.1 / 2.7
// for 1 / 2 / 4 lanes).
int staged_lanes() {
  static const int v = [] {
    const char* e = getenv("B200PHI_STAGED_LANES");
    const int n = e ? atoi(e) : 1;
    return (n == 1 || n == 2 || n == 4) ? n : 1;
  }();
  return v;
}

static int run_reduce_impl(const b200phi_reduce_desc* d, int allow_narrow_out, cudaStream_t stream);
static int run_reduce_stage(const b200phi_reduce_desc* d, int second, cudaStream_t stream) {
  return run_reduce_impl(d, second, stream);
}
int run_reduce(const b200phi_reduce_desc* d, cudaStream_t stream) { return run_reduce_impl(d, 0, stream); }

static int run_reduce_impl(const b200phi_reduce_desc* d, int allow_narrow_out, cudaStream_t stream) {
  ReducePlan plan;
  std::string err;
  Tunables tun = tunables();
  tun.allow_narrow_out = allow_narrow_out;
  int rc = plan_reduce(d, tun, &plan, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  if (!d->x || !d->out) return set_error(B200PHI_INVALID_ARGUMENT, "null tensor pointer");
  if (plan.ws_total_bytes > 0) {
    if (!d->workspace || d->workspace_bytes < plan.ws_total_bytes) {
      char buf[160];
      snprintf(buf, sizeof(buf), "workspace too small: need %zu bytes, got %zu",
               plan.ws_total_bytes, d->workspace ? d->workspace_bytes : (size_t)0);
      return set_error(B200PHI_WORKSPACE_TOO_SMALL, buf);
    }
    if (reinterpret_cast<uintptr_t>(d->workspace) % 256 != 0)
      return set_error(B200PHI_INVALID_ARGUMENT, "workspace must be 256-byte aligned");
  }
  if (plan.kind == kCastFirst) {
    // the reference's own Cast pre-pass (gpu/reduce.h:48-69): x -> scratch in the output dtype
    // (a reduce over an appended extent-1 dim is an element-wise cast), then the reduce proper
    unsigned char* ws = static_cast<unsigned char*>(d->workspace);
    b200phi_reduce_desc c = b200phi_reduce_desc();
    c.op = B200PHI_SUM;
    c.in_dtype = d->in_dtype;
    c.out_dtype = plan.ts_tmp_dtype;
    c.rank = 2;
    c.dims[0] = plan.in_numel;
    c.dims[1] = 1;
    c.naxes = 1;
    c.axes[0] = 1;
    c.x = d->x;
    c.out = ws;
    rc = run_reduce_stage(&c, 1, stream);
    if (rc != B200PHI_OK) return rc;
    b200phi_reduce_desc r = *d;
    r.in_dtype = plan.ts_tmp_dtype;
    r.out_dtype = plan.ts_tmp_dtype;
    r.x = ws;
    r.workspace = ws + plan.ts_tmp_bytes;
    r.workspace_bytes = plan.ts_ws2_bytes;
    return run_reduce_stage(&r, 0, stream);
  }
  if (plan.kind == kTwoStage) {
    // one reduced dim peeled by a streaming row / column kernel into the scratch, then the rest
    // (itself possibly two stages): [A,K,B] -> row sums of [A*K, B], then the column reduce of [A, K]
    unsigned char* ws = static_cast<unsigned char*>(d->workspace);
    b200phi_reduce_desc d1, d2;
    two_stage_descs(plan, d, ws, &d1, &d2);
    d1.workspace = ws + plan.ts_tmp_bytes;
    d1.workspace_bytes = plan.ts_ws1_bytes;
    d2.workspace = ws + plan.ts_tmp_bytes + plan.ts_ws1_bytes;
    d2.workspace_bytes = plan.ts_ws2_bytes;
    rc = run_reduce_stage(&d1, 0, stream);
    if (rc != B200PHI_OK) return rc;
    return run_reduce_stage(&d2, 1, stream);
  }
  if (plan.red_numel == 1 && plan.in_dtype == plan.out_dtype && d->op <= B200PHI_AMIN &&
      d->op != B200PHI_ANY && d->op != B200PHI_ALL) {
    // nothing to combine and nothing to convert (reduce_function.h:1087-1092 runs an identity
    // transform kernel here): a device-to-device copy on the caller's stream
    if (d->out != d->x) {
      cudaError_t e = cudaMemcpyAsync(d->out, d->x, static_cast<size_t>(plan.in_numel) * dtype_size(plan.in_dtype),
                                      cudaMemcpyDeviceToDevice, stream);
      if (e != cudaSuccess) return cuda_status(e);
    }
    return B200PHI_OK;
  }
  ReduceArgs a{&plan, d->x, d->out, d->workspace, stream};
  const bool tiled = plan.kind == kTiledKeep || plan.kind == kTiledRed;
  switch (d->op) {
    case B200PHI_SUM:
    case B200PHI_MEAN:
      rc = tiled ? launch_sum_tiled(a) : launch_sum(a);
      break;
    case B200PHI_PROD:
      rc = tiled ? launch_prod_tiled(a) : launch_prod(a);
      break;
    case B200PHI_MAX:
    case B200PHI_MIN:
    case B200PHI_AMAX:
    case B200PHI_AMIN:
      rc = tiled ? launch_minmax_tiled(a) : launch_minmax(a);
      break;
    case B200PHI_ANY:
    case B200PHI_ALL:
      rc = tiled ? launch_logical_tiled(a) : launch_logical(a);
      break;
    default:
      rc = tiled ? launch_norm_tiled(a) : launch_norm(a);
      break;
  }
  if (rc == B200PHI_UNSUPPORTED) return set_error(rc, "dtype/op combination not built");
  return rc;
}

// The plan may depend on whether x is 16-byte aligned (vector or scalar column walk); the query
// must not: it answers for the worse of the two, whatever d->x holds (NULL included).
size_t reduce_workspace_bytes(const b200phi_reduce_desc* d) {
  if (!d) return 0;
  const Tunables t = tunables();
  size_t need = 0;
  const size_t es = dtype_size(d->in_dtype) ? dtype_size(d->in_dtype) : 1;
  for (int k = 0; k < 2; ++k) {
    b200phi_reduce_desc probe = *d;
    probe.x = reinterpret_cast<const void*>(static_cast<uintptr_t>(k == 0 ? 256 : 256 + es));
    ReducePlan plan;
    std::string err;
    if (plan_reduce(&probe, t, &plan, &err) != B200PHI_OK) return 0;
    if (plan.ws_total_bytes > need) need = plan.ws_total_bytes;
    if (es >= 16) break;
  }
  return need;
}

int run_argreduce(const b200phi_argreduce_desc* d, cudaStream_t stream);
size_t argreduce_workspace_bytes(const b200phi_argreduce_desc* d);
int run_rowstats(const b200phi_rowstats_desc* d, cudaStream_t stream);
int run_elementwise(const b200phi_elementwise_desc* d, cudaStream_t stream);
int run_sharded_reduce(const b200phi_reduce_desc* d, long long global_count, void* comm,
                       cudaStream_t stream);
size_t sharded_reduce_workspace_bytes(const b200phi_reduce_desc* d);
int comm_unique_id(unsigned char id[128]);
int comm_init_rank(const unsigned char id[128], int nranks, int rank, void** comm);
int comm_destroy(void* comm);

}  // namespace b200phi

using namespace b200phi;

extern "C" {

const char* b200phi_last_error(void) { return g_last_error.c_str(); }

const char* b200phi_version(void) { return "b200phi 0.1 sm_100a"; }

size_t b200phi_reduce_workspace(const b200phi_reduce_desc* d) {
  return reduce_workspace_bytes(d);
}

int b200phi_reduce_plan_kind(const b200phi_reduce_desc* d) {
  ReducePlan plan;
  std::string err;
  if (plan_reduce(d, tunables(), &plan, &err) != B200PHI_OK) return -1;
  return plan.kind;
}

int b200phi_reduce(const b200phi_reduce_desc* d, void* stream) {
  return run_reduce(d, static_cast<cudaStream_t>(stream));
}

size_t b200phi_argreduce_workspace(const b200phi_argreduce_desc* d) { return argreduce_workspace_bytes(d); }

int b200phi_argreduce(const b200phi_argreduce_desc* d, void* stream) {
  return run_argreduce(d, static_cast<cudaStream_t>(stream));
}

int b200phi_rowstats(const b200phi_rowstats_desc* d, void* stream) {
  return run_rowstats(d, static_cast<cudaStream_t>(stream));
}

int b200phi_elementwise(const b200phi_elementwise_desc* d, void* stream) {
  return run_elementwise(d, static_cast<cudaStream_t>(stream));
}

int b200phi_comm_unique_id(unsigned char id[128]) { return comm_unique_id(id); }

int b200phi_comm_init_rank(const unsigned char id[128], int nranks, int rank, void** comm) {
  return comm_init_rank(id, nranks, rank, comm);
}

int b200phi_comm_destroy(void* comm) { return comm_destroy(comm); }

size_t b200phi_sharded_reduce_workspace(const b200phi_reduce_desc* d) {
  return sharded_reduce_workspace_bytes(d);
}

int b200phi_sharded_reduce(const b200phi_reduce_desc* d, int64_t global_reduce_count, void* comm,
                           void* stream) {
  return run_sharded_reduce(d, global_reduce_count, comm, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
