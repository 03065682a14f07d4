// b200phi_reduce_noalloc: b200phi_reduce for callers that have no way to give it scratch
// memory.
//
// The ordinary entry point, b200phi_reduce, asks its caller for a workspace whenever the
// reduced extent is cut across thread blocks (partials + tickets for the last-block finalise
// of a column reduce or of a few very long rows) or the reduce runs as two stages through an
// intermediate ([A, K, B] with A and B reduced).  Some hosts cannot provide one: a C or Go service that links this library
// without any device allocator of its own, an inference server whose memory plan is frozen
// before the first request so that nothing on the request path may call cudaMalloc, a
// callback that runs where taking the allocator's lock is forbidden.  For them the scratch
// lives inside the library, in the fixed array below.
//
// What a caller has to know:
//   * Capacity is B200PHI_NOALLOC_SCRATCH_BYTES (4 MiB).  A reduce whose plan needs more is
//     refused with B200PHI_WORKSPACE_TOO_SMALL and a message naming both numbers; nothing
//     is launched.  (The fp32 [524288, 4096] axis-0 sum of config C5 needs 2.4 MB.)
//   * Every device that loads the library carries these 4 MiB until the process exits.
//   * The first call on a device looks up the array's device address
//     (cudaGetSymbolAddress) and keeps it.  Make that first call before starting any
//     stream capture on that device.
//   * All calls on one device share the one array and it keeps whatever the previous call
//     left in it; the tickets are cleared at the start of every call, the partials are
//     fully overwritten before they are read.  Two calls whose kernels overlap on the GPU
//     (different streams, or different host threads on different streams) race on it:
//     keep such calls on one stream or order them with events.
//   * No cudaMalloc, cudaFree, or other allocating routine is called, on any path.
#include <cuda_runtime.h>
#include <stdio.h>

#include <mutex>
#include <string>

#include "../../include/b200phi.h"
#include "reduce_plan.h"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int current_sm_count();
Tunables tunables();
int run_reduce(const b200phi_reduce_desc* d, cudaStream_t stream);

namespace {

// In host code this name is only a handle for the runtime's symbol table, not an address a
// kernel could use; the device address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200PHI_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_scratch_ptr[kMaxDevices] = {nullptr};

// Device address of the scratch array on the CURRENT device; looked up once per device.
int scratch_for_current_device(void** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_status(e);
  if (dev < 0 || dev >= kMaxDevices)
    return set_error(B200PHI_UNSUPPORTED, "b200phi_reduce_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return cuda_status(e);
    if (p == nullptr)
      return set_error(B200PHI_CUDA_ERROR, "b200phi_reduce_noalloc: scratch symbol has no device address");
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return B200PHI_OK;
}

}  // namespace
}  // namespace b200phi

using namespace b200phi;

extern "C" int b200phi_reduce_noalloc(const b200phi_reduce_desc* d, void* stream) {
  ReducePlan plan;
  std::string err;
  int rc = plan_reduce(d, tunables(), &plan, &err);
  if (rc != B200PHI_OK) return set_error(rc, err);
  // every shape is served: the ones that need no scratch simply do not touch the array
  if (plan.ws_total_bytes > static_cast<size_t>(B200PHI_NOALLOC_SCRATCH_BYTES)) {
    char buf[200];
    snprintf(buf, sizeof(buf),
             "b200phi_reduce_noalloc: this reduce needs %zu bytes of scratch, the built-in "
             "scratch holds %zu; use b200phi_reduce with a caller workspace",
             plan.ws_total_bytes, static_cast<size_t>(B200PHI_NOALLOC_SCRATCH_BYTES));
    return set_error(B200PHI_WORKSPACE_TOO_SMALL, buf);
  }
  void* scratch = nullptr;
  rc = scratch_for_current_device(&scratch);
  if (rc != B200PHI_OK) return rc;
  b200phi_reduce_desc local = *d;
  local.workspace = scratch;
  local.workspace_bytes = static_cast<size_t>(B200PHI_NOALLOC_SCRATCH_BYTES);
  // run_reduce clears the tickets (cudaMemsetAsync on `stream`) and launches the same
  // kernels that b200phi_reduce launches; it allocates nothing
  return run_reduce(&local, static_cast<cudaStream_t>(stream));
}
