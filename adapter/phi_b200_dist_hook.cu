// Drop-in adapter, third translation unit: the auto-parallel hook.
//
// Today a reduction over the sharded axis of a DistTensor runs as two steps chosen by the SPMD
// rule (paddle/phi/infermeta/spmd_rules/reduction.cc:112-149): phi::SumKernel on the local shard,
// which leaves the output Partial(sum), then PToRReshardFunction::Eval
// (paddle/phi/core/distributed/auto_parallel/reshard/p_to_r_reshard_function.cc:45-93), which
// all-reduces it through NCCLCommContext (and, for Partial(avg), follows with Full + Divide).
// With this hook both steps are one C-ABI call on the compute stream: local reduce, ncclAllReduce
// of the small partial, and the mean's division by the GLOBAL count folded in.
//
// Compile-checked against the reference headers with -DPADDLE_WITH_NCCL
// (adapter/check_compile.sh); it cannot be linked without libphi.
#include <vector>

#include "b200phi.h"
#include "paddle/phi/backends/gpu/gpu_context.h"
#include "paddle/phi/core/dense_tensor.h"
#include "paddle/phi/core/distributed/nccl_comm_context.h"

namespace phi {
namespace b200 {

// x: this rank's shard (rows [rank*R, (rank+1)*R) of the global tensor); dims: the reduced axes
// (they include the sharded axis 0); out: replicated result.  `is_mean`: Partial(avg).
// global_reduce_count = product of the GLOBAL extents of the reduced axes.
inline void ShardedReduce(const phi::GPUContext& dev_ctx, distributed::NCCLCommContext* comm_ctx,
                          int op, const DenseTensor& x, const std::vector<int64_t>& dims,
                          int64_t global_reduce_count, DenseTensor* out) {
  dev_ctx.Alloc(out, out->dtype());
  b200phi_reduce_desc d = {};
  d.op = op;
  d.in_dtype = static_cast<int>(x.dtype());
  d.out_dtype = static_cast<int>(out->dtype());
  d.rank = x.dims().size();
  for (int i = 0; i < d.rank; ++i) d.dims[i] = x.dims()[i];
  d.naxes = static_cast<int>(dims.size());
  for (int i = 0; i < d.naxes; ++i) d.axes[i] = static_cast<int>(dims[i]);
  d.x = x.data();
  d.out = out->data();
  DenseTensor scratch;
  const size_t ws = b200phi_sharded_reduce_workspace(&d);  // + the fp32 partial for 16-bit sums
  if (ws > 0) {
    scratch.Resize({static_cast<int64_t>(ws)});
    d.workspace = dev_ctx.template Alloc<uint8_t>(&scratch);
    d.workspace_bytes = ws;
  }
  // the communicator phi already holds for this process group; the allreduce is enqueued on
  // the compute stream right behind the local reduce, so no event hand-off is needed
  void* comm = comm_ctx != nullptr ? static_cast<void*>(comm_ctx->GetNcclComm()) : nullptr;
  const int rc = b200phi_sharded_reduce(&d, global_reduce_count, comm, dev_ctx.stream());
  if (rc != B200PHI_OK) {
    PADDLE_THROW(common::errors::InvalidArgument("%s", b200phi_last_error()));
  }
}

}  // namespace b200
}  // namespace phi
