// Drop-in adapter, fourth translation unit: the Partial -> Replicate reshard of the auto-parallel
// runtime on the C ABI.
//
// When a DistTensor reduction leaves its output Partial(sum|avg|max|min|prod)
// (paddle/phi/infermeta/spmd_rules/reduction.cc:112-149), the reshard registry walks its list for
// the first function whose IsSuitable() holds (reshard_function_registry.cc:41-58) and calls Eval.
// The reference's PToRReshardFunction::Eval (reshard/p_to_r_reshard_function.cc:45-93) runs the
// AllReduce kernel through NCCLCommContext and, for Partial(avg), a Full and a Divide kernel
// behind it.  B200PToRReshardFunction answers the same IsSuitable() question for GPU tensors and
// makes ONE b200phi_sharded_reduce call on the compute stream: the partial viewed as a one-row
// shard [1, numel] (its "local reduce" is the copy into `out`), the all-reduce of it, and the
// mean's division by the number of ranks folded in.  A maintainer registers it in front of
// PToRReshardFunction:
//     REGISTER_RESHARD_FUNC(B200PToRReshardFunction);   // reshard_function_registry.cc, before :68
// (the list is walked in registration order).  Where the local reduce itself should fuse with
// the exchange -- paddle.sum of a row-sharded DistTensor -- b200::ShardedReduce
// (phi_b200_dist_hook.cu) is the call the SPMD rule's kernel makes instead of SumKernel.
//
// Compile-checked against the reference headers with -DPADDLE_WITH_NCCL
// (adapter/check_compile.sh); it cannot be linked without libphi.
#include <memory>
#include <vector>

#include "b200phi.h"
#include "paddle/phi/backends/gpu/gpu_context.h"
#include "paddle/phi/common/reduce_type.h"
#include "paddle/phi/core/distributed/auto_parallel/dist_tensor.h"
#include "paddle/phi/core/distributed/auto_parallel/reshard/reshard_function.h"
#include "paddle/phi/core/distributed/auto_parallel/reshard/reshard_function_registry.h"
#include "paddle/phi/core/distributed/auto_parallel/reshard/reshard_utils.h"
#include "paddle/phi/core/distributed/nccl_comm_context.h"

namespace phi {
namespace distributed {

class B200PToRReshardFunction final : public ReshardFunction {
 public:
  bool IsSuitable(const DistTensor& in, const TensorDistAttr& out_dist_attr) override {
    // the reference's conditions (p_to_r_reshard_function.cc:30-43) ...
    RESHARD_SHORTCUT_IF_FALSE(in.dist_attr().is_partial());
    RESHARD_SHORTCUT_IF_FALSE(out_dist_attr.is_replicated());
    const auto& in_mesh = in.dist_attr().process_mesh();
    const auto& out_mesh = out_dist_attr.process_mesh();
    RESHARD_SHORTCUT_IF_FALSE(in_mesh.ndim() == 1);
    RESHARD_SHORTCUT_IF_FALSE(out_mesh.ndim() == 1);
    RESHARD_SHORTCUT_IF_FALSE(in_mesh == out_mesh);
    // ... and what this library serves: a GPU tensor, at most the 8 ranks of one box, a reduce
    // type with an NCCL red-op (gpu/all_reduce_kernel.cu:46-65)
    RESHARD_SHORTCUT_IF_FALSE(in.value().place().GetType() == phi::AllocationType::GPU);
    RESHARD_SHORTCUT_IF_FALSE(in.value().numel() > 0);
    const auto type = in.dist_attr().partial_status().at(0);
    RESHARD_SHORTCUT_IF_FALSE(type == ReduceType::kRedSum || type == ReduceType::kRedAvg ||
                              type == ReduceType::kRedMax || type == ReduceType::kRedMin ||
                              type == ReduceType::kRedProd);
    return true;
  }

  void Eval(DeviceContext* dev_ctx, const DistTensor& in, const TensorDistAttr& out_dist_attr,
            DistTensor* out) override {
    const auto& process_ids = in.dist_attr().process_mesh().process_ids();
    int op = B200PHI_SUM;
    switch (in.dist_attr().partial_status().at(0)) {
      case ReduceType::kRedAvg: op = B200PHI_MEAN; break;
      case ReduceType::kRedMax: op = B200PHI_MAX; break;
      case ReduceType::kRedMin: op = B200PHI_MIN; break;
      case ReduceType::kRedProd: op = B200PHI_PROD; break;
      default: break;
    }
    auto* gpu_ctx = static_cast<phi::GPUContext*>(dev_ctx);
    auto* comm_ctx = static_cast<NCCLCommContext*>(CreateOrGetCommContext(*dev_ctx, process_ids));
    const DenseTensor& partial = in.value();
    DenseTensor* result = GetMutableTensor(out);
    result->Resize(partial.dims());
    gpu_ctx->Alloc(result, partial.dtype());

    // the partial as a one-row shard [1, numel]: reducing axis 0 locally is the copy into
    // `result`; the exchange and, for Partial(avg), the division by the rank count follow on the
    // same stream inside the call
    b200phi_reduce_desc d = {};
    d.op = op;
    d.in_dtype = d.out_dtype = static_cast<int>(partial.dtype());
    d.rank = 2;
    d.dims[0] = 1;
    d.dims[1] = partial.numel();
    d.naxes = 1;
    d.axes[0] = 0;
    d.x = partial.data();
    d.out = result->data();
    DenseTensor scratch;
    const size_t ws = b200phi_sharded_reduce_workspace(&d);
    if (ws > 0) {
      scratch.Resize({static_cast<int64_t>(ws)});
      d.workspace = gpu_ctx->template Alloc<uint8_t>(&scratch);
      d.workspace_bytes = ws;
    }
    const int rc = b200phi_sharded_reduce(&d, static_cast<int64_t>(process_ids.size()),
                                          static_cast<void*>(comm_ctx->GetNcclComm()), gpu_ctx->stream());
    if (rc != B200PHI_OK) {
      PADDLE_THROW(common::errors::InvalidArgument("%s", b200phi_last_error()));
    }
    SetDistProps(out, in.dims(), out_dist_attr);
  }

  std::string Name() override { return "B200PToRReshard"; }
};

// in a Paddle build this line sits in reshard_function_registry.cc in front of
// REGISTER_RESHARD_FUNC(PToRReshardFunction); here it proves the class is complete and registrable
REGISTER_RESHARD_FUNC(B200PToRReshardFunction);

}  // namespace distributed
}  // namespace phi
