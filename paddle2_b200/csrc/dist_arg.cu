// argmax / argmin over the SHARDED leading axis: NCCL has no arg-reduce, so
// every rank reduces its rows to (value, local index) per column, turns the index global with its
// row offset, all-gathers the small [C] pairs, and selects with the reference's tie rule --
// better value, else lowest index (cub::ArgMax / Eigen ArgMaxTupleReducer).
// Everything is enqueued on the caller's stream; scratch comes from the caller.
#include <string>

#include "../../include/b200phi.h"
#include "prim.cuh"

namespace b200phi {

int set_error(int code, const std::string& msg);
int cuda_status(cudaError_t e);
int run_argreduce(const b200phi_argreduce_desc* d, cudaStream_t stream);
size_t argreduce_workspace_bytes(const b200phi_argreduce_desc* d);
int current_sm_count();
int comm_all_gather_bytes(const void* send, void* recv, size_t bytes_per_rank, void* comm,
                          cudaStream_t stream);

static size_t align256(size_t v) { return (v + 255) / 256 * 256; }

static size_t elem_size(int dt) {
  switch (dt) {
    case B200PHI_INT16:
    case B200PHI_FLOAT16:
    case B200PHI_BFLOAT16:
      return 2;
    case B200PHI_INT32:
    case B200PHI_FLOAT32:
      return 4;
    case B200PHI_INT64:
    case B200PHI_FLOAT64:
      return 8;
    default:
      return 0;
  }
}

struct ShardArgLayout {
  long long cols;
  size_t off_idx_local, off_val_local, off_idx_all, off_val_all, off_local_ws, local_ws_bytes, total;
};

static bool shard_arg_layout(const b200phi_argreduce_desc* d, int nranks, ShardArgLayout* L) {
  if (!d || d->rank < 1 || d->rank > B200PHI_MAX_RANK || nranks < 1) return false;
  const size_t es = elem_size(d->in_dtype);
  if (es == 0) return false;
  long long cols = 1;
  for (int i = 1; i < d->rank; ++i) cols *= d->dims[i];
  L->cols = cols;
  size_t off = 0;
  L->off_idx_local = off;
  off = align256(off + static_cast<size_t>(cols) * 8);
  L->off_val_local = off;
  off = align256(off + static_cast<size_t>(cols) * es);
  L->off_idx_all = off;
  off = align256(off + static_cast<size_t>(cols) * 8 * nranks);
  L->off_val_all = off;
  off = align256(off + static_cast<size_t>(cols) * es * nranks);
  // the local step's own scratch: the shard's rows are split across thread blocks
  b200phi_argreduce_desc local = *d;
  local.index_dtype = B200PHI_INT64;
  L->off_local_ws = off;
  L->local_ws_bytes = argreduce_workspace_bytes(&local);
  off = align256(off + L->local_ws_bytes);
  L->total = off;
  return true;
}

__global__ void __launch_bounds__(kBlock) add_offset_kernel(long long* idx, long long n, long long off) {
  for (long long i = static_cast<long long>(blockIdx.x) * kBlock + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * kBlock)
    idx[i] += off;
}

// pairs[r][c] for r < nranks -> the winner per column
template <typename T, typename V, bool kMax>
__global__ void __launch_bounds__(kBlock)
    arg_combine_kernel(const T* __restrict__ vals, const long long* __restrict__ idx, int nranks,
                       long long cols, void* out, int index_dtype, T* val_out) {
  for (long long c = static_cast<long long>(blockIdx.x) * kBlock + threadIdx.x; c < cols;
       c += static_cast<long long>(gridDim.x) * kBlock) {
    T bt = vals[c];
    V bv = Cvt<V, T>::to(bt);
    long long bi = idx[c];
    for (int r = 1; r < nranks; ++r) {
      const T t = vals[static_cast<long long>(r) * cols + c];
      const V v = Cvt<V, T>::to(t);
      const long long i = idx[static_cast<long long>(r) * cols + c];
      const bool better = kMax ? (v > bv) : (v < bv);
      if (better || (v == bv && i < bi)) {
        bv = v;
        bt = t;
        bi = i;
      }
    }
    if (index_dtype == B200PHI_INT32)
      reinterpret_cast<int*>(out)[c] = static_cast<int>(bi);
    else
      reinterpret_cast<long long*>(out)[c] = bi;
    if (val_out) val_out[c] = bt;
  }
}

template <typename T, typename V>
static void launch_combine(bool is_max, const void* vals, const long long* idx, int nranks, long long cols,
                           void* out, int index_dtype, void* val_out, cudaStream_t s) {
  long long blocks = (cols + kBlock - 1) / kBlock;
  if (blocks > current_sm_count() * 8LL) blocks = current_sm_count() * 8LL;
  const unsigned g = static_cast<unsigned>(blocks);
  if (is_max)
    arg_combine_kernel<T, V, true><<<g, kBlock, 0, s>>>(static_cast<const T*>(vals), idx, nranks, cols, out,
                                                        index_dtype, static_cast<T*>(val_out));
  else
    arg_combine_kernel<T, V, false><<<g, kBlock, 0, s>>>(static_cast<const T*>(vals), idx, nranks, cols, out,
                                                         index_dtype, static_cast<T*>(val_out));
}

}  // namespace b200phi

using namespace b200phi;

extern "C" size_t b200phi_sharded_argreduce_workspace(const b200phi_argreduce_desc* d, int nranks) {
  ShardArgLayout L;
  if (!shard_arg_layout(d, nranks, &L)) return 0;
  return L.total;
}

extern "C" int b200phi_sharded_argreduce(const b200phi_argreduce_desc* d, int64_t row_offset, int nranks,
                                         void* comm, void* workspace, size_t workspace_bytes,
                                         void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!d) return set_error(B200PHI_INVALID_ARGUMENT, "null descriptor");
  if (nranks <= 1 || !comm) {  // single GPU: no exchange (north star)
    if (row_offset != 0) return set_error(B200PHI_INVALID_ARGUMENT, "row_offset without a communicator");
    b200phi_argreduce_desc one = *d;
    if (!one.workspace) {  // the scratch handed to this entry point serves the local reduce
      one.workspace = workspace;
      one.workspace_bytes = workspace_bytes;
    }
    return run_argreduce(&one, stream);
  }
  if (d->flatten || d->rank < 1 || (d->axis != 0 && d->axis != -d->rank))
    return set_error(B200PHI_INVALID_ARGUMENT, "sharded argreduce reduces the sharded leading axis (axis 0)");
  ShardArgLayout L;
  if (!shard_arg_layout(d, nranks, &L))
    return set_error(B200PHI_UNSUPPORTED, "sharded argreduce: dtype not built");
  if (!workspace || workspace_bytes < L.total) {
    return set_error(B200PHI_WORKSPACE_TOO_SMALL,
                     "sharded argreduce: workspace too small (see b200phi_sharded_argreduce_workspace)");
  }
  if (reinterpret_cast<uintptr_t>(workspace) % 256 != 0)
    return set_error(B200PHI_INVALID_ARGUMENT, "workspace must be 256-byte aligned");
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  long long* idx_local = reinterpret_cast<long long*>(ws + L.off_idx_local);
  void* val_local = ws + L.off_val_local;
  long long* idx_all = reinterpret_cast<long long*>(ws + L.off_idx_all);
  void* val_all = ws + L.off_val_all;

  b200phi_argreduce_desc local = *d;
  local.index_dtype = B200PHI_INT64;
  local.out = idx_local;
  local.max_out = val_local;
  local.workspace = L.local_ws_bytes ? ws + L.off_local_ws : nullptr;
  local.workspace_bytes = L.local_ws_bytes;
  int rc = run_argreduce(&local, stream);
  if (rc != B200PHI_OK) return rc;
  if (row_offset != 0) {
    long long blocks = (L.cols + kBlock - 1) / kBlock;
    if (blocks > current_sm_count() * 8LL) blocks = current_sm_count() * 8LL;
    add_offset_kernel<<<static_cast<unsigned>(blocks), kBlock, 0, stream>>>(idx_local, L.cols, row_offset);
  }
  const size_t es = elem_size(d->in_dtype);
  rc = comm_all_gather_bytes(idx_local, idx_all, static_cast<size_t>(L.cols) * 8, comm, stream);
  if (rc != B200PHI_OK) return rc;
  rc = comm_all_gather_bytes(val_local, val_all, static_cast<size_t>(L.cols) * es, comm, stream);
  if (rc != B200PHI_OK) return rc;
  const int index_dtype = d->index_dtype == B200PHI_UNDEFINED ? B200PHI_INT64 : d->index_dtype;
  const bool is_max = d->op == B200PHI_ARGMAX;
  switch (d->in_dtype) {
    case B200PHI_FLOAT32:
      launch_combine<float, float>(is_max, val_all, idx_all, nranks, L.cols, d->out, index_dtype, d->max_out, stream);
      break;
    case B200PHI_FLOAT64:
      launch_combine<double, double>(is_max, val_all, idx_all, nranks, L.cols, d->out, index_dtype, d->max_out, stream);
      break;
    case B200PHI_BFLOAT16:
      launch_combine<__nv_bfloat16, float>(is_max, val_all, idx_all, nranks, L.cols, d->out, index_dtype, d->max_out, stream);
      break;
    case B200PHI_FLOAT16:
      launch_combine<__half, float>(is_max, val_all, idx_all, nranks, L.cols, d->out, index_dtype, d->max_out, stream);
      break;
    case B200PHI_INT16:
      launch_combine<int16_t, int>(is_max, val_all, idx_all, nranks, L.cols, d->out, index_dtype, d->max_out, stream);
      break;
    case B200PHI_INT32:
      launch_combine<int32_t, int>(is_max, val_all, idx_all, nranks, L.cols, d->out, index_dtype, d->max_out, stream);
      break;
    case B200PHI_INT64:
      launch_combine<long long, long long>(is_max, val_all, idx_all, nranks, L.cols, d->out, index_dtype, d->max_out, stream);
      break;
    default:
      return set_error(B200PHI_UNSUPPORTED, "sharded argreduce: dtype not built");
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}
