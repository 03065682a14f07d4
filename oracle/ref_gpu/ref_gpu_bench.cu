// TEST / MEASUREMENT INFRASTRUCTURE -- not product code.
// Times the REFERENCE's own GPU kernels, instantiated from its headers where they lie under
// /root/reference and compiled for sm_100a ("the kernel to beat"):
//   ReduceAnyKernel        paddle/phi/kernels/funcs/reduce_function.h:573-699   (C1, last-dim reduce)
//   ReduceHigherDimKernel  paddle/phi/kernels/funcs/reduce_function.h:706-788   (C5-like, two passes)
// launched with the geometry ReduceConfig would choose (SetBlockDimForReduceAny :434-498,
// SetBlockDimForHigher :504-531), re-derived here by those formulas because ReduceConfig::Run
// needs a live GPUContext (libphi) that cannot be linked in this image.
// Output: oracle/_ref/ref_gpu_bench (git-ignored, travels to the GPU box).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "paddle/phi/kernels/funcs/reduce_function.h"
#include "paddle/phi/kernels/funcs/broadcast_function.h"
#include "paddle/phi/kernels/funcs/elementwise_functor.h"
#include "paddle/phi/kernels/funcs/functors.h"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

namespace kps = phi::kps;
using phi::funcs::OneDimIndexCal;

template <typename F> float time_it(F fn, int iters) {
  for (int i = 0; i < 3; ++i) fn();
  CK(cudaDeviceSynchronize());
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  for (int i = 0; i < iters; ++i) fn();
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  CK(cudaGetLastError());
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms / iters;
}

__global__ void fill(float* x, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    x[i] = (float)((i * 2654435761u) & 1023) / 1024.0f;
}

int main() {
  using MPType = float;
  using ReduceOp = kps::AddFunctor<MPType>;
  using TransformOp = kps::IdentityFunctor<float, MPType>;
  // ---------------- C1: [32768, 4096] fp32, reduce the last dim --------------------------------
  {
    const int left_num = 32768, reduce_num = 4096;
    float *x, *y;
    CK(cudaMalloc(&x, (size_t)left_num * reduce_num * 4)); CK(cudaMalloc(&y, left_num * 4));
    fill<<<1184, 256>>>(x, (long long)left_num * reduce_num);
    // SetBlockDimForReduceAny: block.x = GetBlockDim(reduce_num) capped at kReduceMaxThread = 128,
    // block.y = 1; grid.x = left_num; enough blocks -> grid.y = 1 (one pass)
    dim3 block(128, 1, 1), grid(left_num, 1, 1);
    kps::DimConfig dim(grid.x, grid.y, grid.z, block.x, block.y, 0);
    dim.SetRem(reduce_num % block.x, 0, 0);
    OneDimIndexCal reduce_cal(1), left_cal(reduce_num);
    ReduceOp reducer;
    float ms = time_it([&] {
      phi::funcs::ReduceAnyKernel<float, float, MPType, ReduceOp, TransformOp, OneDimIndexCal>
          <<<grid, block>>>(x, y, reducer, TransformOp(), reducer.initial(), reduce_num, left_num, true,
                            reduce_cal, left_cal, dim, false, nullptr, false);
    }, 20);
    float h[4];
    CK(cudaMemcpy(h, y, 16, cudaMemcpyDeviceToHost));
    printf("REF C1 ReduceAnyKernel   sum fp32 [32768,4096] axis -1   %8.2f us  %7.1f GB/s   (y[0]=%.3f)\n",
           ms * 1e3, 537001984.0 / ms / 1e6, h[0]);
    cudaFree(x); cudaFree(y);
  }
  // ---------------- C5-like: [R, 4096] fp32, reduce axis 0 (int-indexed kernels: R*C <= 2^30) ------
  {
    const int reduce_num = 262144, left_num = 4096;
    float *x, *y, *tmp;
    CK(cudaMalloc(&x, (size_t)left_num * reduce_num * 4)); CK(cudaMalloc(&y, left_num * 4));
    fill<<<1184, 256>>>(x, (long long)left_num * reduce_num);
    int sms = 0, tpm = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&tpm, cudaDevAttrMaxThreadsPerMultiProcessor, 0);
    const int max_threads = sms * tpm;
    const int num_block = max_threads / left_num;
    dim3 block(128, 1, 1), grid((left_num + 127) / 128, 1, 1);
    int blocking_size = reduce_num;
    bool again = false;
    if (num_block > 1 && reduce_num >= 512) {
      blocking_size = phi::funcs::details::GetLastPow2(reduce_num / num_block);
      if (blocking_size <= 1) blocking_size = phi::funcs::details::GetLastPow2((int)sqrt((double)reduce_num));
      else if (blocking_size * 2 < reduce_num) blocking_size *= 2;
      again = true;
      grid.y = (reduce_num + blocking_size - 1) / blocking_size;
    }
    CK(cudaMalloc(&tmp, (size_t)grid.y * left_num * 4));
    ReduceOp reducer;
    kps::DimConfig dim(grid.x, grid.y, grid.z, block.x, blocking_size, 0);
    dim.SetRem(left_num % block.x, reduce_num % blocking_size, 0);
    dim3 grid2(grid.x, 1, 1);
    kps::DimConfig dim2(grid2.x, grid2.y, grid2.z, block.x, grid.y, 0);
    dim2.SetRem(left_num % block.x, 0, 0);
    float ms = time_it([&] {
      phi::funcs::ReduceHigherDimKernel<float, float, MPType, ReduceOp, TransformOp>
          <<<grid, block>>>(x, y, reducer, TransformOp(), reducer.initial(), reduce_num, left_num, blocking_size,
                            dim, reduce_num, false, tmp, again);
      if (again)
        phi::funcs::ReduceHigherDimKernel<MPType, float, MPType, ReduceOp, kps::IdentityFunctor<MPType, MPType>>
            <<<grid2, block>>>(tmp, y, reducer, kps::IdentityFunctor<MPType, MPType>(), reducer.initial(),
                               (int)grid.y, left_num, (int)grid.y, dim2, reduce_num, false, tmp, false);
    }, 10);
    float h[4];
    CK(cudaMemcpy(h, y, 16, cudaMemcpyDeviceToHost));
    const double bytes = (double)left_num * reduce_num * 4 + left_num * 4;
    printf("REF C5' ReduceHigherDimKernel x2  sum fp32 [262144,4096] axis 0 (grid %d x %d, blocking %d)  %8.2f us  %7.1f GB/s  (y[0]=%.1f)\n",
           grid.x, grid.y, blocking_size, ms * 1e3, bytes / ms / 1e6, h[0]);
    cudaFree(x); cudaFree(y); cudaFree(tmp);
  }
  // ---------------- C2: relu(x + y) fp32 [256,64,112,112] + [64,1,1] ------------------------------
  // The reference's current route is two launches: add through
  // VectorizedBroadcastKernel<AddFunctor,float,2,1,4,kMixed> (broadcast_function.h:345-435,
  // launched as LaunchBroadcastKernel :437-512 does, geometry from GetGpuLaunchConfig1D,
  // gpu_launch_config.h:107-161), then relu through VectorizedElementwiseKernel
  // (elementwise_base.h:690-722).  (Its 2018 fused kernel, grid = 64 CTAs, needs a cblas header
  // that is not in the tree and cannot be instantiated here.)
  {
    const int64_t numel = (int64_t)256 * 64 * 112 * 112;
    float *x, *y, *t, *out;
    CK(cudaMalloc(&x, numel * 4)); CK(cudaMalloc(&t, numel * 4)); CK(cudaMalloc(&out, numel * 4)); CK(cudaMalloc(&y, 64 * 4));
    fill<<<1184, 256>>>(x, numel);
    fill<<<1, 64>>>(y, 64);
    constexpr int VecSize = 4;
    const int threads = 512;
    const int blocks = (int)(((numel + VecSize - 1) / VecSize + threads - 1) / threads);
    // BroadcastDimsSimplifier result for this shape (dims reversed, fastest first)
    std::vector<int64_t> out_dims = {12544, 64, 256}, x_dims = {12544, 64, 256}, y_dims = {1, 64, 1};
    phi::Array<const char* __restrict__, 2> ins;
    ins[0] = reinterpret_cast<const char*>(x);
    ins[1] = reinterpret_cast<const char*>(y);
    phi::Array<float*, 1> outs;
    outs[0] = t;
    phi::Array<bool, 2> use_bc;
    use_bc[0] = false;
    use_bc[1] = true;
    phi::Array<kps::details::BroadcastConfig, 2> configs;
    configs[0] = kps::details::BroadcastConfig(out_dims, x_dims, 3);
    configs[1] = kps::details::BroadcastConfig(out_dims, y_dims, 3);
    const int main_offset = (int)((numel / (VecSize * threads)) * VecSize * threads);
    const int tail_tid = (int)(numel % (VecSize * threads));
    phi::Array<const char* __restrict__, 1> ins2;
    ins2[0] = reinterpret_cast<const char*>(t);
    phi::Array<float*, 1> outs2;
    outs2[0] = out;
    float ms_add = 0, ms_both = 0;
    auto add = [&] {
      phi::funcs::VectorizedBroadcastKernel<phi::funcs::AddFunctor<float>, float, 2, 1, VecSize, phi::funcs::kMixed>
          <<<blocks, threads>>>(ins, outs, use_bc, (uint32_t)numel, configs, main_offset, tail_tid, VecSize,
                                phi::funcs::AddFunctor<float>());
    };
    auto relu = [&] {
      phi::funcs::VectorizedElementwiseKernel<float, phi::funcs::ReluFunctor<float>, 1, 1, VecSize>
          <<<blocks, threads>>>(ins2, outs2, numel, (int64_t)main_offset, VecSize, phi::funcs::ReluFunctor<float>());
    };
    ms_add = time_it(add, 10);
    ms_both = time_it([&] { add(); relu(); }, 10);
    printf("REF C2 VectorizedBroadcastKernel add alone (grid %d x %d)          %8.2f us  %7.1f GB/s\n", blocks, threads,
           ms_add * 1e3, 1644167424.0 / ms_add / 1e6);
    printf("REF C2 add + relu, two launches as the reference runs them        %8.2f us  %7.1f GB/s (of the fused op's algorithmic bytes)\n",
           ms_both * 1e3, 1644167424.0 / ms_both / 1e6);
    cudaFree(x); cudaFree(y); cudaFree(t); cudaFree(out);
  }
  printf("done\n");
  return 0;
}
