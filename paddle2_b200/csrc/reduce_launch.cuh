// Typed launchers: ReducePlan -> kernel<<<>>> on the caller's stream.
#pragma once

#include "reduce_kernels.cuh"
#include "reduce_plan.h"

namespace b200phi {

struct ReduceArgs {
  const ReducePlan* plan;
  const void* x;
  void* out;
  void* ws;
  cudaStream_t stream;
};

int cuda_status(cudaError_t e);  // api.cu: records the message, returns B200PHI_CUDA_ERROR

template <typename T, typename Acc, typename Op>
int launch_reduce_typed(const ReduceArgs& a) {
  const ReducePlan& p = *a.plan;
  const T* x = static_cast<const T*>(a.x);
  const Epilogue ep = make_epilogue(a.out, p.out_dtype, p.op == B200PHI_MEAN, p.red_numel);
  Acc* partial = static_cast<Acc*>(a.ws);
  unsigned* tickets =
      reinterpret_cast<unsigned*>(static_cast<unsigned char*>(a.ws) + p.ws_partial_bytes);
  if (p.ws_ticket_bytes) {
    // caller scratch arrives uninitialised; tickets must start at zero
    cudaError_t e = cudaMemsetAsync(tickets, 0, p.ws_ticket_bytes, a.stream);
    if (e != cudaSuccess) return cuda_status(e);
  }
  switch (p.kind) {
    case kRowWarp:
    case kRowBlock:
#define B200PHI_ROW(GG)                                                                        \
  if (p.aligned)                                                                               \
    row_reduce_kernel<T, Acc, Op, GG, 8, true><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, p.cols, ep); \
  else                                                                                         \
    row_reduce_kernel<T, Acc, Op, GG, 8, false><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, p.cols, ep);
      switch (p.G) {
        case 32:
          B200PHI_ROW(32) break;
        case 64:
          B200PHI_ROW(64) break;
        case 128:
          B200PHI_ROW(128) break;
        default:
          B200PHI_ROW(256) break;
      }
#undef B200PHI_ROW
      break;
    case kRowBatch: {
      const long long nvec = p.cols / VecOf<T>::kElems;
#define B200PHI_ROWB(GG, NN, RR) \
  row_batch_kernel<T, Acc, Op, GG, NN, RR><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, nvec, ep)
      if (p.G == 8)
        B200PHI_ROWB(8, 4, 2);
      else if (p.G == 16 && p.nv == 4)
        B200PHI_ROWB(16, 4, 2);
      else if (p.G == 16)
        B200PHI_ROWB(16, 8, 1);
      else if (p.G == 32)
        B200PHI_ROWB(32, 8, 1);
      else if (p.G == 64)
        B200PHI_ROWB(64, 8, 1);
      else
        B200PHI_ROWB(128, 8, 1);
#undef B200PHI_ROWB
      break;
    }
    case kRowTiny: {
      const int nvec = static_cast<int>(p.cols / VecOf<T>::kElems);
#define B200PHI_ROWT(GG) \
  row_tiny_kernel<T, Acc, Op, GG, 8><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, nvec, ep)
      if (p.G == 1)
        B200PHI_ROWT(1);
      else if (p.G == 2)
        B200PHI_ROWT(2);
      else if (p.G == 4)
        B200PHI_ROWT(4);
      else if (p.G == 8)
        B200PHI_ROWT(8);
      else
        B200PHI_ROWT(16);
#undef B200PHI_ROWT
      break;
    }
    case kRowScalar: {
      const int c = static_cast<int>(p.cols);
      if (p.G == 8 && p.nv == 8)
        row_scalar_kernel<T, Acc, Op, 8, 8><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
      else if (p.G == 8)
        row_scalar_kernel<T, Acc, Op, 8, 16><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
      else if (p.G == 16)
        row_scalar_kernel<T, Acc, Op, 16, 16><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
      else if (p.nv == 16)
        row_scalar_kernel<T, Acc, Op, 32, 16><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
      else
        row_scalar_kernel<T, Acc, Op, 32, 32><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
      break;
    }
    case kRowSmall: {
      const int c = static_cast<int>(p.cols);
      switch (p.G) {
        case 1:
          row_small_kernel<T, Acc, Op, 1><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
          break;
        case 2:
          row_small_kernel<T, Acc, Op, 2><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
          break;
        case 4:
          row_small_kernel<T, Acc, Op, 4><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
          break;
        case 8:
          row_small_kernel<T, Acc, Op, 8><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
          break;
        default:
          row_small_kernel<T, Acc, Op, 16><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
          break;
      }
      break;
    }
    case kRowSplit:
      row_split_kernel<T, Acc, Op, 8><<<dim3(p.grid_x, p.grid_y), kBlock, 0, a.stream>>>(
          x, p.rows, p.cols, p.seg, partial, tickets, ep);
      break;
    case kCol: {
      const size_t smem = static_cast<size_t>(kBlock) * p.vec * sizeof(Acc);
      if (p.vec > 1) {
        col_reduce_kernel<T, Acc, Op, VecOf<T>::kElems, 8>
            <<<dim3(p.grid_x, p.grid_y), kBlock, smem, a.stream>>>(
                x, p.R, p.inner, p.TX, p.tiles, p.rows_per_split, partial, tickets, ep);
      } else {
        col_reduce_kernel<T, Acc, Op, 1, 8><<<dim3(p.grid_x, p.grid_y), kBlock, smem, a.stream>>>(
            x, p.R, p.inner, p.TX, p.tiles, p.rows_per_split, partial, tickets, ep);
      }
      break;
    }
    default: {
      GenericDims gd;
      gd.nkeep = p.nkeep;
      gd.nred = p.nred;
      for (int i = 0; i < kMaxRank; ++i) {
        gd.keep_extent[i] = p.keep_extent[i];
        gd.keep_stride[i] = p.keep_stride[i];
        gd.red_extent[i] = p.red_extent[i];
        gd.red_stride[i] = p.red_stride[i];
      }
      gd.out_numel = p.out_numel;
      gd.red_numel = p.red_numel;
      generic_reduce_kernel<T, Acc, Op><<<p.grid_x, kBlock, 0, a.stream>>>(x, gd, p.TX, ep);
      break;
    }
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

// in_dtype -> storage type
template <typename Acc, template <typename> class OpT>
int dispatch_reduce_in(const ReduceArgs& a) {
  using Op = OpT<Acc>;
  switch (a.plan->in_dtype) {
    case B200PHI_BOOL:
      return launch_reduce_typed<Bool8, Acc, Op>(a);
    case B200PHI_UINT8:
      return launch_reduce_typed<uint8_t, Acc, Op>(a);
    case B200PHI_INT8:
      return launch_reduce_typed<int8_t, Acc, Op>(a);
    case B200PHI_INT16:
      return launch_reduce_typed<int16_t, Acc, Op>(a);
    case B200PHI_INT32:
      return launch_reduce_typed<int32_t, Acc, Op>(a);
    case B200PHI_INT64:
      return launch_reduce_typed<long long, Acc, Op>(a);
    case B200PHI_FLOAT32:
      return launch_reduce_typed<float, Acc, Op>(a);
    case B200PHI_FLOAT64:
      return launch_reduce_typed<double, Acc, Op>(a);
    case B200PHI_FLOAT16:
      return launch_reduce_typed<__half, Acc, Op>(a);
    case B200PHI_BFLOAT16:
      return launch_reduce_typed<__nv_bfloat16, Acc, Op>(a);
    default:
      return B200PHI_UNSUPPORTED;
  }
}

// Natural accumulation type only (max/min/prod keep the input dtype).
template <template <typename> class OpT>
int dispatch_reduce_natural(const ReduceArgs& a) {
  switch (a.plan->in_dtype) {
    case B200PHI_BOOL:
      return launch_reduce_typed<Bool8, long long, OpT<long long>>(a);
    case B200PHI_UINT8:
      return launch_reduce_typed<uint8_t, long long, OpT<long long>>(a);
    case B200PHI_INT8:
      return launch_reduce_typed<int8_t, long long, OpT<long long>>(a);
    case B200PHI_INT16:
      return launch_reduce_typed<int16_t, long long, OpT<long long>>(a);
    case B200PHI_INT32:
      return launch_reduce_typed<int32_t, long long, OpT<long long>>(a);
    case B200PHI_INT64:
      return launch_reduce_typed<long long, long long, OpT<long long>>(a);
    case B200PHI_FLOAT32:
      return launch_reduce_typed<float, float, OpT<float>>(a);
    case B200PHI_FLOAT64:
      return launch_reduce_typed<double, double, OpT<double>>(a);
    case B200PHI_FLOAT16:
      return launch_reduce_typed<__half, float, OpT<float>>(a);
    case B200PHI_BFLOAT16:
      return launch_reduce_typed<__nv_bfloat16, float, OpT<float>>(a);
    default:
      return B200PHI_UNSUPPORTED;
  }
}

// one translation unit per op family keeps nvcc parallel
int launch_sum(const ReduceArgs& a);      // SUM, MEAN
int launch_prod(const ReduceArgs& a);     // PROD
int launch_minmax(const ReduceArgs& a);   // MAX, MIN, AMAX, AMIN
int launch_logical(const ReduceArgs& a);  // ANY, ALL
int launch_norm(const ReduceArgs& a);     // SUM_ABS, SUM_SQUARE, L2_NORM, MAX_ABS, MIN_ABS, COUNT_NONZERO

}  // namespace b200phi
