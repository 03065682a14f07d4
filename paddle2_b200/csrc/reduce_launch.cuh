// Typed launchers: ReducePlan -> kernel<<<>>> on the caller's stream.
#pragma once

#include "reduce_kernels.cuh"
#include "reduce_plan.h"
#include "reduce_tiled.cuh"
#include "rows_staged.cuh"

namespace b200phi {

struct ReduceArgs {
  const ReducePlan* plan;
  const void* x;
  void* out;
  void* ws;
  cudaStream_t stream;
};

int cuda_status(cudaError_t e);  // api.cu: records the message, returns B200PHI_CUDA_ERROR

// The tiled kinds (general keep/reduce patterns).  Each op family instantiates them in a TU of its
// own (reduce_*_tiled.cu defines B200PHI_TILED_TU) so the build stays parallel; the typed
// launchers have internal linkage because the two TUs of a family give them different bodies.
template <typename T, typename Acc, typename Op>
static int launch_reduce_tiled(const ReduceArgs& a) {
  const ReducePlan& p = *a.plan;
  const T* x = static_cast<const T*>(a.x);
  const Epilogue ep = make_epilogue(a.out, p.out_dtype, p.op == B200PHI_MEAN, p.red_numel);
  Acc* partial = static_cast<Acc*>(a.ws);
  unsigned* tickets =
      reinterpret_cast<unsigned*>(static_cast<unsigned char*>(a.ws) + p.ws_partial_bytes);
  if (p.ws_ticket_bytes) {
    cudaError_t e = cudaMemsetAsync(tickets, 0, p.ws_ticket_bytes, a.stream);
    if (e != cudaSuccess) return cuda_status(e);
  }
  TiledDims td;
  td.nkeep = p.td_nkeep;
  td.nred = p.td_nred;
  for (int k = 0; k < kTiledMaxDims; ++k) {  // right-aligned: the innermost dim sits at the last slot
    const int ik = p.td_nkeep - kTiledMaxDims + k, ir = p.td_nred - kTiledMaxDims + k;
    td.keep_div[k] = make_fast_div(ik >= 0 ? p.td_keep_extent[ik] : 1u);
    td.keep_stride[k] = ik >= 0 ? p.td_keep_stride[ik] : 0;
    td.red_div[k] = make_fast_div(ir >= 0 ? p.td_red_extent[ir] : 1u);
    td.red_stride[k] = ir >= 0 ? p.td_red_stride[ir] : 0;
  }
  constexpr int VE = VecOf<T>::kElems;
  const dim3 grid(p.grid_x, p.grid_y);
  if (p.kind == kTiledKeep) {
    const unsigned R = static_cast<unsigned>(p.red_numel);
    const unsigned rps = static_cast<unsigned>(p.rows_per_split);
    const size_t smem = static_cast<size_t>(kBlock) * p.vec * sizeof(Acc);
    if (p.vec > 1)
      tile_keep_kernel<T, Acc, Op, VE, 8><<<grid, kBlock, smem, a.stream>>>(x, td, p.nq, p.TX, R, rps, partial, tickets, ep);
    else
      tile_keep_kernel<T, Acc, Op, 1, 8><<<grid, kBlock, smem, a.stream>>>(x, td, p.nq, p.TX, R, rps, partial, tickets, ep);
  } else {
    const FastDiv lv = make_fast_div(p.Lv);
    const unsigned on = static_cast<unsigned>(p.out_numel);
    if (p.vec > 1 && p.G == 256)
      tile_red_kernel<T, Acc, Op, VE, 256, 8><<<grid, kBlock, 0, a.stream>>>(x, td, lv, p.W, p.chunk, on, partial, tickets, ep);
    else if (p.vec > 1)
      tile_red_kernel<T, Acc, Op, VE, 32, 8><<<grid, kBlock, 0, a.stream>>>(x, td, lv, p.W, p.chunk, on, partial, tickets, ep);
    else if (p.G == 256)
      tile_red_kernel<T, Acc, Op, 1, 256, 8><<<grid, kBlock, 0, a.stream>>>(x, td, lv, p.W, p.chunk, on, partial, tickets, ep);
    else
      tile_red_kernel<T, Acc, Op, 1, 32, 8><<<grid, kBlock, 0, a.stream>>>(x, td, lv, p.W, p.chunk, on, partial, tickets, ep);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_status(e);
  return B200PHI_OK;
}

template <typename T, typename Acc, typename Op>
static int launch_reduce_typed(const ReduceArgs& a) {
#ifdef B200PHI_TILED_TU
  return launch_reduce_tiled<T, Acc, Op>(a);
#else
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
    case kRowStaged:
      return launch_rows_staged<T, StagedReduce<T, Acc, Op>>(x, p.rows, static_cast<int>(p.cols), ep, a.stream);
    case kRowPeel: {
      if constexpr (sizeof(T) <= 2) {  // (the planner picks this kind for 1- and 2-byte elements only)
        const int c = static_cast<int>(p.cols);
        constexpr int VE = VecOf<T>::kElems;
        if constexpr (VE <= 8) {
          // (two or four rows per lane group and iteration -- the RB parameter -- measured no
          // better: 101 bf16 columns sum 4.69 / 4.83 / 3.87 TB/s, max 4.79 / 4.01 / 4.00 for RB 1 / 2 / 4)
          if (p.G == 8 && p.nv == 2)
            row_peel_kernel<T, Acc, Op, 8, 2><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
          else if (p.G == 8 && p.nv == 4)
            row_peel_kernel<T, Acc, Op, 8, 4><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
          else if (p.G == 8)
            row_peel_kernel<T, Acc, Op, 8, 8><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
        }
        if (p.G == 16 && p.nv == 4)
          row_peel_kernel<T, Acc, Op, 16, 4><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
        else if (p.G == 16)
          row_peel_kernel<T, Acc, Op, 16, 8><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
        else if (p.G == 32 && p.nv == 16)
          row_peel_kernel<T, Acc, Op, 32, 16><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
        else if (p.G == 32)
          row_peel_kernel<T, Acc, Op, 32, 8><<<p.grid_x, kBlock, 0, a.stream>>>(x, p.rows, c, ep);
      } else {
        return B200PHI_UNSUPPORTED;
      }
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
#endif
}

// in_dtype -> storage type
template <typename Acc, template <typename> class OpT>
static int dispatch_reduce_in(const ReduceArgs& a) {
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
static int dispatch_reduce_natural(const ReduceArgs& a) {
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

// one translation unit per op family (two: classic kinds, tiled kinds) keeps nvcc parallel
#ifdef B200PHI_TILED_TU
#define B200PHI_LAUNCH_NAME(n) n##_tiled
#else
#define B200PHI_LAUNCH_NAME(n) n
#endif
int launch_sum(const ReduceArgs& a);      // SUM, MEAN
int launch_prod(const ReduceArgs& a);     // PROD
int launch_minmax(const ReduceArgs& a);   // MAX, MIN, AMAX, AMIN
int launch_logical(const ReduceArgs& a);  // ANY, ALL
int launch_norm(const ReduceArgs& a);     // SUM_ABS, SUM_SQUARE, L2_NORM, MAX_ABS, MIN_ABS, COUNT_NONZERO
int launch_sum_tiled(const ReduceArgs& a);
int launch_prod_tiled(const ReduceArgs& a);
int launch_minmax_tiled(const ReduceArgs& a);
int launch_logical_tiled(const ReduceArgs& a);
int launch_norm_tiled(const ReduceArgs& a);

}  // namespace b200phi
