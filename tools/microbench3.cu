// Probe #3 (not product code): the library's own row_reduce_kernel template in several
// configurations next to the bare microbench loop, on the C1 shape.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "../include/b200phi.h"
#include "../paddle2_b200/csrc/reduce_kernels.cuh"

using namespace b200phi;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ float4 ldg16f(const void* p) {
  float4 r;
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
template <int G, int U>
__global__ void __launch_bounds__(256) bare_row_sum(const float* __restrict__ x, long long rows, int cols, float* __restrict__ out) {
  __shared__ float sm[8];
  constexpr int GROUPS = 256 / G;
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  const int nvec = cols / 4;
  for (long long row = (long long)blockIdx.x * GROUPS + grp; row < rows; row += (long long)gridDim.x * GROUPS) {
    const float4* vp = reinterpret_cast<const float4*>(x + row * cols);
    float4 a = make_float4(0, 0, 0, 0);
    int v = g;
    for (; v + (U - 1) * G < nvec; v += U * G) {
      float4 b[U];
#pragma unroll
      for (int u = 0; u < U; ++u) b[u] = ldg16f(vp + v + u * G);
#pragma unroll
      for (int u = 0; u < U; ++u) { a.x += b[u].x; a.y += b[u].y; a.z += b[u].z; a.w += b[u].w; }
    }
    for (; v < nvec; v += G) { float4 b = ldg16f(vp + v); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
    float s = (a.x + a.y) + (a.z + a.w);
    for (int m = 16; m > 0; m >>= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
    if (G == 32) { if (g == 0) out[row] = s; }
    else {
      const int w = threadIdx.x >> 5;
      if ((threadIdx.x & 31) == 0) sm[w] = s;
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
      if (g == 0) { float t = 0; for (int i = 0; i < G / 32; ++i) t += sm[grp * (G / 32) + i]; out[row] = t; }
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
    }
  }
}

// library span, but the store is a plain float store (no dtype switch / mean division)
template <int G, int U>
__global__ void __launch_bounds__(kBlock) lib_span_plain_store(const float* __restrict__ x, long long rows, long long cols, float* out) {
  __shared__ float smem[kWarps];
  constexpr int kRowsPerCta = kBlock / G;
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  for (long long row = (long long)blockIdx.x * kRowsPerCta + grp; row < rows; row += (long long)gridDim.x * kRowsPerCta) {
    float v = reduce_span<float, float, SumOp<float>, G, U>(x + row * cols, cols, g);
    v = group_combine<float, SumOp<float>, G>(v, smem);
    if (g == 0) out[row] = v;
  }
}

// progressive variants: which ingredient of the library kernel costs what
template <int G, int U, int VAR>
__global__ void __launch_bounds__(kBlock) variant(const float* __restrict__ x, long long rows, long long cols, float* out, Epilogue ep) {
  __shared__ float smem[2 * kWarps];
  constexpr int kRowsPerCta = kBlock / G;
  const int g = threadIdx.x % G, grp = threadIdx.x / G;
  int parity = 0;
  for (long long row = (long long)blockIdx.x * kRowsPerCta + grp; row < rows; row += (long long)gridDim.x * kRowsPerCta) {
    float v;
    if (VAR & 1) {  // library aligned span
      v = reduce_span_aligned<float, float, SumOp<float>, G, U>(x + row * cols, cols / 4, g);
    } else {        // bare loop
      const float4* vp = reinterpret_cast<const float4*>(x + row * cols);
      const int nvec = (int)(cols / 4);
      float4 a = make_float4(0, 0, 0, 0);
      int vv = g;
      for (; vv + (U - 1) * G < nvec; vv += U * G) {
        float4 b[U];
#pragma unroll
        for (int u = 0; u < U; ++u) b[u] = ldg16f(vp + vv + u * G);
#pragma unroll
        for (int u = 0; u < U; ++u) { a.x += b[u].x; a.y += b[u].y; a.z += b[u].z; a.w += b[u].w; }
      }
      for (; vv < nvec; vv += G) { float4 b = ldg16f(vp + vv); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
      v = (a.x + a.y) + (a.z + a.w);
    }
    if (VAR & 2) {  // library leader combine
      v = group_combine_leader<float, SumOp<float>, G>(v, smem, parity);
      parity ^= 1;
    } else {        // bare combine
      for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
      const int w = threadIdx.x >> 5;
      if ((threadIdx.x & 31) == 0) smem[w] = v;
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
      if (g == 0) { float t = 0; for (int i = 0; i < G / 32; ++i) t += smem[grp * (G / 32) + i]; v = t; }
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
    }
    if (g == 0) {
      if (VAR & 4) store_result(ep, row, v); else out[row] = v;
    }
  }
}

template <typename F> float time_it(F fn, int iters = 20) {
  for (int i = 0; i < 5; ++i) fn();
  CK(cudaDeviceSynchronize());
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  for (int i = 0; i < iters; ++i) fn();
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms; cudaEventElapsedTime(&ms, a, b);
  CK(cudaGetLastError());
  return ms / iters;
}

int main() {
  const long long rows = 32768; const int cols = 4096;
  float *x, *out;
  CK(cudaMalloc(&x, rows * cols * 4)); CK(cudaMalloc(&out, rows * 4));
  CK(cudaMemset(x, 0, rows * cols * 4));
  const double bytes = (double)rows * cols * 4 + rows * 4;
  Epilogue ep = make_epilogue(out, B200PHI_FLOAT32, false, cols);
#define P(name, ms) printf("%-44s %7.2f us %7.1f GB/s\n", name, ms * 1e3, bytes / ms / 1e6)
#define VARIANT(G, U, V, desc) P(desc, time_it([&] { variant<G, U, V><<<(unsigned)(rows / (256 / G)), 256>>>(x, rows, cols, out, ep); }))
  VARIANT(256, 4, 0, "v0 bare span, bare combine, plain store");
  VARIANT(256, 4, 1, "v1 LIB span, bare combine, plain store");
  VARIANT(256, 4, 2, "v2 bare span, LIB combine, plain store");
  VARIANT(256, 4, 4, "v4 bare span, bare combine, LIB store");
  VARIANT(256, 4, 3, "v3 LIB span, LIB combine, plain store");
  VARIANT(256, 4, 7, "v7 all LIB");
  VARIANT(256, 8, 7, "v7 all LIB U=8");
  VARIANT(128, 8, 0, "G=128 v0");
  VARIANT(128, 8, 7, "G=128 v7 all LIB");
  for (int rep = 0; rep < 1; ++rep) {
    P("bare G=256 U=4 ctas=32768", time_it([&] { bare_row_sum<256, 4><<<32768, 256>>>(x, rows, cols, out); }));
    P("bare G=128 U=8 ctas=16384", time_it([&] { bare_row_sum<128, 8><<<16384, 256>>>(x, rows, cols, out); }));
    P("bare G=32  U=8 ctas=4096", time_it([&] { bare_row_sum<32, 8><<<4096, 256>>>(x, rows, cols, out); }));
    P("lib kernel G=256 U=8 ctas=32768", time_it([&] { row_reduce_kernel<float, float, SumOp<float>, 256, 8, true><<<32768, 256>>>(x, rows, cols, ep); }));
    P("lib kernel G=256 U=4 ctas=32768", time_it([&] { row_reduce_kernel<float, float, SumOp<float>, 256, 4, true><<<32768, 256>>>(x, rows, cols, ep); }));
    P("lib kernel G=128 U=8 ctas=16384", time_it([&] { row_reduce_kernel<float, float, SumOp<float>, 128, 8, true><<<16384, 256>>>(x, rows, cols, ep); }));
    P("lib kernel G=64  U=8 ctas=8192", time_it([&] { row_reduce_kernel<float, float, SumOp<float>, 64, 8, true><<<8192, 256>>>(x, rows, cols, ep); }));
    P("lib kernel G=32  U=8 ctas=4096", time_it([&] { row_reduce_kernel<float, float, SumOp<float>, 32, 8, true><<<4096, 256>>>(x, rows, cols, ep); }));
    P("lib span + plain store G=256 U=4", time_it([&] { lib_span_plain_store<256, 4><<<32768, 256>>>(x, rows, cols, out); }));
    P("lib span + plain store G=256 U=8", time_it([&] { lib_span_plain_store<256, 8><<<32768, 256>>>(x, rows, cols, out); }));
    P("lib span + plain store G=128 U=8", time_it([&] { lib_span_plain_store<128, 8><<<16384, 256>>>(x, rows, cols, out); }));
  }
  printf("done\n");
  return 0;
}
