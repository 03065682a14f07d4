// Design-space probe #2 (not product code): geometry of the row reduce (C1), the fused
// broadcast add+relu (C2) and the bf16 row statistics (C4).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/microbench2 tools/microbench2.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ float4 ldg16(const void* p) {
  float4 r;
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg16(void* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float warp_sum(float v) {
  for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}

// ---- C1: fp32 row sum, G threads per row, 256-thread CTAs --------------------------------------
template <int G, int U>
__global__ void __launch_bounds__(256) row_sum(const float* __restrict__ x, long long rows, int cols, float* __restrict__ out) {
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
      for (int u = 0; u < U; ++u) b[u] = ldg16(vp + v + u * G);
#pragma unroll
      for (int u = 0; u < U; ++u) { a.x += b[u].x; a.y += b[u].y; a.z += b[u].z; a.w += b[u].w; }
    }
    for (; v < nvec; v += G) { float4 b = ldg16(vp + v); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
    float s = warp_sum((a.x + a.y) + (a.z + a.w));
    if (G == 32) {
      if (g == 0) out[row] = s;
    } else {
      // G/32 warps of this group combine through smem; barrier id = 1 + grp, G threads
      const int w = threadIdx.x >> 5;
      if ((threadIdx.x & 31) == 0) sm[w] = s;
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
      if (g == 0) { float t = 0; for (int i = 0; i < G / 32; ++i) t += sm[grp * (G / 32) + i]; out[row] = t; }
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
    }
  }
}

// ---- C2: out = relu(x + y[channel]); x viewed as [rows, inner], channel = row % nchan ----------------
// variant A: one CTA per (row, chunk)
template <int U>
__global__ void __launch_bounds__(256) addrelu_rows(const float* __restrict__ x, const float* __restrict__ y, float* __restrict__ out,
                                                    int inner_v, int nchan, int chunk_v, int chunks) {
  const int row = blockIdx.x / chunks, chunk = blockIdx.x % chunks;
  const float yv = __ldg(y + row % nchan);
  const float4* xp = reinterpret_cast<const float4*>(x) + (long long)row * inner_v;
  float4* op = reinterpret_cast<float4*>(out) + (long long)row * inner_v;
  const int v0 = chunk * chunk_v;
  int v1 = v0 + chunk_v; if (v1 > inner_v) v1 = inner_v;
  for (int vb = v0 + threadIdx.x; vb < v1; vb += U * 256) {
    float4 b[U];
#pragma unroll
    for (int u = 0; u < U; ++u) if (vb + u * 256 < v1) b[u] = ldg16(xp + vb + u * 256);
#pragma unroll
    for (int u = 0; u < U; ++u) if (vb + u * 256 < v1) {
      float4 r;
      r.x = fmaxf(b[u].x + yv, 0.f); r.y = fmaxf(b[u].y + yv, 0.f); r.z = fmaxf(b[u].z + yv, 0.f); r.w = fmaxf(b[u].w + yv, 0.f);
      stg16(op + vb + u * 256, r);
    }
  }
}
// variant B: flat grid-stride over vectors, row recovered by one division per vector
template <int U>
__global__ void __launch_bounds__(256) addrelu_flat(const float* __restrict__ x, const float* __restrict__ y, float* __restrict__ out,
                                                    long long nvec, unsigned inner_v, unsigned nchan) {
  const float4* xp = reinterpret_cast<const float4*>(x);
  float4* op = reinterpret_cast<float4*>(out);
  for (long long vb = (long long)blockIdx.x * (256 * U) + threadIdx.x; vb < nvec; vb += (long long)gridDim.x * 256 * U) {
    float4 b[U];
    float yv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long v = vb + u * 256;
      if (v < nvec) { b[u] = ldg16(xp + v); yv[u] = __ldg(y + (unsigned)(v / inner_v) % nchan); }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long v = vb + u * 256;
      if (v < nvec) {
        float4 r;
        r.x = fmaxf(b[u].x + yv[u], 0.f); r.y = fmaxf(b[u].y + yv[u], 0.f); r.z = fmaxf(b[u].z + yv[u], 0.f); r.w = fmaxf(b[u].w + yv[u], 0.f);
        stg16(op + v, r);
      }
    }
  }
}
// plain copy for reference
template <int U>
__global__ void __launch_bounds__(256) copy_flat(const float4* __restrict__ x, float4* __restrict__ out, long long nvec) {
  for (long long vb = (long long)blockIdx.x * (256 * U) + threadIdx.x; vb < nvec; vb += (long long)gridDim.x * 256 * U) {
    float4 b[U];
#pragma unroll
    for (int u = 0; u < U; ++u) if (vb + u * 256 < nvec) b[u] = ldg16(x + vb + u * 256);
#pragma unroll
    for (int u = 0; u < U; ++u) if (vb + u * 256 < nvec) stg16(out + vb + u * 256, b[u]);
  }
}

// ---- C4: bf16 row mean+var, G threads per row, NV vectors per thread in registers ----------------
template <int G, int NV>
__global__ void __launch_bounds__(256) rowstats(const __nv_bfloat16* __restrict__ x, long long rows, int cols,
                                                __nv_bfloat16* __restrict__ mean_out, __nv_bfloat16* __restrict__ var_out) {
  __shared__ float sm[8];
  constexpr int GROUPS = 256 / G;
  const int g = threadIdx.x % G, grp = threadIdx.x / G, w = threadIdx.x >> 5;
  const int nvec = cols / 8;
  for (long long row = (long long)blockIdx.x * GROUPS + grp; row < rows; row += (long long)gridDim.x * GROUPS) {
    const float4* vp = reinterpret_cast<const float4*>(x + row * cols);
    float4 b[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) if (g + k * G < nvec) b[k] = ldg16(vp + g + k * G);
    float s = 0;
#pragma unroll
    for (int k = 0; k < NV; ++k) if (g + k * G < nvec) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&b[k]);
#pragma unroll
      for (int e = 0; e < 4; ++e) { float2 f = __bfloat1622float2(h[e]); s += f.x + f.y; }
    }
    s = warp_sum(s);
    if (G > 32) {
      if ((threadIdx.x & 31) == 0) sm[w] = s;
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
      s = 0; for (int i = 0; i < G / 32; ++i) s += sm[grp * (G / 32) + i];
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
    }
    const float mean = s / cols;
    float q = 0;
#pragma unroll
    for (int k = 0; k < NV; ++k) if (g + k * G < nvec) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&b[k]);
#pragma unroll
      for (int e = 0; e < 4; ++e) { float2 f = __bfloat1622float2(h[e]); float d0 = f.x - mean, d1 = f.y - mean; q += d0 * d0 + d1 * d1; }
    }
    q = warp_sum(q);
    if (G > 32) {
      if ((threadIdx.x & 31) == 0) sm[w] = q;
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
      q = 0; for (int i = 0; i < G / 32; ++i) q += sm[grp * (G / 32) + i];
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G));
    }
    if (g == 0) { mean_out[row] = __float2bfloat16_rn(mean); var_out[row] = __float2bfloat16_rn(q / (cols - 1)); }
  }
}

template <typename F> float time_it(F fn, int iters = 20) {
  for (int i = 0; i < 3; ++i) fn();
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
  // ---------------- C1 ----------------
  {
    const long long rows = 32768; const int cols = 4096;
    float *x, *out;
    CK(cudaMalloc(&x, rows * cols * 4)); CK(cudaMalloc(&out, rows * 4));
    CK(cudaMemset(x, 0, rows * cols * 4));
    const double bytes = (double)rows * cols * 4 + rows * 4;
#define ROW(G, U, CTAS) { int ctas = CTAS; float ms = time_it([&] { row_sum<G, U><<<ctas, 256>>>(x, rows, cols, out); }); \
    printf("C1 row_sum G=%-3d U=%d ctas=%-6d %7.2f us %7.1f GB/s\n", G, U, ctas, ms * 1e3, bytes / ms / 1e6); }
    ROW(32, 8, 4096) ROW(32, 8, 148 * 8) ROW(32, 8, 148 * 4) ROW(32, 8, 148 * 16) ROW(32, 4, 4096) ROW(32, 16, 4096)
    ROW(64, 8, 8192) ROW(64, 8, 148 * 8) ROW(64, 8, 148 * 4) ROW(64, 4, 8192)
    ROW(128, 8, 16384) ROW(128, 8, 148 * 8) ROW(128, 8, 148 * 4) ROW(128, 4, 148 * 8)
    ROW(256, 4, 32768) ROW(256, 4, 148 * 8) ROW(256, 4, 148 * 4) ROW(256, 4, 148 * 16) ROW(256, 2, 148 * 8)
    cudaFree(x); cudaFree(out);
  }
  // ---------------- C2 ----------------
  {
    const int rows = 256 * 64, inner = 112 * 112, nchan = 64;
    const long long n = (long long)rows * inner;
    float *x, *y, *out;
    CK(cudaMalloc(&x, n * 4)); CK(cudaMalloc(&out, n * 4)); CK(cudaMalloc(&y, nchan * 4));
    CK(cudaMemset(x, 0, n * 4)); CK(cudaMemset(y, 0, nchan * 4));
    const double bytes = 2.0 * n * 4 + nchan * 4;
    const int inner_v = inner / 4;
#define ADDR(U, CH) { int chunks = CH; int chunk_v = (inner_v + chunks - 1) / chunks; float ms = time_it([&] { addrelu_rows<U><<<rows * chunks, 256>>>(x, y, out, inner_v, nchan, chunk_v, chunks); }); \
    printf("C2 addrelu_rows U=%d chunks/row=%d ctas=%-7d %7.2f us %7.1f GB/s\n", U, chunks, rows * chunks, ms * 1e3, bytes / ms / 1e6); }
    ADDR(4, 1) ADDR(4, 2) ADDR(4, 4) ADDR(8, 1) ADDR(8, 2) ADDR(2, 2) ADDR(2, 4) ADDR(1, 4)
#define ADDF(U, CTAS) { int ctas = CTAS; float ms = time_it([&] { addrelu_flat<U><<<ctas, 256>>>(x, y, out, n / 4, inner_v, nchan); }); \
    printf("C2 addrelu_flat U=%d ctas=%-7d %7.2f us %7.1f GB/s\n", U, ctas, ms * 1e3, bytes / ms / 1e6); }
    ADDF(4, 148 * 4) ADDF(4, 148 * 8) ADDF(4, 148 * 16) ADDF(8, 148 * 4) ADDF(8, 148 * 8) ADDF(2, 148 * 8) ADDF(2, 148 * 16) ADDF(4, 148 * 64) ADDF(4, (int)((n / 4 + 1023) / 1024)) ADDF(8, (int)((n / 4 + 2047) / 2048)) ADDF(1, 148 * 32)
#define CPY(U, CTAS) { int ctas = CTAS; float ms = time_it([&] { copy_flat<U><<<ctas, 256>>>((const float4*)x, (float4*)out, n / 4); }); \
    printf("C2 copy_flat    U=%d ctas=%-7d %7.2f us %7.1f GB/s\n", U, ctas, ms * 1e3, bytes / ms / 1e6); }
    CPY(4, 148 * 4) CPY(4, 148 * 8) CPY(4, 148 * 16) CPY(8, 148 * 8) CPY(4, (int)((n / 4 + 1023) / 1024)) CPY(2, 148 * 16) CPY(1, 148 * 32)
    { float ms = time_it([&] { cudaMemcpyAsync(out, x, n * 4, cudaMemcpyDeviceToDevice); }); printf("C2 cudaMemcpy D2D %7.2f us %7.1f GB/s\n", ms * 1e3, bytes / ms / 1e6); }
    cudaFree(x); cudaFree(y); cudaFree(out);
  }
  // ---------------- C4 ----------------
  {
    const long long rows = 16384; const int cols = 8192;
    __nv_bfloat16 *x, *m, *v;
    CK(cudaMalloc(&x, rows * cols * 2)); CK(cudaMalloc(&m, rows * 2)); CK(cudaMalloc(&v, rows * 2));
    CK(cudaMemset(x, 0, rows * cols * 2));
    const double bytes = (double)rows * cols * 2 + rows * 4;
#define RS(G, NV, CTAS) { int ctas = CTAS; float ms = time_it([&] { rowstats<G, NV><<<ctas, 256>>>(x, rows, cols, m, v); }); \
    printf("C4 rowstats G=%-3d NV=%-2d ctas=%-6d %7.2f us %7.1f GB/s\n", G, NV, ctas, ms * 1e3, bytes / ms / 1e6); }
    RS(256, 4, 16384) RS(256, 4, 148 * 8) RS(256, 4, 148 * 4) RS(128, 8, 8192) RS(128, 8, 148 * 8) RS(128, 8, 148 * 4)
    RS(64, 16, 4096) RS(64, 16, 148 * 8) RS(64, 16, 148 * 4) RS(32, 32, 2048) RS(32, 32, 148 * 4)
    cudaFree(x); cudaFree(m); cudaFree(v);
  }
  printf("done\n");
  return 0;
}
