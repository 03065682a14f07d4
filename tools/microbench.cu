// Design-space probe for the streaming reduce kernels (not product code, not linked into
// libb200phi.so): times LDG.128 and TMA-bulk (cp.async.bulk + mbarrier ring) variants of the
// C5 column reduce [R, 4096] fp32 -> [4096] and prints algorithmic GB/s for each.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/microbench tools/microbench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ float4 ldg16(const float4* p) {
  float4 r;
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}

// ---------------- LDG variant: block = TX threads over float4 columns, grid (tiles, splits) ----
template <int THREADS, int U>
__global__ void __launch_bounds__(THREADS) col_ldg(const float4* __restrict__ x, long long R, int cols_v,
                                                   long long rows_per_split, float4* __restrict__ partial) {
  const int cv = blockIdx.x * THREADS + threadIdx.x;
  if (cv >= cols_v) return;
  long long r0 = (long long)blockIdx.y * rows_per_split, r1 = r0 + rows_per_split;
  if (r1 > R) r1 = R;
  float4 acc = make_float4(0, 0, 0, 0);
  const float4* base = x + cv;
  long long r = r0;
  for (; r + U <= r1; r += U) {
    float4 b[U];
#pragma unroll
    for (int u = 0; u < U; ++u) b[u] = ldg16(base + (r + u) * cols_v);
#pragma unroll
    for (int u = 0; u < U; ++u) { acc.x += b[u].x; acc.y += b[u].y; acc.z += b[u].z; acc.w += b[u].w; }
  }
  for (; r < r1; ++r) { float4 b = ldg16(base + r * cols_v); acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w; }
  partial[(long long)blockIdx.y * cols_v + cv] = acc;
}

// second pass: sum `splits` partial rows
__global__ void fold(const float4* __restrict__ partial, int splits, int cols_v, float4* __restrict__ out) {
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  if (cv >= cols_v) return;
  float4 acc = make_float4(0, 0, 0, 0);
  for (int s = 0; s < splits; ++s) { float4 b = partial[(long long)s * cols_v + cv]; acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w; }
  out[cv] = acc;
}

// ---------------- TMA bulk variant ----------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\nbra WAIT_LOOP;\nDONE:\n}\n" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// One CTA streams rows [r0, r1) of a [R, COLS] fp32 matrix, full width, STAGE_ROWS rows per stage.
// Warp CONSUMERS/32 is the producer (one elected lane); CONSUMERS threads own float4 columns.
template <int CONSUMERS, int STAGES, int STAGE_ROWS, int COLS>
__global__ void __launch_bounds__(CONSUMERS + 32) col_tma(const float* __restrict__ x, long long R,
                                                           long long rows_per_cta, float4* __restrict__ partial) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int ROW_BYTES = COLS * 4;
  constexpr int STAGE_BYTES = STAGE_ROWS * ROW_BYTES;
  constexpr int VPT = COLS / 4 / CONSUMERS;  // float4 columns per thread
  float* data = reinterpret_cast<float*>(smem);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  long long r0 = (long long)blockIdx.x * rows_per_cta, r1 = r0 + rows_per_cta;
  if (r1 > R) r1 = R;
  const long long nrows = r1 > r0 ? r1 - r0 : 0;
  const long long nstage = (nrows + STAGE_ROWS - 1) / STAGE_ROWS;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], CONSUMERS / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x >= CONSUMERS) {
    if (threadIdx.x == CONSUMERS) {
      for (long long it = 0; it < nstage; ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        if (it >= STAGES) mbar_wait(&empty[s], ph ^ 1);
        long long rows = nrows - it * STAGE_ROWS;
        if (rows > STAGE_ROWS) rows = STAGE_ROWS;
        const uint32_t bytes = (uint32_t)(rows * ROW_BYTES);
        mbar_expect_tx(&full[s], bytes);
        bulk_g2s(data + (size_t)s * (STAGE_BYTES / 4), x + (r0 + it * STAGE_ROWS) * COLS, bytes, &full[s]);
      }
    }
    return;
  }
  float4 acc[VPT];
#pragma unroll
  for (int v = 0; v < VPT; ++v) acc[v] = make_float4(0, 0, 0, 0);
  for (long long it = 0; it < nstage; ++it) {
    const int s = it % STAGES;
    const uint32_t ph = (it / STAGES) & 1;
    mbar_wait(&full[s], ph);
    long long rows = nrows - it * STAGE_ROWS;
    if (rows > STAGE_ROWS) rows = STAGE_ROWS;
    const float4* sd = reinterpret_cast<const float4*>(data + (size_t)s * (STAGE_BYTES / 4));
    if (rows == STAGE_ROWS) {
#pragma unroll
      for (int rr = 0; rr < STAGE_ROWS; ++rr)
#pragma unroll
        for (int v = 0; v < VPT; ++v) {
          float4 b = sd[rr * (COLS / 4) + v * CONSUMERS + threadIdx.x];
          acc[v].x += b.x; acc[v].y += b.y; acc[v].z += b.z; acc[v].w += b.w;
        }
    } else {
      for (int rr = 0; rr < rows; ++rr)
#pragma unroll
        for (int v = 0; v < VPT; ++v) {
          float4 b = sd[rr * (COLS / 4) + v * CONSUMERS + threadIdx.x];
          acc[v].x += b.x; acc[v].y += b.y; acc[v].z += b.z; acc[v].w += b.w;
        }
    }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(&empty[s]);
  }
#pragma unroll
  for (int v = 0; v < VPT; ++v) partial[(long long)blockIdx.x * (COLS / 4) + v * CONSUMERS + threadIdx.x] = acc[v];
}

// ---------------- pure read ceiling: every CTA streams a contiguous slab, sums into one float --------
template <int THREADS, int U>
__global__ void __launch_bounds__(THREADS) read_all(const float4* __restrict__ x, long long nvec, float* out) {
  float4 acc = make_float4(0, 0, 0, 0);
  const long long per = (nvec + gridDim.x - 1) / gridDim.x;
  long long v0 = per * blockIdx.x, v1 = v0 + per;
  if (v1 > nvec) v1 = nvec;
  long long v = v0 + threadIdx.x;
  for (; v + (long long)(U - 1) * THREADS < v1; v += (long long)U * THREADS) {
    float4 b[U];
#pragma unroll
    for (int u = 0; u < U; ++u) b[u] = ldg16(x + v + (long long)u * THREADS);
#pragma unroll
    for (int u = 0; u < U; ++u) { acc.x += b[u].x; acc.y += b[u].y; acc.z += b[u].z; acc.w += b[u].w; }
  }
  for (; v < v1; v += THREADS) { float4 b = ldg16(x + v); acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w; }
  float s = acc.x + acc.y + acc.z + acc.w;
  for (int m = 16; m > 0; m >>= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
  if ((threadIdx.x & 31) == 0) atomicAdd(out, s);
}

__global__ void fill_random(float* x, long long n, unsigned seed) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    unsigned long long z = (unsigned long long)i * 0x9E3779B97F4A7C15ull + seed;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    x[i] = (float)(z >> 40) * (1.0f / 16777216.0f);
  }
}

static float time_it(void (*fn)(void*), void* ctx, int iters) {
  for (int i = 0; i < 3; ++i) fn(ctx);
  CK(cudaDeviceSynchronize());
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  for (int i = 0; i < iters; ++i) fn(ctx);
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms / iters;
}

struct Ctx { const float* x; long long R; float4* partial; float4* out; int a, b, c; float* scalar; };
static const int COLS = 4096;

template <int THREADS, int U> void run_ldg(void* p) {
  Ctx* c = (Ctx*)p;
  const int cols_v = COLS / 4, tiles = (cols_v + THREADS - 1) / THREADS, splits = c->a;
  long long rps = (c->R + splits - 1) / splits;
  col_ldg<THREADS, U><<<dim3(tiles, splits), THREADS>>>((const float4*)c->x, c->R, cols_v, rps, c->partial);
  fold<<<(cols_v + 255) / 256, 256>>>(c->partial, splits, cols_v, c->out);
}
template <int CONS, int STAGES, int SR> void run_tma(void* p) {
  Ctx* c = (Ctx*)p;
  const int ctas = c->a;
  long long rpc = (c->R + ctas - 1) / ctas;
  size_t smem = (size_t)STAGES * SR * COLS * 4 + 2 * STAGES * 8;
  static bool set = false;
  if (!set) { CK(cudaFuncSetAttribute(col_tma<CONS, STAGES, SR, COLS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); set = true; }
  col_tma<CONS, STAGES, SR, COLS><<<ctas, CONS + 32, smem>>>(c->x, c->R, rpc, c->partial);
  fold<<<(COLS / 4 + 255) / 256, 256>>>(c->partial, ctas, COLS / 4, c->out);
}
template <int THREADS, int U> void run_read(void* p) {
  Ctx* c = (Ctx*)p;
  read_all<THREADS, U><<<c->a, THREADS>>>((const float4*)c->x, c->R * COLS / 4, c->scalar);
}

int main(int argc, char** argv) {
  long long R = argc > 1 ? atoll(argv[1]) : 524288;
  const int fill = argc > 2 ? atoi(argv[2]) : 1;  // 0 zeros, 1 low-entropy pattern, 2 random
  const int quick = argc > 3 ? atoi(argv[3]) : 0;
  printf("fill mode %d\n", fill);
  const double bytes = (double)R * COLS * 4;
  float* x;
  CK(cudaMalloc(&x, (size_t)bytes));
  CK(cudaMemset(x, 0, (size_t)bytes));
  if (fill == 2) { fill_random<<<148 * 8, 256>>>(x, R * COLS, 123u); CK(cudaDeviceSynchronize()); }
  if (fill == 1) {  // x[r][c] = ((131 r + 7 c) mod 1024)/1024 -> every column sums to R/1024*511.5
    std::vector<float> row(COLS * 1024);
    for (int r = 0; r < 1024; ++r) for (int c = 0; c < COLS; ++c) row[(size_t)r * COLS + c] = ((131 * r + 7 * c) % 1024) / 1024.0f;
    for (long long r0 = 0; r0 < R; r0 += 1024) CK(cudaMemcpy(x + r0 * COLS, row.data(), sizeof(float) * COLS * 1024, cudaMemcpyHostToDevice));
  }
  float4 *partial, *out; float* scalar;
  CK(cudaMalloc(&partial, (size_t)8192 * COLS * 4));
  CK(cudaMalloc(&out, COLS * 4));
  CK(cudaMalloc(&scalar, 4));
  Ctx c{x, R, partial, out, 0, 0, 0, scalar};
  const double want = (double)(R / 1024) * 511.5;
  auto check = [&](const char* name) {
    std::vector<float> h(COLS);
    CK(cudaMemcpy(h.data(), out, COLS * 4, cudaMemcpyDeviceToHost));
    double worst = 0;
    for (int i = 0; i < COLS; ++i) { double e = fabs(h[i] - want) / want; if (e > worst) worst = e; }
    if (fill == 1 && worst > 1e-4) printf("   !! %s wrong result: rel err %.3g\n", name, worst);
  };
  const int iters = 10;
#define LDG(T, U, S) { c.a = S; float ms = time_it(run_ldg<T, U>, &c, iters); printf("ldg  threads=%-4d U=%-2d splits=%-5d ctas=%-6d  %8.3f ms  %7.1f GB/s\n", T, U, S, S * ((COLS / 4 + T - 1) / T), ms, bytes / ms / 1e6); check("ldg"); }
#define TMA(C_, ST, SR, N) { c.a = N; float ms = time_it(run_tma<C_, ST, SR>, &c, iters); printf("tma  consumers=%-4d stages=%d stage_rows=%d ctas=%-5d smem=%3dKB  %8.3f ms  %7.1f GB/s\n", C_, ST, SR, N, (ST * SR * 16), ms, bytes / ms / 1e6); check("tma"); }
#define RD(T, U, N) { c.a = N; float ms = time_it(run_read<T, U>, &c, iters); printf("read threads=%-4d U=%-2d ctas=%-6d  %8.3f ms  %7.1f GB/s\n", T, U, N, ms, bytes / ms / 1e6); }
  if (quick) {
    RD(256, 8, 148 * 32) LDG(256, 8, 148) LDG(1024, 4, 148) LDG(256, 8, 296) TMA(256, 6, 2, 148)
    printf("done\n");
    return 0;
  }
  RD(256, 8, 148 * 8) RD(256, 16, 148 * 8) RD(512, 8, 148 * 4) RD(256, 8, 148 * 32) RD(256, 4, 148 * 8) RD(1024, 4, 148 * 2) RD(256, 8, 148 * 64)
  LDG(256, 8, 296) LDG(256, 8, 148) LDG(256, 8, 592) LDG(256, 8, 1184) LDG(256, 8, 2368)
  LDG(256, 16, 296) LDG(256, 16, 592) LDG(256, 4, 296) LDG(256, 4, 592)
  LDG(512, 8, 296) LDG(512, 8, 592) LDG(1024, 4, 148) LDG(1024, 4, 296) LDG(1024, 8, 296) LDG(128, 8, 296) LDG(128, 16, 592)
  TMA(256, 6, 2, 148) TMA(256, 3, 4, 148) TMA(256, 12, 1, 148) TMA(512, 6, 2, 148)
  TMA(256, 3, 2, 296) TMA(256, 6, 1, 296) TMA(512, 3, 2, 296) TMA(256, 4, 1, 444) TMA(256, 2, 2, 444) TMA(256, 6, 2, 1480)
  TMA(512, 6, 2, 592)
  printf("done\n");
  return 0;
}
