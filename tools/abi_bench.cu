// C++ driver of the C ABI (no Python, no torch): times b200phi_* calls on the five configs
// with cudaMalloc'd buffers and CUDA events.  Cross-checks bench.py's numbers.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/abi_bench tools/abi_bench.cu -Iinclude -Lpaddle2_b200 -lb200phi -Xlinker -rpath -Xlinker '$ORIGIN/../paddle2_b200'
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200phi.h"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
#define OK(x) do { int rc = (x); if (rc) { printf("b200phi status %d: %s (line %d)\n", rc, b200phi_last_error(), __LINE__); exit(1);} } while (0)

__global__ void fill_random(unsigned* x, long long n, unsigned seed, unsigned mask, unsigned orv) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    unsigned long long z = (unsigned long long)i * 0x9E3779B97F4A7C15ull + seed;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    x[i] = ((unsigned)(z >> 32) & mask) | orv;
  }
}

template <typename F> float time_it(F fn, int iters) {
  // ABI_BENCH_ONCE=1: one launch per kernel and no warm-up, for captures under ncu
  static const bool once = getenv("ABI_BENCH_ONCE") != nullptr;
  if (once) iters = 1;
  for (int i = 0; i < (once ? 0 : 5); ++i) fn();
  CK(cudaDeviceSynchronize());
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  for (int i = 0; i < iters; ++i) fn();
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms / iters;
}

static void* dalloc(size_t bytes, unsigned mask, unsigned orv) {
  void* p; CK(cudaMalloc(&p, bytes));
  fill_random<<<148 * 8, 256>>>((unsigned*)p, (long long)(bytes / 4), 123u, mask, orv);
  CK(cudaDeviceSynchronize());
  return p;
}

int main(int argc, char** argv) {
  const char* only = argc > 1 ? argv[1] : "";
  printf("%s\n", b200phi_version());
  const unsigned F01_MASK = 0x007fffff, F01_OR = 0x3f000000;  // floats in [0.5, 1)
  if (!*only || strstr(only, "C1")) {
    float* x = (float*)dalloc((size_t)32 * 1024 * 4096 * 4, F01_MASK, F01_OR);
    float* out; CK(cudaMalloc(&out, 32 * 1024 * 4));
    b200phi_reduce_desc d = {};
    d.op = B200PHI_SUM; d.in_dtype = d.out_dtype = B200PHI_FLOAT32; d.rank = 3;
    d.dims[0] = 32; d.dims[1] = 1024; d.dims[2] = 4096; d.naxes = 1; d.axes[0] = 2; d.x = x; d.out = out;
    float ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 20);
    printf("C1 sum fp32 [32,1024,4096] axis -1      %8.2f us  %7.1f GB/s\n", ms * 1e3, 537001984.0 / ms / 1e6);
    cudaFree(x); cudaFree(out);
  }
  if (!*only || strstr(only, "C2")) {
    const size_t n = (size_t)256 * 64 * 112 * 112;
    float* x = (float*)dalloc(n * 4, 0x807fffff, 0x3f000000);
    float* y = (float*)dalloc(64 * 4, 0x807fffff, 0x3f000000);
    float* out; CK(cudaMalloc(&out, n * 4));
    b200phi_elementwise_desc e = {};
    e.binary = B200PHI_ADD; e.act = B200PHI_ACT_RELU; e.dtype = B200PHI_FLOAT32; e.arity = 2; e.axis = -1;
    e.in[0].rank = 4; e.in[0].dims[0] = 256; e.in[0].dims[1] = 64; e.in[0].dims[2] = 112; e.in[0].dims[3] = 112; e.in[0].data = x;
    e.in[1].rank = 3; e.in[1].dims[0] = 64; e.in[1].dims[1] = 1; e.in[1].dims[2] = 1; e.in[1].data = y;
    e.out_rank = 4; memcpy(e.out_dims, e.in[0].dims, sizeof(e.out_dims)); e.out = out;
    float ms = time_it([&] { OK(b200phi_elementwise(&e, 0)); }, 20);
    printf("C2 relu(x+y) fp32 [256,64,112,112]+[64,1,1] %8.2f us  %7.1f GB/s\n", ms * 1e3, 1644167424.0 / ms / 1e6);
    cudaFree(x); cudaFree(y); cudaFree(out);
  }
  if (!*only || strstr(only, "C3")) {
    const size_t n = (size_t)8192 * 32000;
    void* x = dalloc(n * 8, 0xffffffff, 0);
    long long *idx, *val; CK(cudaMalloc(&idx, 8192 * 8)); CK(cudaMalloc(&val, 8192 * 8));
    b200phi_argreduce_desc a = {};
    a.op = B200PHI_ARGMAX; a.in_dtype = B200PHI_INT64; a.index_dtype = B200PHI_INT64; a.rank = 2;
    a.dims[0] = 8192; a.dims[1] = 32000; a.axis = 1; a.x = x; a.out = idx; a.max_out = val;
    float ms = 0;
    const bool only_b = strstr(only, "C3b") != nullptr;
    if (!only_b) {
      ms = time_it([&] { OK(b200phi_argreduce(&a, 0)); }, 20);
      printf("C3a argmax+max int64 [8192,32000]        %8.2f us  %7.1f GB/s\n", ms * 1e3, 2097283072.0 / ms / 1e6);
    }
    a.in_dtype = B200PHI_BFLOAT16;   // reuse the buffer as bf16 bits; mask NaN/Inf exponents is not needed for timing
    ms = time_it([&] { OK(b200phi_argreduce(&a, 0)); }, 20);
    printf("C3b argmax+max bf16 [8192,32000]         %8.2f us  %7.1f GB/s\n", ms * 1e3, 524369920.0 / ms / 1e6);
    if (!only_b) {
      b200phi_reduce_desc d = {};
      d.op = B200PHI_MAX; d.in_dtype = d.out_dtype = B200PHI_INT64; d.rank = 2; d.dims[0] = 8192; d.dims[1] = 32000;
      d.naxes = 1; d.axes[0] = 1; d.x = x; d.out = val;
      ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 20);
      printf("C3a max alone int64 [8192,32000]         %8.2f us  %7.1f GB/s\n", ms * 1e3, 2097217536.0 / ms / 1e6);
    }
    cudaFree(x); cudaFree(idx); cudaFree(val);
  }
  if (!*only || strstr(only, "C4")) {
    const size_t n = (size_t)16384 * 8192;
    void* x = dalloc(n * 2, 0x3f7f3f7f, 0x3c003c00);  // bf16 pairs of moderate magnitude
    void *m, *v; CK(cudaMalloc(&m, 16384 * 2)); CK(cudaMalloc(&v, 16384 * 2));
    b200phi_rowstats_desc r = {};
    r.in_dtype = r.out_dtype = B200PHI_BFLOAT16; r.rows = 16384; r.cols = 8192; r.unbiased = 1; r.x = x; r.mean = m; r.var = v;
    float ms = time_it([&] { OK(b200phi_rowstats(&r, 0)); }, 20);
    printf("C4 row mean+var bf16 [16384,8192]        %8.2f us  %7.1f GB/s\n", ms * 1e3, 268500992.0 / ms / 1e6);
    b200phi_reduce_desc d = {};
    d.op = B200PHI_MEAN; d.in_dtype = d.out_dtype = B200PHI_BFLOAT16; d.rank = 2; d.dims[0] = 16384; d.dims[1] = 8192;
    d.naxes = 1; d.axes[0] = 1; d.x = x; d.out = m;
    ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 20);
    printf("C4 mean alone bf16 [16384,8192]          %8.2f us  %7.1f GB/s\n", ms * 1e3, 268468224.0 / ms / 1e6);
    cudaFree(x); cudaFree(m); cudaFree(v);
  }
  if (!*only || strstr(only, "C5")) {
    const long long R = 524288, C = 4096;
    float* x = (float*)dalloc((size_t)R * C * 4, F01_MASK, F01_OR);
    float* out; CK(cudaMalloc(&out, C * 4));
    b200phi_reduce_desc d = {};
    d.op = B200PHI_SUM; d.in_dtype = d.out_dtype = B200PHI_FLOAT32; d.rank = 2; d.dims[0] = R; d.dims[1] = C;
    d.naxes = 1; d.axes[0] = 0; d.x = x; d.out = out;
    size_t ws = b200phi_reduce_workspace(&d);
    void* w; CK(cudaMalloc(&w, ws ? ws : 256));
    d.workspace = w; d.workspace_bytes = ws;
    float ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 10);
    printf("C5 sum fp32 [524288,4096] axis 0 (ws %zu B)  %8.2f us  %7.1f GB/s\n", ws, ms * 1e3, 8589950976.0 / ms / 1e6);
    d.naxes = 0; d.reduce_all = 1;
    ws = b200phi_reduce_workspace(&d); d.workspace_bytes = ws;
    ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 10);
    printf("   full reduce of the same 2^31 floats        %8.2f us  %7.1f GB/s\n", ms * 1e3, 8589934596.0 / ms / 1e6);
    d.naxes = 1; d.axes[0] = 1; d.reduce_all = 0; d.workspace_bytes = 0; d.workspace = nullptr;
    float* out2; CK(cudaMalloc(&out2, R * 4)); d.out = out2;
    ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 10);
    printf("   row sums of the same tensor (axis 1)       %8.2f us  %7.1f GB/s\n", ms * 1e3, (8589934592.0 + R * 4) / ms / 1e6);
    cudaFree(x); cudaFree(out); cudaFree(w); cudaFree(out2);
  }
  if (strstr(only, "M")) {  // functor plug-ins (b200phi_map); not part of the default run
    const size_t n = (size_t)256 * 64 * 112 * 112;
    float* x = (float*)dalloc(n * 4, 0x807fffff, 0x3f000000);
    float* y = (float*)dalloc(n * 4, 0x807fffff, 0x3f000000);
    float* out; CK(cudaMalloc(&out, n * 4));
    auto view = [](b200phi_tensor_view* v, const void* data, int rank, long long a, long long b, long long c, long long e) {
      v->rank = rank; v->dims[0] = a; v->dims[1] = b; v->dims[2] = c; v->dims[3] = e; v->data = data;
    };
    b200phi_map_desc m = {};
    m.in_dtype = B200PHI_FLOAT32; m.axis = -1; m.out_rank = 4;
    m.out_dims[0] = 256; m.out_dims[1] = 64; m.out_dims[2] = 112; m.out_dims[3] = 112; m.out = out;
    view(&m.in[0], x, 4, 256, 64, 112, 112);
    // unary
    struct { int op; const char* name; } un[] = {{B200PHI_MAP_RELU, "relu"}, {B200PHI_MAP_SILU, "silu"},
                                                  {B200PHI_MAP_EXP, "exp"}, {B200PHI_MAP_TANH, "tanh"}};
    for (auto& u : un) {
      m.op = u.op; m.arity = 1;
      float ms = time_it([&] { OK(b200phi_map(&m, 0)); }, 20);
      printf("M1 %-6s fp32 [256,64,112,112]              %8.2f us  %7.1f GB/s\n", u.name, ms * 1e3, 2.0 * n * 4 / ms / 1e6);
    }
    // compare, same shape and channel broadcast: 2 (or 1) fp32 reads, 1 byte written per element
    m.op = B200PHI_MAP_LESS_THAN; m.arity = 2;
    view(&m.in[1], y, 4, 256, 64, 112, 112);
    float ms = time_it([&] { OK(b200phi_map(&m, 0)); }, 20);
    printf("M2 less_than fp32 same shape -> bool          %8.2f us  %7.1f GB/s\n", ms * 1e3, 9.0 * n / ms / 1e6);
    view(&m.in[1], y, 3, 64, 1, 1, 0);
    ms = time_it([&] { OK(b200phi_map(&m, 0)); }, 20);
    printf("M2 less_than fp32 [..]<[64,1,1] -> bool       %8.2f us  %7.1f GB/s\n", ms * 1e3, 5.0 * n / ms / 1e6);
    m.op = B200PHI_MAP_EQUAL;
    ms = time_it([&] { OK(b200phi_map(&m, 0)); }, 20);
    printf("M2 equal fp32 [..]==[64,1,1] -> bool          %8.2f us  %7.1f GB/s\n", ms * 1e3, 5.0 * n / ms / 1e6);
    // max_grad over the last axis of [8192, 32000] bf16 / fp32: x streamed, out and dout per row
    for (int pass = 0; pass < 2; ++pass) {
      const int dt = pass ? B200PHI_FLOAT32 : B200PHI_BFLOAT16;
      const size_t es = pass ? 4 : 2;
      const size_t nn = (size_t)8192 * (pass ? 16000 : 32000);
      b200phi_map_desc g = {};
      g.op = B200PHI_MAP_EQUAL_MUL; g.in_dtype = dt; g.arity = 3; g.axis = -1; g.out_rank = 2;
      g.out_dims[0] = 8192; g.out_dims[1] = (long long)(nn / 8192); g.out = out;
      view(&g.in[0], x, 2, 8192, (long long)(nn / 8192), 0, 0);
      view(&g.in[1], y, 2, 8192, 1, 0, 0);
      view(&g.in[2], y + 65536, 2, 8192, 1, 0, 0);
      ms = time_it([&] { OK(b200phi_map(&g, 0)); }, 20);
      printf("M3 max_grad %s [8192,%zu] axis -1          %8.2f us  %7.1f GB/s\n", pass ? "fp32" : "bf16", nn / 8192,
             ms * 1e3, 2.0 * nn * es / ms / 1e6);
    }
    {  // max_grad over axis 0 of [32768, 4096] fp32: all three operands streamed along the row
      b200phi_map_desc g = {};
      g.op = B200PHI_MAP_EQUAL_MUL; g.in_dtype = B200PHI_FLOAT32; g.arity = 3; g.axis = -1; g.out_rank = 2;
      g.out_dims[0] = 32768; g.out_dims[1] = 4096; g.out = out;
      view(&g.in[0], x, 2, 32768, 4096, 0, 0);
      view(&g.in[1], y, 1, 4096, 0, 0, 0);
      view(&g.in[2], y + 4096, 1, 4096, 0, 0, 0);
      ms = time_it([&] { OK(b200phi_map(&g, 0)); }, 20);
      printf("M3 max_grad fp32 [32768,4096] axis 0          %8.2f us  %7.1f GB/s\n", ms * 1e3, 2.0 * 32768 * 4096 * 4 / ms / 1e6);
    }
    cudaFree(x); cudaFree(y); cudaFree(out);
  }
  if (strstr(only, "G")) {  // per-channel reductions: [N,C,H,W] over (0,2,3) -- batch-norm statistics, bias gradients
    const size_t n = (size_t)256 * 64 * 112 * 112;
    float* x = (float*)dalloc(n * 4, 0x807fffff, 0x3f000000);
    float* out; CK(cudaMalloc(&out, 4096 * 4));
    b200phi_reduce_desc d = {};
    d.op = B200PHI_SUM; d.in_dtype = d.out_dtype = B200PHI_FLOAT32; d.rank = 4;
    d.dims[0] = 256; d.dims[1] = 64; d.dims[2] = 112; d.dims[3] = 112;
    d.naxes = 3; d.axes[0] = 0; d.axes[1] = 2; d.axes[2] = 3; d.x = x; d.out = out;
    size_t ws = b200phi_reduce_workspace(&d);
    void* w; CK(cudaMalloc(&w, ws ? ws : 256)); d.workspace = w; d.workspace_bytes = ws;
    float ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 10);
    printf("G1 sum fp32 [256,64,112,112] over (0,2,3) (ws %zu) %8.2f us  %7.1f GB/s\n", ws, ms * 1e3, n * 4.0 / ms / 1e6);
    cudaFree(w);
    d.in_dtype = d.out_dtype = B200PHI_BFLOAT16; d.op = B200PHI_MEAN; d.rank = 3;
    d.dims[0] = 512; d.dims[1] = 1024; d.dims[2] = 784; d.naxes = 2; d.axes[0] = 0; d.axes[1] = 2;
    ws = b200phi_reduce_workspace(&d);
    CK(cudaMalloc(&w, ws ? ws : 256)); d.workspace = w; d.workspace_bytes = ws;
    ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 10);
    printf("G2 mean bf16 [512,1024,784] over (0,2) (ws %zu)   %8.2f us  %7.1f GB/s\n", ws, ms * 1e3, 512.0 * 1024 * 784 * 2 / ms / 1e6);
    cudaFree(x); cudaFree(out); cudaFree(w);
  }
  if (strstr(only, "A")) {  // mixed-precision add: fp32 accumulator += bf16 gradient, in place
    const size_t n = (size_t)1 << 28;
    float* acc = (float*)dalloc(n * 4, 0x807fffff, 0x3f000000);
    void* gr = dalloc(n * 2, 0xbf7fbf7f, 0x3c003c00);
    b200phi_mixed_add_desc d = {};
    d.y_dtype = B200PHI_BFLOAT16;
    d.x.rank = d.y.rank = d.out_rank = 2;
    d.x.dims[0] = d.y.dims[0] = d.out_dims[0] = 16384; d.x.dims[1] = d.y.dims[1] = d.out_dims[1] = 16384;
    d.x.data = acc; d.y.data = gr; d.out = acc;
    float ms = time_it([&] { OK(b200phi_mixed_add(&d, 0)); }, 20);
    printf("A1 fp32 += bf16 [16384,16384] in place        %8.2f us  %7.1f GB/s\n", ms * 1e3, n * 10.0 / ms / 1e6);
    cudaFree(acc); cudaFree(gr);
  }
  if (strstr(only, "L")) {  // reduce-side plug-ins: logsumexp rows and norm reductions
    const size_t n = (size_t)8192 * 32000;
    void* x = dalloc(n * 4, 0x40ff40ff, 0x3c003c00);  // as bf16: pairs of moderate magnitude; as fp32: same range
    void* out; CK(cudaMalloc(&out, 8192 * 8));
    b200phi_logsumexp_desc l = {};
    l.rows = 8192; l.cols = 32000; l.x = x; l.out = out;
    l.dtype = B200PHI_BFLOAT16;
    float ms = time_it([&] { OK(b200phi_logsumexp(&l, 0)); }, 20);
    printf("L1 logsumexp bf16 [8192,32000]             %8.2f us  %7.1f GB/s\n", ms * 1e3, (n * 2.0 + 8192 * 2) / ms / 1e6);
    l.dtype = B200PHI_FLOAT32;
    ms = time_it([&] { OK(b200phi_logsumexp(&l, 0)); }, 20);
    printf("L1 logsumexp fp32 [8192,32000]             %8.2f us  %7.1f GB/s\n", ms * 1e3, (n * 4.0 + 8192 * 4) / ms / 1e6);
    l.dtype = B200PHI_BFLOAT16; l.rows = 16384 * 4; l.cols = 4096;
    ms = time_it([&] { OK(b200phi_logsumexp(&l, 0)); }, 20);
    printf("L1 logsumexp bf16 [65536,4096]             %8.2f us  %7.1f GB/s\n", ms * 1e3, (65536.0 * 4096 * 2) / ms / 1e6);
    b200phi_reduce_desc d = {};
    d.op = B200PHI_L2_NORM; d.in_dtype = d.out_dtype = B200PHI_FLOAT32; d.rank = 2; d.dims[0] = 32768; d.dims[1] = 4096;
    d.naxes = 1; d.axes[0] = 1; d.x = x; d.out = out;
    ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 20);
    printf("L2 p_norm(2) fp32 [32768,4096] axis -1      %8.2f us  %7.1f GB/s\n", ms * 1e3, 537001984.0 / ms / 1e6);
    d.op = B200PHI_SUM_SQUARE; d.in_dtype = d.out_dtype = B200PHI_BFLOAT16; d.dims[0] = 65536; d.naxes = 0; d.reduce_all = 1;
    size_t ws = b200phi_reduce_workspace(&d);
    void* w; CK(cudaMalloc(&w, ws ? ws : 256)); d.workspace = w; d.workspace_bytes = ws;
    ms = time_it([&] { OK(b200phi_reduce(&d, 0)); }, 20);
    printf("L2 squared_l2_norm bf16 2^28 elements       %8.2f us  %7.1f GB/s\n", ms * 1e3, 65536.0 * 4096 * 2 / ms / 1e6);
    cudaFree(x); cudaFree(out); cudaFree(w);
  }
  printf("done\n");
  return 0;
}
