// Micro-benchmark: L1TEX data-pipe wavefronts per access pattern (run under ncu --metrics ...).
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

// each warp covers `span` bytes per iteration; lane l touches `bytes` at l*stride
template <int MODE> __global__ void k(double* __restrict__ p, long n_iter, long warp_span_doubles) {
  const int lane = threadIdx.x & 31;
  const long warp = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
  double* base = p + warp * n_iter * warp_span_doubles;
  double acc = 0;
  for (long i = 0; i < n_iter; ++i, base += warp_span_doubles) {
    if (MODE == 0) {        // STG.256, lane stride 64 B, two instructions fill 2 KB
      double* q = base + lane * 8;
      asm volatile("st.global.v4.f64 [%0], {%1,%1,%1,%1};" ::"l"(q), "d"(1.0) : "memory");
      asm volatile("st.global.v4.f64 [%0], {%1,%1,%1,%1};" ::"l"(q + 4), "d"(2.0) : "memory");
    } else if (MODE == 1) { // STG.256, lane stride 32 B (contiguous), two instructions fill 2 KB
      double* q = base + lane * 4;
      asm volatile("st.global.v4.f64 [%0], {%1,%1,%1,%1};" ::"l"(q), "d"(1.0) : "memory");
      asm volatile("st.global.v4.f64 [%0], {%1,%1,%1,%1};" ::"l"(q + 128), "d"(2.0) : "memory");
    } else if (MODE == 2) { // STG.128 contiguous, four instructions fill 2 KB
      double* q = base + lane * 2;
      for (int j = 0; j < 4; ++j) asm volatile("st.global.v2.f64 [%0], {%1,%1};" ::"l"(q + 64 * j), "d"(1.0) : "memory");
    } else if (MODE == 3) { // LDG.128 x4, lane stride 64 B (2 KB)
      const double* q = base + lane * 8;
      for (int j = 0; j < 4; ++j) { double a, b; asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(a), "=d"(b) : "l"(q + 2 * j)); acc += a + b; }
    } else if (MODE == 4) { // LDG.256 x2, lane stride 64 B (2 KB)
      const double* q = base + lane * 8;
      for (int j = 0; j < 2; ++j) { double a, b, c, d; asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(q + 4 * j)); acc += a + b + c + d; }
    } else if (MODE == 5) { // LDG.128 x4 contiguous (2 KB)
      const double* q = base + lane * 2;
      for (int j = 0; j < 4; ++j) { double a, b; asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(a), "=d"(b) : "l"(q + 64 * j)); acc += a + b; }
    } else if (MODE == 6) { // LDG.256 x2 contiguous (2 KB)
      const double* q = base + lane * 4;
      for (int j = 0; j < 2; ++j) { double a, b, c, d; asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(q + 128 * j)); acc += a + b + c + d; }
    } else if (MODE == 7) { // STG.64 x8 transposed-coalesced (each instr 256 B contiguous), 2 KB
      double* q = base + lane;
      for (int j = 0; j < 8; ++j) asm volatile("st.global.f64 [%0], %1;" ::"l"(q + 32 * j), "d"(1.0) : "memory");
    }
  }
  if (acc == 123.456) p[0] = acc;
}
template <int MODE> int run(double* d, const char* name) {
  const long n_iter = 64, span = 256;             // 2 KB per warp-iteration
  const int blocks = 148 * 8, threads = 128;       // 4736 warps x 64 x 2 KB = 620 MB
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<blocks, threads>>>(d, n_iter, span);
  cudaEventRecord(a);
  k<MODE><<<blocks, threads>>>(d, n_iter, span);
  cudaEventRecord(b);
  CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double bytes = (double)blocks * threads / 32 * n_iter * 2048;
  printf("mode %d %-44s %.3f ms  %.0f GB/s\n", MODE, name, ms, bytes / ms / 1e6);
  return 0;
}
int main() {
  double* d; CK(cudaMalloc(&d, (size_t)148 * 8 * 4 * 64 * 2048 + 4096));
  CK(cudaMemset(d, 0, (size_t)148 * 8 * 4 * 64 * 2048));
  run<0>(d, "STG.256 lane stride 64B"); run<1>(d, "STG.256 contiguous"); run<2>(d, "STG.128 contiguous");
  run<7>(d, "STG.64 contiguous");
  run<3>(d, "LDG.128 lane stride 64B"); run<4>(d, "LDG.256 lane stride 64B"); run<5>(d, "LDG.128 contiguous"); run<6>(d, "LDG.256 contiguous");
  return 0;
}
