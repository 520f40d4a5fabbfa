// Achievable HBM bandwidth for the fused kernel's traffic pattern with no arithmetic: per point
// read 16 B (timestamp + value), write NOUT x 8 B, rows of 5760 points, one CTA per series, each
// thread moves 8 consecutive points per tile with 128-bit loads and 256-bit stores.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
__device__ __forceinline__ void st4(double* p, double a, double b, double c, double d) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
template <int NOUT>
__global__ void __launch_bounds__(128, 5) k(const int64_t* __restrict__ ts, const double* __restrict__ val, int64_t T,
                                          double* o0, double* o1, double* o2, double* o3, double* o4, double* o5) {
  const int64_t s = blockIdx.x;
  double* outs[6] = {o0, o1, o2, o3, o4, o5};
  for (int64_t j = threadIdx.x * 8; j < T; j += 128 * 8) {
    const int64_t i = s * T + j;
    double v[8]; long long t[8];
#pragma unroll
    for (int r = 0; r < 8; r += 2) {
      const longlong2 a = *reinterpret_cast<const longlong2*>(ts + i + r);
      const double2 b = *reinterpret_cast<const double2*>(val + i + r);
      t[r] = a.x; t[r + 1] = a.y; v[r] = b.x; v[r + 1] = b.y;
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) v[r] += (double)(t[r] & 1);
#pragma unroll
    for (int o = 0; o < NOUT; ++o) {
      st4(outs[o] + i, v[0], v[1], v[2], v[3]);
      st4(outs[o] + i + 4, v[4], v[5], v[6], v[7]);
    }
  }
}
template <int NOUT> void run(const int64_t* ts, const double* val, int64_t S, int64_t T, double** o) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; ++i) k<NOUT><<<(unsigned)S, 128>>>(ts, val, T, o[0], o[1], o[2], o[3], o[4], o[5]);
  CK(cudaEventRecord(a));
  for (int i = 0; i < 20; ++i) k<NOUT><<<(unsigned)S, 128>>>(ts, val, T, o[0], o[1], o[2], o[3], o[4], o[5]);
  CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
  float ms; CK(cudaEventElapsedTime(&ms, a, b)); ms /= 20;
  const double bytes = (double)S * T * (16 + 8 * NOUT);
  printf("read 16 B + write %d x 8 B per point: %.3f ms  %.0f GB/s\n", NOUT, ms, bytes / ms / 1e6);
}
int main() {
  const int64_t S = 6144, T = 5760, N = S * T;
  int64_t* ts; double* val; double* o[6];
  CK(cudaMalloc(&ts, N * 8)); CK(cudaMalloc(&val, N * 8));
  CK(cudaMemset(ts, 0, N * 8)); CK(cudaMemset(val, 0, N * 8));
  for (int i = 0; i < 6; ++i) { CK(cudaMalloc(&o[i], N * 8)); }
  run<1>(ts, val, S, T, o); run<2>(ts, val, S, T, o); run<4>(ts, val, S, T, o); run<6>(ts, val, S, T, o);
  return 0;
}
