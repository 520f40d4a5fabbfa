// hist_variants.cu -- how to add one 11-bit digit per value to a CTA-private shared-memory histogram
// on B200: plain shared atomics against warp-level folding, on digit patterns from spread to constant.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o hist_variants hist_variants.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int BINS = 2048;

__device__ __forceinline__ void add_rounds(unsigned* h, bool hit, unsigned d, int lane, int& spread) {
  unsigned todo = __ballot_sync(0xffffffffu, hit);
  if (spread < 4) {
#pragma unroll
    for (int round = 0; round < 2; ++round) {
      if (todo == 0) return;
      const int leader = __ffs(todo) - 1;
      const unsigned dl = __shfl_sync(0xffffffffu, d, leader);
      const unsigned peers = __ballot_sync(0xffffffffu, hit && d == dl) & todo;
      if (lane == leader) atomicAdd(&h[dl], (unsigned)__popc(peers));
      if (round == 0) spread = (__popc(peers) <= 2 && __popc(todo) >= 16) ? spread + 1 : 0;
      todo &= ~peers;
    }
  }
  if ((todo >> lane) & 1) atomicAdd(&h[d], 1u);
}
__device__ __forceinline__ void add_plain(unsigned* h, bool hit, unsigned d) { if (hit) atomicAdd(&h[d], 1u); }
__device__ __forceinline__ void add_match(unsigned* h, bool hit, unsigned d, int lane) {
  const unsigned peers = __match_any_sync(0xffffffffu, hit ? d : 0xffffffffu);
  if (hit && lane == __ffs(peers) - 1) atomicAdd(&h[d], (unsigned)__popc(peers));
}
// one round: if the whole warp agrees one lane adds for all, else plain
__device__ __forceinline__ void add_uniform(unsigned* h, bool hit, unsigned d, int lane) {
  const unsigned todo = __ballot_sync(0xffffffffu, hit);
  if (todo == 0) return;
  const int leader = __ffs(todo) - 1;
  const unsigned dl = __shfl_sync(0xffffffffu, d, leader);
  const unsigned peers = __ballot_sync(0xffffffffu, hit && d == dl);
  if (peers == todo) { if (lane == leader) atomicAdd(&h[dl], (unsigned)__popc(peers)); }
  else if (hit) atomicAdd(&h[d], 1u);
}

template <int V>
__global__ void __launch_bounds__(256) hist(const double* __restrict__ x, int64_t n, int64_t per_cta, unsigned long long* __restrict__ out) {
  __shared__ unsigned sh[BINS];
  for (int i = threadIdx.x; i < BINS; i += 256) sh[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t a = (int64_t)blockIdx.x * per_cta, b = a + per_cta < n ? a + per_cta : n;
  int spread = 0;
  if (V == 3) {
    // four consecutive values per lane, equal neighbours merged before the atomics
    for (int64_t j = a + 4 * threadIdx.x; j + 3 < b; j += 1024) {
      const double2 p = *reinterpret_cast<const double2*>(x + j), q = *reinterpret_cast<const double2*>(x + j + 2);
      const double v[4] = {p.x, p.y, q.x, q.y};
      unsigned d[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) d[i] = (unsigned)((unsigned long long)__double_as_longlong(v[i]) >> 41) & (BINS - 1);
      unsigned c = 1;
#pragma unroll
      for (int i = 1; i < 4; ++i) {
        if (d[i] == d[i - 1]) ++c; else { atomicAdd(&sh[d[i - 1]], c); c = 1; }
      }
      atomicAdd(&sh[d[3]], c);
    }
  } else {
    for (int64_t j = a + 2 * threadIdx.x; j + 513 < b; j += 1024) {
      const double2 p = *reinterpret_cast<const double2*>(x + j), q = *reinterpret_cast<const double2*>(x + j + 512);
      const double v[4] = {p.x, p.y, q.x, q.y};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const unsigned d = (unsigned)((unsigned long long)__double_as_longlong(v[i]) >> 41) & (BINS - 1);
        const bool hit = v[i] == v[i];
        if (V == 0) add_rounds(sh, hit, d, lane, spread);
        if (V == 1) add_plain(sh, hit, d);
        if (V == 2) add_match(sh, hit, d, lane);
        if (V == 4) add_uniform(sh, hit, d, lane);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < BINS; i += 256) { const unsigned c = sh[i]; if (c) atomicAdd(&out[i], (unsigned long long)c); }
}

int main() {
  const int64_t n = 6144LL * 5760, per_cta = 17280;   // 2048 CTAs
  std::vector<double> h(n);
  double* d; unsigned long long* o;
  cudaMalloc(&d, n * 8); cudaMalloc(&o, BINS * 8);
  const char* names[] = {"uniform digits", "ramp (runs of ~64 equal digits)", "64 distinct values", "constant", "slow ramp (runs of ~2000)"};
  for (int pat = 0; pat < 5; ++pat) {
    uint64_t s = 88172645463325252ULL;
    for (int64_t i = 0; i < n; ++i) {
      s ^= s << 13; s ^= s >> 7; s ^= s << 17;
      double v;
      if (pat == 0) v = 1.0 + (double)(s >> 11) / 9007199254740992.0;
      else if (pat == 1) v = 1.0 + (double)(i % 131072) / 131072.0;
      else if (pat == 2) v = 1.0 + (double)((s >> 20) & 63) / 64.0;
      else if (pat == 3) v = 1.5;
      else v = 1.0 + (double)(i % 4194304) / 4194304.0;
      h[i] = v;
    }
    cudaMemcpy(d, h.data(), n * 8, cudaMemcpyHostToDevice);
    printf("%s\n", names[pat]);
    for (int V = 0; V < 5; ++V) {
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      float best = 1e9f;
      for (int rep = 0; rep < 4; ++rep) {
        cudaMemset(o, 0, BINS * 8);
        cudaEventRecord(e0);
        const unsigned grid = (unsigned)((n + per_cta - 1) / per_cta);
        if (V == 0) hist<0><<<grid, 256>>>(d, n, per_cta, o);
        if (V == 1) hist<1><<<grid, 256>>>(d, n, per_cta, o);
        if (V == 2) hist<2><<<grid, 256>>>(d, n, per_cta, o);
        if (V == 3) hist<3><<<grid, 256>>>(d, n, per_cta, o);
        if (V == 4) hist<4><<<grid, 256>>>(d, n, per_cta, o);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
      }
      unsigned long long tot = 0; std::vector<unsigned long long> ho(BINS);
      cudaMemcpy(ho.data(), o, BINS * 8, cudaMemcpyDeviceToHost);
      for (auto c : ho) tot += c;
      const char* vn[] = {"two leader rounds + spread heuristic (product)", "plain shared atomics", "match_any", "4 consecutive per lane, runs merged", "uniform-warp check, else plain"};
      printf("  %-48s %8.1f us  (%.0f GB/s)  counted %llu\n", vn[V], best * 1000, n * 8 / (best * 1e6), tot);
    }
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return 0;
}
