// window_fast_impl.cuh -- the roofline kernel: fused align + rolling min/max/mean + increase/rate.
// (Included by window_fast.cu and window_fast_dense.cu, which instantiate it for sample input and
// for dense input; two files so that they build in parallel.)
//
// Same definitions and the same block decomposition as window.cu (suffix of the first chunk +
// whole 8-point chunks + prefix of the own chunk; nothing is ever subtracted), re-cut so that a
// grid point costs ~150 thread instructions instead of ~840:
//   * results leave the SM straight from registers as 256-bit stores (STG.E.ENL2.256: one full
//     32-byte sector per lane); the tile origin is shifted by 0..3 halo points so that every
//     thread's 8 points start on a 32-byte boundary of every output row;
//   * the increase is accumulated as a sum of per-sample deltas (x - prev, or x itself after a
//     drop; the telescoped form of (last - first) + pre-reset values, equal within a few ulp of
//     the increments), so a window's increase is suffix + chunk totals + prefix like the sum, and
//     no first/last value has to be fetched.  Junction deltas between chunks are formed on the
//     fly while the chunk totals are folded; the fold also keeps the variant that starts inside the
//     folded chunks, which is what a window needs whose first chunk part is empty (NaN runs and
//     the start of every series);
//   * divisions are replaced by two per-CTA tables in shared memory (1/n and 1000/(k*step));
//   * the align stage starts from "one sample per step" as its guess and only binary-searches the
//     staged timestamps when the guess is wrong; indices are 32-bit, shared memory is addressed
//     directly (no generic pointers).
// Requests it does not cover (window_steps < 9 or > NE/2 with window outputs, n_resets on sample
// input, extrapolated rates, source indices together with window outputs, output rows not 32-byte
// aligned) go to the general kernel in window.cu.
//
// Shared memory per CTA (NT threads, NE = 8 NT points): four suffix arrays S_*[NE] (aliased with
// the sample staging area and the aligned-value hand-off), chunk totals T_*[NT], the two tables.
#include <math.h>
#include <stdlib.h>
#include "window_common.cuh"

#ifndef PMX_FAST_CTAS_128
#define PMX_FAST_CTAS_128 5
#endif

namespace pmxw {
namespace {

template <int NT> struct FTile {
  static constexpr int NE = NT * E;
  static constexpr int NS = NE + E;
  static constexpr int NSP = NS + NS / 8 + 2;
  static constexpr size_t S_BYTES = size_t(4) * NE * 8;          // sum, min, max, delta-sum
  static constexpr size_t SM_BYTES = size_t(NE) * 2;
  static constexpr size_t T_BYTES = size_t(6) * NT * 8 + size_t(NT) * 4;
  static constexpr size_t FIXED = S_BYTES + SM_BYTES + T_BYTES + 16;
  // staging (values + 32-bit times) sits in S_sum/S_mn, the aligned values in S_d meanwhile
  static_assert(size_t(NSP) * 12 <= size_t(2) * NE * 8, "staging must fit in two suffix arrays");
};

__device__ __forceinline__ int pad8(int i) { return i + (i >> 3); }
__device__ __forceinline__ double dmin(double a, double b) { return b < a ? b : a; }
__device__ __forceinline__ double dmax(double a, double b) { return b > a ? b : a; }

__device__ __forceinline__ void st4(double* p, double a, double b, double c, double d) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
// tile edges: the group of four straddles the output range [e_lo, e_hi).  Cold.
__device__ __noinline__ void store4_edge(double* __restrict__ row, int e, int e_lo, int e_hi,
                                         double a, double b, double c, double d) {
  if (e >= e_lo && e < e_hi) row[e] = a;
  if (e + 1 >= e_lo && e + 1 < e_hi) row[e + 1] = b;
  if (e + 2 >= e_lo && e + 2 < e_hi) row[e + 2] = c;
  if (e + 3 >= e_lo && e + 3 < e_hi) row[e + 3] = d;
}
// four consecutive outputs of one array; e is a multiple of 4; [e_lo, e_hi) is the output range
__device__ __forceinline__ void store4(double* __restrict__ row, int e, int e_lo, int e_hi,
                                       double a, double b, double c, double d) {
  if (row == nullptr) return;
  if (e >= e_lo && e + 3 < e_hi) st4(row + e, a, b, c, d);
  else if (e + 3 >= e_lo && e < e_hi) store4_edge(row, e, e_lo, e_hi, a, b, c, d);
}
__device__ __forceinline__ void store4i(int32_t* __restrict__ row, int e, int e_lo, int e_hi,
                                        int a, int b, int c, int d) {
  if (row == nullptr) return;
  if (e >= e_lo && e + 3 < e_hi) { *reinterpret_cast<int4*>(row + e) = make_int4(a, b, c, d); return; }
  if (e >= e_lo && e < e_hi) row[e] = a;
  if (e + 1 >= e_lo && e + 1 < e_hi) row[e + 1] = b;
  if (e + 2 >= e_lo && e + 2 < e_hi) row[e + 2] = c;
  if (e + 3 >= e_lo && e + 3 < e_hi) row[e + 3] = d;
}

// largest index in [lo, n) of the padded staged (tile-relative, 32-bit) timestamps with ts <= g,
// or lo-1.  Only runs when the one-sample-per-step guess is wrong.
__device__ __noinline__ int floor_staged(const int* s_t, int lo, int n, int g) {
  int hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (s_t[pad8(mid)] <= g) lo = mid + 1; else hi = mid;
  }
  return lo - 1;
}

__device__ int64_t floor_global(const int64_t* __restrict__ ts, int64_t n, int64_t g) {
  if (n == 0) return -1;
  const int64_t t0 = ts[0];
  if (g < t0) return -1;
  const int64_t tl = ts[n - 1];
  if (g >= tl) return n - 1;
  int64_t guess = (int64_t)((double)(g - t0) * (double)(n - 1) / (double)(tl - t0));
  if (guess < 0) guess = 0;
  if (guess > n - 2) guess = n - 2;
  if (ts[guess] <= g && g < ts[guess + 1]) return guess;
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (ts[mid] <= g) lo = mid + 1; else hi = mid;
  }
  return lo - 1;
}

// One grid point of a tile whose samples do not fit the staging area (many samples per grid
// step): binary search of the series in global memory.  Cold.
__device__ __noinline__ double align_one_global(const int64_t* __restrict__ ts, const double* __restrict__ val,
                                                int64_t cur, int64_t len, int64_t g, int64_t lookback,
                                                int64_t* idx_out, int32_t* src_out) {
  int64_t lo = cur, hi = len;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (ts[mid] <= g) lo = mid + 1; else hi = mid; }
  const int64_t idx = lo - 1;
  *idx_out = idx;
  *src_out = -1;
  if (idx < 0) return pmx_nan();
  const double sample = val[idx];
  if (g - ts[idx] < lookback && (unsigned long long)__double_as_longlong(sample) != PMX_STALE_NAN_BITS) {
    *src_out = (int32_t)idx;
    return sample;
  }
  return pmx_nan();
}

__device__ __forceinline__ double delta_of(double cur, double prev) { return cur < prev ? cur : cur - prev; }

// General tile (gaps, several samples per grid step, series ends): stage the samples that can
// serve the tile in shared memory as 32-bit tile-relative timestamps + values, then guess / verify /
// binary-search per thread and walk forward.  Out of line: regular tiles never come here, and the
// hot loop stays small.  All threads of the CTA call it together (it synchronises).  Writes the
// aligned values to A and returns the cursor for the next tile.
template <int NT>
__device__ __noinline__ int64_t align_general_tile(const int64_t* __restrict__ gts0, const double* __restrict__ gv0,
                                                   int64_t len, int64_t cur, int shift, int elo, int64_t tbase,
                                                   int64_t step_ms, int64_t lookback_ms, int n_el, int OUT, int e0,
                                                   int t, double* s_val, int* s_t, double* A, int64_t* s_cursor,
                                                   int32_t* IA) {
  using TL = FTile<NT>;
  const int step = (int)step_ms, lookback = (int)lookback_ms;
  const int64_t rem = len - cur;
  const int ns = (int)(rem < TL::NS ? rem : TL::NS);
  {
    const int64_t* gts = gts0 + cur;
    const double* gv = gv0 + cur;
    for (int i = t; i < ns; i += NT) {
      int d32;                                                    // saturate: order and lookback tests stay exact
      asm("cvt.sat.s32.s64 %0, %1;" : "=r"(d32) : "l"(gts[i] - tbase));
      s_t[pad8(i)] = d32 < -0x40000000 ? -0x40000000 : d32;
      s_val[pad8(i)] = gv[i];
    }
  }
  __syncthreads();
  const int ghi = (n_el - 1) * step;
  const bool covered = (cur + ns >= len) || (ns > 0 && s_t[pad8(ns - 1)] > ghi);
  const int e_cur = OUT;                                          // its floor seeds the next tile
  if (!covered) {
#pragma unroll 1
    for (int r = 0; r < E; ++r) {
      const int e = e0 + r;
      double v = pmx_nan();
      int32_t src = -1;
      if (e >= elo && e < n_el) {
        int64_t idx;
        v = align_one_global(gts0, gv0, cur, len, tbase + (int64_t)e * step_ms, lookback_ms, &idx, &src);
        if (e == e_cur) *s_cursor = idx;
      }
      A[r * NT + t] = v;
      if (IA) IA[r * NT + t] = src;
    }
  } else {
    // guess: one sample per grid step; staged sample 0 is the floor of element elo (or the sample
    // after it when the series starts later)
    int idx = e0 - elo - shift;
    if (idx > ns - 1) idx = ns - 1;
    {
      const int ef = e0 > elo ? e0 : elo;                         // first element this thread aligns
      const int g = ef * step;
      if (idx < 0) idx = (ns > 0 && s_t[0] <= g) ? floor_staged(s_t, 0, ns, g) : -1;
      else if (!(s_t[pad8(idx)] <= g && (idx + 1 >= ns || s_t[pad8(idx + 1)] > g))) idx = floor_staged(s_t, 0, ns, g);
    }
#pragma unroll 1
    for (int r = 0; r < E; ++r) {
      const int e = e0 + r;
      const int g = e * step;
      double v = pmx_nan();
      int32_t src = -1;
      if (e >= elo && e < n_el) {
        if (idx + 1 < ns && s_t[pad8(idx + 1)] <= g) {
          ++idx;
          if (idx + 1 < ns && s_t[pad8(idx + 1)] <= g) idx = floor_staged(s_t, idx + 1, ns, g);
        }
        if (idx >= 0) {
          const double sample = s_val[pad8(idx)];
          if (g - s_t[pad8(idx)] < lookback &&
              (unsigned long long)__double_as_longlong(sample) != PMX_STALE_NAN_BITS) {
            v = sample;
            src = (int32_t)(cur + idx);
          }
        }
        if (e == e_cur) *s_cursor = idx >= 0 ? cur + idx : -1;
      }
      A[r * NT + t] = v;
      if (IA) IA[r * NT + t] = src;
    }
  }
  __syncthreads();                                                // staging dead, cursor visible
  return *s_cursor;
}

// QB = (window_steps - 1) & 7 as a compile-time constant (the output loops then unroll with every
// "far / near" choice and every shared-memory row offset folded), or -1 for any window.
// IDX: also write the source index of every aligned point (pmx_align; no window outputs then).
// NRES: also count the counter resets of every window (pmx_rate's n_resets): wherever a delta is
// formed, "current < previous" is counted beside it -- in-chunk counts ride in spare bits of the
// meta words (S_meta 11..13, T_meta 16..18), the folds keep integer twins of their sums.
template <int NT, bool ALIGN, int QB, bool IDX = false, bool NRES = false>
__global__ void __launch_bounds__(NT, (NT == 256 ? 2 : PMX_FAST_CTAS_128))
window_fast_kernel(const WinParams p) {
  using TL = FTile<NT>;
  constexpr int NE = TL::NE;
  extern __shared__ __align__(16) unsigned char smem[];
  double* S_sum = reinterpret_cast<double*>(smem);
  double* S_mn = S_sum + NE;
  double* S_mx = S_mn + NE;
  double* S_d = S_mx + NE;
  double* s_val = S_sum;                                          // staging aliases S_sum/S_mn
  int* s_t = reinterpret_cast<int*>(S_sum + TL::NSP);             // tile-relative timestamps
  double* A = S_d;                                                // aligned values, align -> registers
  uint16_t* S_meta = reinterpret_cast<uint16_t*>(S_d + NE);
  double* T_sum = reinterpret_cast<double*>(S_meta + NE);
  double* T_mn = T_sum + NT;
  double* T_mx = T_mn + NT;
  double* T_fv = T_mx + NT;
  double* T_lv = T_fv + NT;
  double* T_di = T_lv + NT;
  uint32_t* T_meta = reinterpret_cast<uint32_t*>(T_di + NT);
  int64_t* s_cursor = reinterpret_cast<int64_t*>(T_meta + NT);
  double* rcp = reinterpret_cast<double*>(s_cursor + 2);           // rcp[n] = 1/n, n <= w
  double* rdt = rcp + (p.w + 1);                                   // rdt[k] = 1000/(k*step_ms), k < w

  const int t = threadIdx.x;
  const int64_t s = blockIdx.x / p.chunks;
  const int ci = blockIdx.x % p.chunks;
  const int64_t c0 = (int64_t)ci * p.chunk_len;
  const int64_t c1 = (c0 + p.chunk_len < p.n_steps) ? c0 + p.chunk_len : p.n_steps;
  if (c0 >= c1) return;
  const int H = p.w - 1;
  const int qa = H >> 3, qb = QB >= 0 ? QB : (H & 7);
  // shift the tile origin back so that tile element 0 sits on a 32-byte boundary of the row
  const int HP = H + (int)(((s * p.n_steps + c0 - H) % 4 + 4) % 4);
  const int OUT = (NE - HP) & ~3;                                 // >= the host's (NE - H - 3) & ~3: a full tile uses all NE points
  const bool do_roll = p.o_min || p.o_max || p.o_mean || p.o_cnt;
  const bool do_rate = p.o_inc || p.o_rate || (NRES && p.o_nres);
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const int need = p.min_periods > 1 ? p.min_periods : 1;

  for (int k = t; k <= p.w; k += NT) {
    rcp[k] = k ? 1.0 / (double)k : 0.0;
    rdt[k] = k ? 1000.0 / (double)((int64_t)k * p.step_ms) : 0.0;
  }

  int64_t base = 0, len = 0, cursor = -1;
  if (ALIGN) { base = p.off[s]; len = p.off[s + 1] - base; }
  // ts and val must share the 16-byte phase for the paired loads
  const bool vec_in = ALIGN && ((reinterpret_cast<uintptr_t>(p.ts) | reinterpret_cast<uintptr_t>(p.val)) & 15) == 0;
  bool first = true;
  const int e0 = t * E;

  for (int64_t jt = c0; jt < c1; jt += OUT) {
    const int64_t j0 = jt - HP;                                   // grid index of tile element 0
    const int n_out = (int)((c1 - jt < OUT) ? (c1 - jt) : OUT);
    const int n_el = HP + n_out;
    const int64_t row = s * p.n_steps + j0;                       // multiple of 4

    double a[E];
    int32_t src[IDX ? E : 1];                                     // series-relative sample index or -1
    if (ALIGN) {
      const int elo = j0 < 0 ? (int)(-j0) : 0;                    // first element with a grid point
      const int64_t tbase = p.start_ms + j0 * p.step_ms;          // grid time of tile element 0
      if (first) {
        cursor = floor_global(p.ts + base, len, tbase + (int64_t)elo * p.step_ms);
        first = false;
      }
      // cursor = index of the last sample at or before the tile's first grid point, or -1
      const int64_t cur = cursor > 0 ? cursor : 0;
      const int shift = cursor < 0 ? 1 : 0;
      // ---- regular tile: exactly one sample per grid step.  Every thread fetches "its" nine
      // timestamps and eight values straight from global memory (full sectors, no staging) and
      // checks ts[k] <= g < ts[k+1] for each of its points; one barrier votes.
      bool ok = true;
      {
        const int64_t k0 = cur + (e0 - elo - shift);
        const int64_t* gts = p.ts + base;
        const double* gv = p.val + base;
        int64_t tk[E + 1]; double vk[E];
        // 128-bit loads when all nine samples exist; the pairing follows the parity of the first
        // sample's address (the same for every thread of the CTA: e0 is a multiple of 8).
        // (256-bit loads of three aligned groups + selection by phase were slower: 48 more live
        // registers, spills; 0.703 ms against 0.659 ms.)
        const bool inside = k0 >= 0 && k0 + E < len;
        const bool odd = ((reinterpret_cast<uintptr_t>(gts + k0) >> 3) & 1) != 0;
        if (inside && vec_in && !odd) {
#pragma unroll
          for (int r = 0; r < E; r += 2) {
            const longlong2 a2 = *reinterpret_cast<const longlong2*>(gts + k0 + r);
            const double2 v2 = *reinterpret_cast<const double2*>(gv + k0 + r);
            tk[r] = a2.x; tk[r + 1] = a2.y; vk[r] = v2.x; vk[r + 1] = v2.y;
          }
          tk[E] = gts[k0 + E];
        } else if (inside && vec_in) {
          tk[0] = gts[k0]; vk[0] = gv[k0];
#pragma unroll
          for (int r = 1; r < E; r += 2) {
            const longlong2 a2 = *reinterpret_cast<const longlong2*>(gts + k0 + r);
            tk[r] = a2.x; tk[r + 1] = a2.y;
          }
#pragma unroll
          for (int r = 1; r + 1 < E; r += 2) {
            const double2 v2 = *reinterpret_cast<const double2*>(gv + k0 + r);
            vk[r] = v2.x; vk[r + 1] = v2.y;
          }
          vk[E - 1] = gv[k0 + E - 1];
        } else {
#pragma unroll
          for (int r = 0; r <= E; ++r) {
            const int64_t k = k0 + r;
            tk[r] = (k >= 0 && k < len) ? gts[k] : (k < 0 ? (int64_t)0x8000000000000000LL : 0x7fffffffffffffffLL);
          }
#pragma unroll
          for (int r = 0; r < E; ++r) {
            const int64_t k = k0 + r;
            vk[r] = (k >= 0 && k < len) ? gv[k] : pmx_nan();
          }
        }
        int64_t g = tbase + (int64_t)e0 * p.step_ms;
        if (inside && e0 >= elo && e0 + E <= n_el) {
          // all nine samples exist and all eight points count: with u_r = g_r - ts[k0+r] the tile is
          // regular iff 0 <= u_r and u_{r+1} < step for every r.  The high words are OR-ed (any bit
          // set: negative or a gap beyond 2^32 ms, both go to the general path) and the low words
          // are compared as 32-bit numbers (step and lookback are < 2^30 here).
          uint32_t hor = 0, lo[E + 1];
#pragma unroll
          for (int r = 0; r <= E; ++r, g += p.step_ms) {
            const int64_t u = g - tk[r];
            lo[r] = (uint32_t)u;
            hor |= (uint32_t)((uint64_t)u >> 32);
          }
          const uint32_t st32 = (uint32_t)p.step_ms, lb32 = (uint32_t)p.lookback_ms;
#pragma unroll
          for (int r = 0; r < E; ++r) {
            ok = ok && lo[r + 1] < st32;
            const bool valid = lo[r] < lb32 && (unsigned long long)__double_as_longlong(vk[r]) != PMX_STALE_NAN_BITS;
            a[r] = valid ? vk[r] : pmx_nan();
            if (IDX) src[r] = valid ? (int32_t)(k0 + r) : -1;
          }
          ok = ok && hor == 0;
        } else {
#pragma unroll
          for (int r = 0; r < E; ++r, g += p.step_ms) {
            const int e = e0 + r;
            double v = pmx_nan();
            if (IDX) src[r] = -1;
            if (e >= elo && e < n_el) {
              ok = ok && tk[r] <= g && tk[r + 1] > g;
              if (k0 + r >= 0 && g - tk[r] < p.lookback_ms &&
                  (unsigned long long)__double_as_longlong(vk[r]) != PMX_STALE_NAN_BITS) {
                v = vk[r];
                if (IDX) src[r] = (int32_t)(k0 + r);
              }
            }
            a[r] = v;
          }
        }
      }
      if (__syncthreads_and(ok)) {
        cursor = cur + (OUT - elo - shift);
      } else {
        cursor = align_general_tile<NT>(p.ts + base, p.val + base, len, cur, shift, elo, tbase, p.step_ms,
                                        p.lookback_ms, n_el, OUT, e0, t, s_val, s_t, A, s_cursor,
                                        IDX ? reinterpret_cast<int32_t*>(S_mx) : nullptr);
#pragma unroll
        for (int r = 0; r < E; ++r) a[r] = A[r * NT + t];         // own elements: written by this thread
        if (IDX) {
#pragma unroll
          for (int r = 0; r < E; ++r) src[r] = reinterpret_cast<int32_t*>(S_mx)[r * NT + t];
        }
      }
      if (jt + OUT < c1) {                                        // pull the next tile's samples towards L2
        const int64_t ncur = cursor > 0 ? cursor : 0;
        const int64_t nrem = len - ncur;
        const int lines = (int)(((nrem < TL::NS ? nrem : TL::NS) * 8 + 127) >> 7);
        if (t < lines) {
          asm volatile("prefetch.global.L2 [%0];" ::"l"(p.ts + base + ncur + t * 16));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(p.val + base + ncur + t * 16));
        }
      }
    } else {
      // dense input: every thread reads its own eight points (256-bit loads when the row allows)
      const double* xr = p.x + row;
      if (e0 + E <= n_el && j0 + e0 >= 0 && (reinterpret_cast<uintptr_t>(xr) & 31) == 0) {
        asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a[0]), "=d"(a[1]), "=d"(a[2]), "=d"(a[3]) : "l"(xr + e0));
        asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a[4]), "=d"(a[5]), "=d"(a[6]), "=d"(a[7]) : "l"(xr + e0 + 4));
      } else {
#pragma unroll
        for (int r = 0; r < E; ++r) a[r] = (e0 + r < n_el && j0 + e0 + r >= 0) ? xr[e0 + r] : pmx_nan();
      }
    }

    if (p.o_aligned) {
      store4(p.o_aligned + row, e0, HP, n_el, a[0], a[1], a[2], a[3]);
      store4(p.o_aligned + row, e0 + 4, HP, n_el, a[4], a[5], a[6], a[7]);
    }
    if (IDX && p.o_idx) {
      store4i(p.o_idx + row, e0, HP, n_el, src[0], src[1], src[2], src[3]);
      store4i(p.o_idx + row, e0 + 4, HP, n_el, src[4], src[5], src[6], src[7]);
    }
    if (!do_roll && !do_rate) continue;

    // ---------------------------------------------------------------- in-chunk suffix scans
    // First as if all eight values were finite (no branches, constant counts); the chunk total
    // tells whether that was true (any NaN or Inf makes the sum non-finite) and only then the
    // general scan runs and overwrites what this thread stored.
    double c_fv = a[0];                                           // chunk's first valid value
    int c_cn = E;
    {
      double ss = a[E - 1], smn = a[E - 1], smx = a[E - 1], sd = 0.0;
      int snr = 0;                                                // resets inside the suffix (NRES)
      {
        const int si = (E - 1) * NT + t;
        S_sum[si] = ss; S_mn[si] = smn; S_mx[si] = smx; S_d[si] = sd;
        S_meta[si] = (uint16_t)(1 | (1 << 4));
      }
#pragma unroll
      for (int r = E - 2; r >= 0; --r) {
        const double v = a[r];
        ss += v; smn = dmin(smn, v); smx = dmax(smx, v);
        sd += delta_of(a[r + 1], v);
        if (NRES) snr += a[r + 1] < v ? 1 : 0;
        const int si = r * NT + t;
        S_sum[si] = ss; S_mn[si] = smn; S_mx[si] = smx; S_d[si] = sd;
        S_meta[si] = (uint16_t)((E - r) | ((E - r) << 4) | (NRES ? snr << 11 : 0));
      }
      T_sum[t] = ss; T_mn[t] = smn; T_mx[t] = smx; T_fv[t] = a[0]; T_lv[t] = a[E - 1]; T_di[t] = sd;
      T_meta[t] = (uint32_t)(E | (E << 4) | ((E - 1) << 12) | (NRES ? snr << 16 : 0));
      if (!pmx_isfinite(ss)) {
        ss = 0.0; smn = INF; smx = -INF; sd = 0.0; snr = 0;
        double sfv = 0.0, lv = 0.0;
        int scf = 0, scn = 0, foff = 0, lpos = 0;
#pragma unroll
        for (int r = E - 1; r >= 0; --r) {
          const double v = a[r];
          if (pmx_isfinite(v)) { ss += v; smn = dmin(smn, v); smx = dmax(smx, v); ++scf; }
          if (v == v) {
            if (scn > 0) { sd += delta_of(sfv, v); if (NRES) snr += sfv < v ? 1 : 0; } else { lv = v; lpos = r; }
            sfv = v; foff = r; ++scn;
          }
          const int si = r * NT + t;
          S_sum[si] = ss; S_mn[si] = smn; S_mx[si] = smx; S_d[si] = sd;
          S_meta[si] = (uint16_t)(scf | (scn << 4) | (((foff - r) & 7) << 8) | (NRES ? snr << 11 : 0));
        }
        T_sum[t] = ss; T_mn[t] = smn; T_mx[t] = smx; T_fv[t] = sfv; T_lv[t] = lv; T_di[t] = sd;
        T_meta[t] = (uint32_t)(scf | (scn << 4) | (foff << 8) | (lpos << 12) | (NRES ? snr << 16 : 0));
        c_fv = sfv; c_cn = scn;
      }
    }
    __syncthreads();

    // ---------------------------------------------------------------- windows
    // Two passes over the thread's 8 outputs, rolling statistics first, increase/rate second:
    // each pass keeps half of the state in registers (5 CTAs per SM instead of 4).
    const int kfirst = t - qa + 1 > 0 ? t - qa + 1 : 0;
    if (do_roll && e0 + E > HP) {
      double ms = 0.0, mmn = INF, mmx = -INF; int mcf = 0, mcn = 0;                 // chunks t-qa+1 .. t-1
      for (int k = kfirst; k < t; ++k) {
        const uint32_t m = T_meta[k];
        ms += T_sum[k]; mmn = dmin(mmn, T_mn[k]); mmx = dmax(mmx, T_mx[k]);
        mcf += m & 0xf; mcn += (m >> 4) & 0xf;
      }
      double ms2 = ms, mmn2 = mmn, mmx2 = mmx; int mcf2 = mcf, mcn2 = mcn;         // ... plus chunk t-qa
      if (t - qa >= 0) {
        const int k = t - qa;
        const uint32_t m = T_meta[k];
        ms2 = T_sum[k] + ms; mmn2 = dmin(T_mn[k], mmn); mmx2 = dmax(T_mx[k], mmx);
        mcf2 += m & 0xf; mcn2 += (m >> 4) & 0xf;
      }
      double ps = 0.0, pmn = INF, pmx = -INF; int pcf = 0, pcn = 0;
      // The first qb outputs start one chunk further back ("far"): they use the fold that includes
      // chunk t-qa and rows 8-qb.. of chunk t-qa-1.  With QB known at compile time the loops are
      // unrolled and `far` folds away.
      const int lcF = t - qa - 1 > 0 ? t - qa - 1 : 0, lcN = t - qa > 0 ? t - qa : 0;   // < 0: halo points, result unused
#pragma unroll(QB >= 0 ? 2 : 1)
      for (int h = 0; h < 2; ++h) {                               // two groups of four outputs
        double omn[4], omx[4], omean[4]; int ocnt[4];
        unsigned bad = 0;                                         // outputs with fewer than `need` values
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int r = 4 * h + q;
          const double v = QB >= 0 ? a[r] : (h ? a[4 + q] : a[q]);
          if (pmx_isfinite(v)) { ps += v; pmn = dmin(pmn, v); pmx = dmax(pmx, v); ++pcf; }
          pcn += (v == v) ? 1 : 0;
          const bool far = r < qb;
          const int si = far ? (r - qb + E) * NT + lcF : (r - qb) * NT + lcN;
          const uint32_t sm = S_meta[si];
          const double ws = S_sum[si] + (far ? ms2 : ms) + ps;
          const int wcf = (int)(sm & 0xf) + (far ? mcf2 : mcf) + pcf;
          omn[q] = dmin(dmin(S_mn[si], far ? mmn2 : mmn), pmn);
          omx[q] = dmax(dmax(S_mx[si], far ? mmx2 : mmx), pmx);
          omean[q] = ws * rcp[wcf];
          bad |= (wcf < need ? 1u : 0u) << q;
          ocnt[q] = (int)((sm >> 4) & 0xf) + (far ? mcn2 : mcn) + pcn;
        }
        if (bad) {                                                // rare: NaN runs
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if ((bad >> q) & 1) { omn[q] = pmx_nan(); omx[q] = pmx_nan(); omean[q] = pmx_nan(); }
        }
        const int e = e0 + 4 * h;
        if (e >= HP && e + 3 < n_el) {                            // whole group inside the output range
          const int64_t o = row + e;
          if (p.o_min) st4(p.o_min + o, omn[0], omn[1], omn[2], omn[3]);
          if (p.o_max) st4(p.o_max + o, omx[0], omx[1], omx[2], omx[3]);
          if (p.o_mean) st4(p.o_mean + o, omean[0], omean[1], omean[2], omean[3]);
          if (p.o_cnt) *reinterpret_cast<int4*>(p.o_cnt + o) = make_int4(ocnt[0], ocnt[1], ocnt[2], ocnt[3]);
        } else if (e + 3 >= HP && e < n_el) {                     // tile edge
          if (p.o_min) store4_edge(p.o_min + row, e, HP, n_el, omn[0], omn[1], omn[2], omn[3]);
          if (p.o_max) store4_edge(p.o_max + row, e, HP, n_el, omx[0], omx[1], omx[2], omx[3]);
          if (p.o_mean) store4_edge(p.o_mean + row, e, HP, n_el, omean[0], omean[1], omean[2], omean[3]);
          store4i(p.o_cnt ? p.o_cnt + row : nullptr, e, HP, n_el, ocnt[0], ocnt[1], ocnt[2], ocnt[3]);
        }
      }
    }
    if (do_rate && e0 + E > HP) {
      // fold the delta sums of the chunks between the window start and the own chunk, forming the
      // junction delta of every non-empty chunk with the last valid value before it
      double md = 0.0, md2 = 0.0;
      // mdB / bfpos: the same fold over the near chunks, started inside them (no junction into the
      // first non-empty one) -- what a window needs whose first chunk part holds no sample
      double mdB = 0.0; int bfpos = -1;
      double accF = 0.0; int fposF = -1;                          // ... and the same for "far" windows
      int mr = 0, mr2 = 0, mrB = 0, mrF = 0;                       // NRES: reset counts beside md, md2, mdB, accF
      double clv = 0.0; int clpos = -1;                            // last valid value / position before the own chunk
      if (t - qa - 1 >= 0) {
        const uint32_t m = T_meta[t - qa - 1];
        if ((m >> 4) & 0xf) { clv = T_lv[t - qa - 1]; clpos = (t - qa - 1) * E + (int)((m >> 12) & 7); }
      }
      if (t - qa >= 0) {
        const int k = t - qa;
        const uint32_t m = T_meta[k];
        if ((m >> 4) & 0xf) {
          accF = T_di[k];
          md2 = accF + (clpos >= 0 ? delta_of(T_fv[k], clv) : 0.0);
          if (NRES) { mrF = (int)((m >> 16) & 7); mr2 = mrF + ((clpos >= 0 && T_fv[k] < clv) ? 1 : 0); }
          fposF = k * E + (int)((m >> 8) & 7);
          clv = T_lv[k]; clpos = k * E + (int)((m >> 12) & 7);
        }
      }
      for (int k = kfirst; k < t; ++k) {
        const uint32_t m = T_meta[k];
        if ((m >> 4) & 0xf) {
          const double di = T_di[k];
          const double j = clpos >= 0 ? delta_of(T_fv[k], clv) : 0.0;
          md += di + j;
          mdB += di + (bfpos >= 0 ? j : 0.0);
          if (NRES) {
            const int nr = (int)((m >> 16) & 7), jr = (clpos >= 0 && T_fv[k] < clv) ? 1 : 0;
            mr += nr + jr;
            mrB += nr + (bfpos >= 0 ? jr : 0);
          }
          if (bfpos < 0) bfpos = k * E + (int)((m >> 8) & 7);
          clv = T_lv[k]; clpos = k * E + (int)((m >> 12) & 7);
        }
      }
      md2 += md;
      if (NRES) { mr2 += mr; if (fposF >= 0) mrF += mr; else mrF = mrB; }
      if (fposF >= 0) accF += md; else { accF = mdB; fposF = bfpos; }
      const double Jc = (c_cn > 0 && clpos >= 0) ? delta_of(c_fv, clv) : 0.0;     // own chunk's junction delta
      const int Jr = (NRES && c_cn > 0 && clpos >= 0 && c_fv < clv) ? 1 : 0;       // ... and whether it is a reset
      double pd = 0.0, pj = 0.0, plv = 0.0;                       // own prefix without / with its junction delta
      int pcn = 0, plpos = 0, pfpos = 0;
      int pr = 0, pjr = 0;                                        // NRES: resets in the own prefix without / with the junction
      const int lcF = t - qa - 1 > 0 ? t - qa - 1 : 0, lcN = t - qa > 0 ? t - qa : 0;
#pragma unroll(QB >= 0 ? 2 : 1)
      for (int h = 0; h < 2; ++h) {
        double oinc[4], orate[4];
        int onres[NRES ? 4 : 1];
        unsigned bad = 0;                                         // outputs with fewer than two samples
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int r = 4 * h + q;
          const double v = QB >= 0 ? a[r] : (h ? a[4 + q] : a[q]);
          if (v == v) {
            if (pcn > 0) {
              const double d = delta_of(v, plv); pd += d; pj += d;
              if (NRES) { const int dr = v < plv ? 1 : 0; pr += dr; pjr += dr; }
            } else { pfpos = e0 + r; pj = Jc; if (NRES) pjr = Jr; }
            plv = v; plpos = e0 + r; ++pcn;
          }
          const bool far = r < qb;
          const int si = far ? (r - qb + E) * NT + lcF : (r - qb) * NT + lcN;
          const double cmd = far ? md2 : md, cacc = far ? accF : mdB;
          const int cfp = far ? fposF : bfpos;
          const uint32_t sm = S_meta[si];
          int fpos; double inc;
          int nres = 0;
          if ((sm >> 4) & 0xf) {                                  // first chunk part holds a sample
            fpos = e0 + r - H + (int)((sm >> 8) & 7);             // window start + offset of its first sample
            inc = S_d[si] + cmd + pj;
            if (NRES) nres = (int)((sm >> 11) & 7) + (far ? mr2 : mr) + pjr;
          } else {                                                // rare: NaN run or series start
            fpos = cfp;
            if (fpos >= 0) { inc = cacc + pj; if (NRES) nres = (far ? mrF : mrB) + pjr; }
            else { fpos = pcn > 0 ? pfpos : 0x7fffffff; inc = pd; if (NRES) nres = pr; }   // window inside the own chunk
          }
          const int lpos = pcn > 0 ? plpos : clpos;
          const int span = lpos - fpos;
          if (NRES) onres[q] = span > 0 ? nres : 0;
          oinc[q] = inc;
          orate[q] = inc * rdt[span > 0 ? span : 0];
          bad |= (span > 0 ? 0u : 1u) << q;
        }
        if (bad) {
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if ((bad >> q) & 1) { oinc[q] = pmx_nan(); orate[q] = pmx_nan(); }
        }
        const int e = e0 + 4 * h;
        if (e >= HP && e + 3 < n_el) {
          const int64_t o = row + e;
          if (p.o_inc) st4(p.o_inc + o, oinc[0], oinc[1], oinc[2], oinc[3]);
          if (p.o_rate) st4(p.o_rate + o, orate[0], orate[1], orate[2], orate[3]);
          if (NRES && p.o_nres) *reinterpret_cast<int4*>(p.o_nres + o) = make_int4(onres[0], onres[1], onres[2], onres[3]);
        } else if (e + 3 >= HP && e < n_el) {
          if (p.o_inc) store4_edge(p.o_inc + row, e, HP, n_el, oinc[0], oinc[1], oinc[2], oinc[3]);
          if (p.o_rate) store4_edge(p.o_rate + row, e, HP, n_el, orate[0], orate[1], orate[2], orate[3]);
          if (NRES) store4i(p.o_nres ? p.o_nres + row : nullptr, e, HP, n_el, onres[0], onres[1], onres[2], onres[3]);
        }
      }
    }
    // S/T must not be overwritten before every thread has read them.  With ALIGN the next tile's
    // vote barrier (__syncthreads_and, before any shared-memory write) already orders that.
    if (!ALIGN) __syncthreads();
  }
}

bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; }

template <int NT, bool ALIGN, int QB, bool IDX = false, bool NRES = false>
int launch_fast(const pmx_ctx* ctx, WinParams p, cudaStream_t st) {
  using TL = FTile<NT>;
  const int out_per_tile = (TL::NE - (p.w - 1) - 3) & ~3;
  const int64_t tiles = (p.n_steps + out_per_tile - 1) / out_per_tile;
  // CTAs wanted: 16 per SM.  Measured on config 2 (6144 series): more, shorter CTAs are slower (every
  // CTA pays ~2 us of table set-up and cursor search: 0.523 / 0.533 / 0.543 / 0.612 ms at 16 / 64 / 128 / 256)
  const int64_t target = (int64_t)ctx->sm_count * 16;
  int64_t chunks = (target + p.n_series - 1) / p.n_series;
  if (chunks > tiles) chunks = tiles;
  if (chunks < 1) chunks = 1;
  const int64_t tiles_per_chunk = (tiles + chunks - 1) / chunks;
  p.chunk_len = tiles_per_chunk * out_per_tile;
  p.chunks = (int32_t)((p.n_steps + p.chunk_len - 1) / p.chunk_len);
  const int64_t grid = p.n_series * p.chunks;
  if (grid > 0x7fffffffLL) return pmx_fail(PMX_E_LIMIT, "window kernel: n_series * chunks exceeds 2^31-1");
  const size_t smem = TL::FIXED + size_t(2) * (p.w + 1) * 8 + 16;
  PMX_CUDA(cudaFuncSetAttribute(window_fast_kernel<NT, ALIGN, QB, IDX, NRES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  window_fast_kernel<NT, ALIGN, QB, IDX, NRES><<<(unsigned)grid, NT, smem, st>>>(p);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // namespace

template <bool ALIGN>
int launch_window_fast(const pmx_ctx* ctx, WinParams p, cudaStream_t st) {
  const int nt = window_block_threads();
  const int H = p.w - 1;
  const bool windows = p.o_min || p.o_max || p.o_mean || p.o_cnt || p.o_inc || p.o_rate;
  if ((windows && H < 8) || H > nt * E / 2 - 4) return 1;        // plain alignment takes any window
  if (p.rate_mode != 0) return 1;
  // reset counts: instantiated for dense input (pmx_rate) only
  if (p.o_nres && (ALIGN || nt != 128 || (reinterpret_cast<uintptr_t>(p.o_nres) & 15))) return 1;
  if (p.o_idx && (windows || !ALIGN || nt != 128 || (reinterpret_cast<uintptr_t>(p.o_idx) & 15))) return 1;
  // tile-relative timestamps are 32-bit: a tile (plus the lookback) must span < 2^30 ms
  if (p.step_ms > 0x3fffffffLL / (nt * E + 8) || p.lookback_ms >= 0x3fffffffLL) return 1;
  if (!aligned32(p.o_aligned) || !aligned32(p.o_min) || !aligned32(p.o_max) || !aligned32(p.o_mean) ||
      !aligned32(p.o_inc) || !aligned32(p.o_rate) || (reinterpret_cast<uintptr_t>(p.o_cnt) & 15))
    return 1;
  if constexpr (ALIGN) {
    if (p.o_idx) return launch_fast<128, true, 0, true>(ctx, p, st);   // pmx_align with source indices
  }
  if constexpr (!ALIGN) {
    if (p.o_nres) {
      switch (H & 7) {
        case 0: return launch_fast<128, false, 0, false, true>(ctx, p, st);
        case 1: return launch_fast<128, false, 1, false, true>(ctx, p, st);
        case 2: return launch_fast<128, false, 2, false, true>(ctx, p, st);
        case 3: return launch_fast<128, false, 3, false, true>(ctx, p, st);
        case 4: return launch_fast<128, false, 4, false, true>(ctx, p, st);
        case 5: return launch_fast<128, false, 5, false, true>(ctx, p, st);
        case 6: return launch_fast<128, false, 6, false, true>(ctx, p, st);
        default: return launch_fast<128, false, 7, false, true>(ctx, p, st);
      }
    }
  }
  if (nt == 128) {
    switch (H & 7) {                                              // one kernel per window phase
      case 0: return launch_fast<128, ALIGN, 0>(ctx, p, st);
      case 1: return launch_fast<128, ALIGN, 1>(ctx, p, st);
      case 2: return launch_fast<128, ALIGN, 2>(ctx, p, st);
      case 3: return launch_fast<128, ALIGN, 3>(ctx, p, st);
      case 4: return launch_fast<128, ALIGN, 4>(ctx, p, st);
      case 5: return launch_fast<128, ALIGN, 5>(ctx, p, st);
      case 6: return launch_fast<128, ALIGN, 6>(ctx, p, st);
      default: return launch_fast<128, ALIGN, 7>(ctx, p, st);
    }
  }
  return launch_fast<256, ALIGN, -1>(ctx, p, st);
}

}  // namespace pmxw
