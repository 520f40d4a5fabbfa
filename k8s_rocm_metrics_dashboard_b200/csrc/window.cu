// window.cu -- K2 align, K3 rolling min/max/mean/count, K4 increase/rate and their fusion.
//
// Work decomposition: one CTA per (series, chunk of grid points); the CTA walks its chunk in
// tiles of NE = 8*NT grid points (the first w-1 of each tile are halo).  Every thread owns 8
// consecutive grid points in registers.
//
//   stage   : the series' samples that can serve the tile are copied with coalesced loads into
//             shared memory (padded 9/8 so the per-thread sequential reads are conflict-free);
//   align   : each thread binary-searches the staged timestamps once, then walks forward
//             (searchsorted(side='right')-1, lookback and stale-marker rules of SURVEY X3);
//   windows : block decomposition without subtraction.  The window (e-w, e] of element e is
//             [suffix of the chunk that holds e-w+1] + [whole 8-chunks in between] + [prefix of
//             e's own chunk].  Suffix aggregates are produced by an in-register scan and
//             exchanged through a transposed shared array, chunk totals through a second one,
//             the prefix stays in registers.  Sums are therefore fresh per window (no prefix
//             differences), min/max are exact, and the rate aggregate (first/last sample,
//             sum of pre-reset values) is an ordered monoid combined the same way;
//   store   : results go through a per-warp shared buffer so that global stores are coalesced.
//
// HBM traffic per grid point in the fused call: 16 B in (ts,val), 40 B out.
#include <math.h>
#include <stdlib.h>
#include "window_common.cuh"
using namespace pmxw;

namespace {

template <int NT> struct Tile {
  static constexpr int NE = NT * E;                 // grid points per tile (halo + outputs)
  static constexpr int NS = NE + E;                 // staged samples per tile
  static constexpr int NSP = NS + NS / 8 + 2;       // ... padded
  static constexpr int NW = NT / 32;
  static constexpr int WB = 256 + 32 + 2;           // per-warp output buffer (doubles), padded
  static constexpr size_t STAGE_BYTES = size_t(2) * NSP * 8;
  static constexpr size_t S_BYTES = size_t(3) * NE * 8 + size_t(NE) * 2;
  static constexpr size_t A_BYTES = ((STAGE_BYTES > S_BYTES ? STAGE_BYTES : S_BYTES) + 15) & ~size_t(15);
  static constexpr size_t T_BYTES = size_t(3) * NT * 8 + size_t(NT) * 4;
  static constexpr size_t W_BYTES = size_t(NW) * WB * 8;
  static constexpr size_t SMEM = A_BYTES + T_BYTES + W_BYTES + 16;
};

__device__ __forceinline__ int pad8(int i) { return i + (i >> 3); }

// samples of one series as seen by a tile: staged (padded shared memory) or global
struct SampleView {
  const int64_t* ts; const double* val; int64_t n; bool padded;
  __device__ __forceinline__ int64_t T(int64_t i) const { return padded ? ts[i + (i >> 3)] : ts[i]; }
  __device__ __forceinline__ double V(int64_t i) const { return padded ? val[i + (i >> 3)] : val[i]; }
};

// largest i in [lo, n) with T(i) <= g, or lo-1
__device__ __forceinline__ int64_t floor_index(const SampleView& v, int64_t lo, int64_t g) {
  int64_t hi = v.n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (v.T(mid) <= g) lo = mid + 1; else hi = mid;
  }
  return lo - 1;
}

// Uniform search in global memory for the first tile of a chunk: guess from the mean sample
// spacing (exact for regularly scraped series), verify, else binary search.
__device__ int64_t floor_index_global(const int64_t* __restrict__ ts, int64_t n, int64_t g) {
  if (n == 0) return -1;
  const int64_t t0 = ts[0];
  if (g < t0) return -1;
  const int64_t tl = ts[n - 1];
  if (g >= tl) return n - 1;
  // n >= 2 here and t0 <= g < tl
  int64_t guess = (int64_t)((double)(g - t0) * (double)(n - 1) / (double)(tl - t0));
  if (guess < 0) guess = 0;
  if (guess > n - 2) guess = n - 2;
  if (ts[guess] <= g && g < ts[guess + 1]) return guess;
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (ts[mid] <= g) lo = mid + 1; else hi = mid;
  }
  return lo - 1;
}

template <typename T>
__device__ __forceinline__ void warp_store(T* wbuf, const T (&v)[E], T* __restrict__ out_row,
                                           int warp_e0, int e_lo, int e_hi, int lane) {
#pragma unroll
  for (int r = 0; r < E; ++r) wbuf[lane * (E + 1) + r] = v[r];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < E; ++k) {
    const int i = k * 32 + lane;
    const int e = warp_e0 + i;
    if (e >= e_lo && e < e_hi) out_row[e] = wbuf[pad8(i)];
  }
  __syncwarp();
}

// ordered monoid of the rate computation over a run of grid points.  rs = sum of the per-sample
// deltas x_i - x_{i-1} (x_i itself after a drop) between the run's first and last sample: the
// telescoped form of (last - first) + pre-reset values, without its cancellation after a reset
struct RateAgg {
  int cnt, nres, fpos, lpos;
  double fv, lv, rs;
};
__device__ __forceinline__ RateAgg rate_identity() { return RateAgg{0, 0, 0, 0, 0.0, 0.0, 0.0}; }
__device__ __forceinline__ RateAgg rate_combine(const RateAgg& A, const RateAgg& B) {
  if (A.cnt == 0) return B;
  if (B.cnt == 0) return A;
  RateAgg C;
  const bool reset = B.fv < A.lv;
  C.cnt = A.cnt + B.cnt;
  C.nres = A.nres + B.nres + (reset ? 1 : 0);
  C.fpos = A.fpos; C.fv = A.fv; C.lpos = B.lpos; C.lv = B.lv;
  C.rs = A.rs + B.rs + (reset ? B.fv : B.fv - A.lv);
  return C;
}
__device__ __forceinline__ void rate_append(RateAgg& P, double v, int pos) {
  if (pmx_isnan(v)) return;
  if (P.cnt > 0) {
    const bool reset = v < P.lv;
    P.rs += reset ? v : v - P.lv;
    P.nres += reset ? 1 : 0;
  } else { P.fv = v; P.fpos = pos; }
  P.lv = v; P.lpos = pos; P.cnt++;
}

template <int NT, bool ALIGN>
__global__ void __launch_bounds__(NT, (NT == 256 ? 2 : 4))
window_kernel(const WinParams p) {
  using TL = Tile<NT>;
  constexpr int NE = TL::NE;
  extern __shared__ __align__(16) unsigned char smem[];
  int64_t* s_ts = reinterpret_cast<int64_t*>(smem);
  double* s_val = reinterpret_cast<double*>(smem) + TL::NSP;
  double* S0 = reinterpret_cast<double*>(smem);
  double* S1 = S0 + NE;
  double* S2 = S1 + NE;
  uint16_t* Sm = reinterpret_cast<uint16_t*>(S2 + NE);
  double* T0 = reinterpret_cast<double*>(smem + TL::A_BYTES);
  double* T1 = T0 + NT;
  double* T2 = T1 + NT;
  uint32_t* Tm = reinterpret_cast<uint32_t*>(T2 + NT);
  double* wbuf_all = reinterpret_cast<double*>(smem + TL::A_BYTES + TL::T_BYTES);
  int64_t* s_cursor = reinterpret_cast<int64_t*>(smem + TL::A_BYTES + TL::T_BYTES + TL::W_BYTES);

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  double* wbuf = wbuf_all + warp * TL::WB;
  const int64_t s = blockIdx.x / p.chunks;
  const int ci = blockIdx.x % p.chunks;
  const int64_t c0 = (int64_t)ci * p.chunk_len;
  const int64_t c1 = (c0 + p.chunk_len < p.n_steps) ? c0 + p.chunk_len : p.n_steps;
  if (c0 >= c1) return;
  const int H = p.w - 1;                 // halo grid points per tile
  const int OUT = NE - H;                // outputs per full tile
  const int qa = H / E, qb = H % E;
  const bool do_roll = p.o_min || p.o_max || p.o_mean || p.o_cnt;
  const bool do_rate = p.o_inc || p.o_rate || p.o_nres;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);

  int64_t base = 0, len = 0, cur = 0;
  if (ALIGN) { base = p.off[s]; len = p.off[s + 1] - base; }
  bool first = true;
  const int e0 = t * E;
  const int warp_e0 = warp * 32 * E;

  for (int64_t jt = c0; jt < c1; jt += OUT) {
    const int64_t j0 = jt - H;                                  // grid index of tile element 0
    const int n_out = (int)((c1 - jt < OUT) ? (c1 - jt) : OUT);
    const int n_el = H + n_out;
    const int64_t row = s * p.n_steps + j0;                     // output index of tile element 0
    double a[E];

    if (ALIGN) {
      const int64_t jlo = j0 > 0 ? j0 : 0;
      if (first) {
        cur = floor_index_global(p.ts + base, len, p.start_ms + jlo * p.step_ms);
        if (cur < 0) cur = 0;
        first = false;
      }
      const int64_t rem = len - cur;
      const int ns = (int)(rem < TL::NS ? rem : TL::NS);
      for (int i = t; i < ns; i += NT) {
        s_ts[pad8(i)] = p.ts[base + cur + i];
        s_val[pad8(i)] = p.val[base + cur + i];
      }
      __syncthreads();
      const int64_t ghi = p.start_ms + (j0 + n_el - 1) * p.step_ms;
      const bool covered = (cur + ns >= len) || (ns > 0 && s_ts[pad8(ns - 1)] > ghi);
      SampleView sv;
      if (covered) { sv.ts = s_ts; sv.val = s_val; sv.n = ns; sv.padded = true; }
      else { sv.ts = p.ts + base + cur; sv.val = p.val + base + cur; sv.n = rem; sv.padded = false; }
      int64_t idx = -2;                                          // -2: not searched yet
      const int e_cur = OUT < NE - 1 ? OUT : NE - 1;             // element whose floor seeds the next tile
#pragma unroll
      for (int r = 0; r < E; ++r) {
        const int e = e0 + r;
        const int64_t j = j0 + e;
        double v = pmx_nan();
        int32_t src = -1;
        if (j >= 0 && e < n_el) {
          const int64_t g = p.start_ms + j * p.step_ms;
          if (idx == -2) idx = floor_index(sv, 0, g);
          else if (idx + 1 < sv.n && sv.T(idx + 1) <= g) {
            ++idx;
            if (idx + 1 < sv.n && sv.T(idx + 1) <= g) idx = floor_index(sv, idx + 1, g);
          }
          if (idx >= 0) {
            const double sample = sv.V(idx);
            if (g - sv.T(idx) < p.lookback_ms &&
                (unsigned long long)__double_as_longlong(sample) != PMX_STALE_NAN_BITS) {
              v = sample;
              src = (int32_t)(cur + idx);
            }
          }
          if (e == e_cur) *s_cursor = cur + (idx > 0 ? idx : 0);
          if (p.o_idx && e >= H) p.o_idx[row + e] = src;
        }
        a[r] = v;
      }
      __syncthreads();                                           // staging dead, cursor visible
      cur = *s_cursor;
    } else {
      const double* xr = p.x + s * p.n_steps;
      for (int i = t; i < n_el; i += NT) {
        const int64_t j = j0 + i;
        s_val[pad8(i)] = j >= 0 ? xr[j] : pmx_nan();
      }
      __syncthreads();
#pragma unroll
      for (int r = 0; r < E; ++r) a[r] = (e0 + r < n_el) ? s_val[pad8(e0 + r)] : pmx_nan();
      __syncthreads();
    }

    if (p.o_aligned) warp_store<double>(wbuf, a, p.o_aligned + row, warp_e0, H, n_el, lane);

    // ------------------------------------------------------------------ rolling min/max/mean
    if (do_roll) {
      {
        double ss = 0.0, smn = INF, smx = -INF; int cf = 0, cn = 0;
#pragma unroll
        for (int r = E - 1; r >= 0; --r) {
          const double v = a[r];
          const bool fin = pmx_isfinite(v);
          if (fin) { ss += v; smn = fmin(smn, v); smx = fmax(smx, v); }
          cf += fin ? 1 : 0; cn += pmx_isnan(v) ? 0 : 1;
          S0[r * NT + t] = ss; S1[r * NT + t] = smn; S2[r * NT + t] = smx;
          Sm[r * NT + t] = (uint16_t)(cf | (cn << 4));
        }
        T0[t] = ss; T1[t] = smn; T2[t] = smx; Tm[t] = (uint32_t)(cf | (cn << 8));
      }
      __syncthreads();
      double omn[E], omx[E], omean[E]; int32_t ocnt[E];
      if (e0 + E > H) {                                          // thread owns at least one output
        double ms = 0.0, mmn = INF, mmx = -INF; int mcf = 0, mcn = 0;
        for (int k = (t - qa + 1 > 0 ? t - qa + 1 : 0); k < t; ++k) {
          ms += T0[k]; mmn = fmin(mmn, T1[k]); mmx = fmax(mmx, T2[k]);
          const uint32_t m = Tm[k]; mcf += m & 0xff; mcn += m >> 8;
        }
        double ms2 = ms, mmn2 = mmn, mmx2 = mmx; int mcf2 = mcf, mcn2 = mcn;
        if (qa >= 1 && t - qa >= 0) {
          const int k = t - qa;
          ms2 = T0[k] + ms; mmn2 = fmin(mmn, T1[k]); mmx2 = fmax(mmx, T2[k]);
          const uint32_t m = Tm[k]; mcf2 += m & 0xff; mcn2 += m >> 8;
        }
        double ps = 0.0, pmn = INF, pmx = -INF; int pcf = 0, pcn = 0;
        const int need = p.min_periods > 1 ? p.min_periods : 1;
#pragma unroll
        for (int r = 0; r < E; ++r) {
          const double v = a[r];
          const bool fin = pmx_isfinite(v);
          if (fin) { ps += v; pmn = fmin(pmn, v); pmx = fmax(pmx, v); }
          pcf += fin ? 1 : 0; pcn += pmx_isnan(v) ? 0 : 1;
          double ws, wmn, wmx; int wcf, wcn;
          if (qa == 0 && r >= qb) {                              // window inside this chunk
            ws = 0.0; wmn = INF; wmx = -INF; wcf = 0; wcn = 0;
#pragma unroll
            for (int d = E - 1; d >= 0; --d) {
              if (d <= r && d <= qb) {
                const double u = a[r - d];
                const bool f = pmx_isfinite(u);
                if (f) { ws += u; wmn = fmin(wmn, u); wmx = fmax(wmx, u); }
                wcf += f ? 1 : 0; wcn += pmx_isnan(u) ? 0 : 1;
              }
            }
          } else {
            const bool far = r < qb;                             // window starts one chunk further back
            const int lchunk = far ? t - qa - 1 : t - qa;
            const int lrow = far ? r - qb + E : r - qb;
            ws = far ? ms2 : ms; wmn = far ? mmn2 : mmn; wmx = far ? mmx2 : mmx;
            wcf = far ? mcf2 : mcf; wcn = far ? mcn2 : mcn;
            if (lchunk >= 0) {
              const int si = lrow * NT + lchunk;
              ws = S0[si] + ws; wmn = fmin(wmn, S1[si]); wmx = fmax(wmx, S2[si]);
              const uint32_t m = Sm[si]; wcf += m & 0xf; wcn += m >> 4;
            }
            ws += ps; wmn = fmin(wmn, pmn); wmx = fmax(wmx, pmx); wcf += pcf; wcn += pcn;
          }
          const bool ok = wcf >= need;
          omean[r] = ok ? ws / (double)wcf : pmx_nan();
          omn[r] = ok ? wmn : pmx_nan();
          omx[r] = ok ? wmx : pmx_nan();
          ocnt[r] = wcn;
        }
      } else {
#pragma unroll
        for (int r = 0; r < E; ++r) { omn[r] = omx[r] = omean[r] = pmx_nan(); ocnt[r] = 0; }
      }
      if (p.o_min) warp_store<double>(wbuf, omn, p.o_min + row, warp_e0, H, n_el, lane);
      if (p.o_max) warp_store<double>(wbuf, omx, p.o_max + row, warp_e0, H, n_el, lane);
      if (p.o_mean) warp_store<double>(wbuf, omean, p.o_mean + row, warp_e0, H, n_el, lane);
      if (p.o_cnt) warp_store<int32_t>(reinterpret_cast<int32_t*>(wbuf), ocnt, p.o_cnt + row, warp_e0, H, n_el, lane);
      __syncthreads();                                           // S/T arrays free again
    }

    // ------------------------------------------------------------------ increase / rate
    if (do_rate) {
      {
        int cnt = 0, nres = 0, fpos = 0, lpos = 0; double fv = 0.0, lv = 0.0, rs = 0.0;
#pragma unroll
        for (int r = E - 1; r >= 0; --r) {
          const double v = a[r];
          if (!pmx_isnan(v)) {
            if (cnt > 0) { const bool reset = fv < v; rs += reset ? fv : fv - v; nres += reset ? 1 : 0; }
            else { lv = v; lpos = r; }
            fv = v; fpos = r; cnt++;
          }
          S0[r * NT + t] = fv; S1[r * NT + t] = rs;
          Sm[r * NT + t] = (uint16_t)(cnt | (nres << 4) | (fpos << 8));
        }
        T0[t] = fv; T1[t] = lv; T2[t] = rs;
        Tm[t] = (uint32_t)(cnt | (nres << 4) | (fpos << 8) | (lpos << 12));
      }
      __syncthreads();
      double oinc[E], orate[E]; int32_t onres[E];
      if (e0 + E > H) {
        RateAgg M = rate_identity();
        for (int k = (t - qa + 1 > 0 ? t - qa + 1 : 0); k < t; ++k) {
          const uint32_t m = Tm[k];
          RateAgg B{(int)(m & 0xf), (int)((m >> 4) & 0xf), k * E + (int)((m >> 8) & 0xf),
                    k * E + (int)((m >> 12) & 0xf), T0[k], T1[k], T2[k]};
          M = rate_combine(M, B);
        }
        RateAgg M2 = M;
        if (qa >= 1 && t - qa >= 0) {
          const int k = t - qa;
          const uint32_t m = Tm[k];
          RateAgg B{(int)(m & 0xf), (int)((m >> 4) & 0xf), k * E + (int)((m >> 8) & 0xf),
                    k * E + (int)((m >> 12) & 0xf), T0[k], T1[k], T2[k]};
          M2 = rate_combine(B, M);
        }
        RateAgg P = rate_identity();
#pragma unroll
        for (int r = 0; r < E; ++r) {
          rate_append(P, a[r], e0 + r);
          RateAgg W;
          if (qa == 0 && r >= qb) {
            W = rate_identity();
#pragma unroll
            for (int d = E - 1; d >= 0; --d)
              if (d <= r && d <= qb) rate_append(W, a[r - d], e0 + r - d);
          } else {
            const bool far = r < qb;
            const int lchunk = far ? t - qa - 1 : t - qa;
            const int lrow = far ? r - qb + E : r - qb;
            W = far ? M2 : M;
            if (lchunk >= 0) {
              const int si = lrow * NT + lchunk;
              const uint32_t m = Sm[si];
              const uint32_t tm = Tm[lchunk];
              RateAgg A{(int)(m & 0xf), (int)((m >> 4) & 0xf), lchunk * E + (int)((m >> 8) & 0xf),
                        lchunk * E + (int)((tm >> 12) & 0xf), S0[si], T1[lchunk], S1[si]};
              W = rate_combine(A, W);
            }
            W = rate_combine(W, P);
          }
          double inc = pmx_nan(), rate = pmx_nan();
          if (W.cnt >= 2) {
            inc = W.rs;
            const double dt = (double)((int64_t)(W.lpos - W.fpos) * p.step_ms) / 1000.0;
            rate = inc / dt;
            if (p.rate_mode == 1) {
              const int e = e0 + r;                          // window (e - w, e] in tile element indices
              pmx_extrapolate(inc, W.fv, W.cnt, dt, (double)((int64_t)(W.fpos - (e - p.w)) * p.step_ms) / 1000.0,
                              (double)((int64_t)(e - W.lpos) * p.step_ms) / 1000.0,
                              (double)((int64_t)p.w * p.step_ms) / 1000.0, &inc, &rate);
            }
          }
          oinc[r] = inc; orate[r] = rate; onres[r] = W.cnt >= 2 ? W.nres : 0;
        }
      } else {
#pragma unroll
        for (int r = 0; r < E; ++r) { oinc[r] = orate[r] = pmx_nan(); onres[r] = 0; }
      }
      if (p.o_inc) warp_store<double>(wbuf, oinc, p.o_inc + row, warp_e0, H, n_el, lane);
      if (p.o_rate) warp_store<double>(wbuf, orate, p.o_rate + row, warp_e0, H, n_el, lane);
      if (p.o_nres) warp_store<int32_t>(reinterpret_cast<int32_t*>(wbuf), onres, p.o_nres + row, warp_e0, H, n_el, lane);
      __syncthreads();
    }
  }
}

// Direct per-output loops for windows too long for the tiled kernel (w-1 > NE/2).  Same
// definitions; reads come from L1/L2.
__global__ void __launch_bounds__(256)
window_naive_kernel(const WinParams p) {
  const int64_t total = p.n_series * p.n_steps;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const int need = p.min_periods > 1 ? p.min_periods : 1;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < total;
       o += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = o / p.n_steps, j = o % p.n_steps;
    const double* xr = p.x + s * p.n_steps;
    const int64_t lo = j - p.w + 1 > 0 ? j - p.w + 1 : 0;
    double ws = 0.0, wmn = INF, wmx = -INF; int wcf = 0, wcn = 0;
    RateAgg W = rate_identity();
    for (int64_t k = lo; k <= j; ++k) {
      const double v = xr[k];
      if (pmx_isfinite(v)) { ws += v; wmn = fmin(wmn, v); wmx = fmax(wmx, v); ++wcf; }
      wcn += pmx_isnan(v) ? 0 : 1;
      rate_append(W, v, (int)(k - lo));
    }
    const bool ok = wcf >= need;
    if (p.o_min) p.o_min[o] = ok ? wmn : pmx_nan();
    if (p.o_max) p.o_max[o] = ok ? wmx : pmx_nan();
    if (p.o_mean) p.o_mean[o] = ok ? ws / (double)wcf : pmx_nan();
    if (p.o_cnt) p.o_cnt[o] = wcn;
    double inc = pmx_nan(), rate = pmx_nan();
    if (W.cnt >= 2) {
      inc = W.rs;
      const double dt = (double)((int64_t)(W.lpos - W.fpos) * p.step_ms) / 1000.0;
      rate = inc / dt;
      if (p.rate_mode == 1) {                              // positions are relative to lo; window (j - w, j]
        const int64_t wstart = j - p.w - lo;
        pmx_extrapolate(inc, W.fv, W.cnt, dt, (double)((W.fpos - wstart) * p.step_ms) / 1000.0,
                        (double)((j - lo - W.lpos) * p.step_ms) / 1000.0,
                        (double)((int64_t)p.w * p.step_ms) / 1000.0, &inc, &rate);
      }
    }
    if (p.o_inc) p.o_inc[o] = inc;
    if (p.o_rate) p.o_rate[o] = rate;
    if (p.o_nres) p.o_nres[o] = W.cnt >= 2 ? W.nres : 0;
  }
}

template <int NT, bool ALIGN>
int launch_window(const pmx_ctx* ctx, WinParams p, cudaStream_t st) {
  using TL = Tile<NT>;
  const int out_per_tile = TL::NE - (p.w - 1);
  const int64_t tiles = (p.n_steps + out_per_tile - 1) / out_per_tile;
  // enough CTAs to fill the machine a few times over even when there are few series
  const int64_t target = (int64_t)ctx->sm_count * 16;
  int64_t chunks = (target + p.n_series - 1) / p.n_series;
  if (chunks > tiles) chunks = tiles;
  if (chunks < 1) chunks = 1;
  const int64_t tiles_per_chunk = (tiles + chunks - 1) / chunks;
  p.chunk_len = tiles_per_chunk * out_per_tile;
  p.chunks = (int32_t)((p.n_steps + p.chunk_len - 1) / p.chunk_len);
  const int64_t grid = p.n_series * p.chunks;
  if (grid > 0x7fffffffLL) return pmx_fail(PMX_E_LIMIT, "window kernel: n_series * chunks exceeds 2^31-1");
  PMX_CUDA(cudaFuncSetAttribute(window_kernel<NT, ALIGN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TL::SMEM));
  window_kernel<NT, ALIGN><<<(unsigned)grid, NT, TL::SMEM, st>>>(p);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

template <bool ALIGN>
int dispatch_window(const pmx_ctx* ctx, WinParams p, cudaStream_t st) {
  if (p.n_series == 0 || p.n_steps == 0) return PMX_OK;
  const int rc = launch_window_fast<ALIGN>(ctx, p, st);      // the roofline kernel when it applies
  if (rc != 1) return rc;
  const int nt = window_block_threads();
  if (nt == 128) return launch_window<128, ALIGN>(ctx, p, st);
  return launch_window<256, ALIGN>(ctx, p, st);
}

}  // namespace

namespace pmxw {
int window_block_threads() { return 128; }                 // (256-thread CTAs: 2 per SM, measured slower)
}  // namespace pmxw

extern "C" {

int pmx_align(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_off,
              int64_t n_series, int64_t start_ms, int64_t step_ms, int64_t n_steps,
              int64_t lookback_ms, double* d_out_val, int32_t* d_out_idx, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0 && step_ms > 0 && lookback_ms > 0, "bad sizes");
  if (n_series == 0 || n_steps == 0) return PMX_OK;
  PMX_REQUIRE(d_off && (d_out_val || d_out_idx), "NULL buffer");
  WinParams p{};
  p.ts = d_ts; p.val = d_val; p.off = d_off; p.n_series = n_series; p.n_steps = n_steps;
  p.start_ms = start_ms; p.step_ms = step_ms; p.lookback_ms = lookback_ms; p.w = 1; p.min_periods = 1;
  p.o_aligned = d_out_val; p.o_idx = d_out_idx;
  return dispatch_window<true>(ctx, p, pmx_st(stream));
}

static int window_dense(pmx_ctx* ctx, WinParams p, cudaStream_t st) {
  if (p.n_series == 0 || p.n_steps == 0) return PMX_OK;
  const int nt = window_block_threads();
  if (p.w - 1 > nt * E / 2) {                       // window longer than half a tile
    const int64_t total = p.n_series * p.n_steps;
    int64_t blocks = (total + 255) / 256;
    if (blocks > (int64_t)ctx->sm_count * 32) blocks = (int64_t)ctx->sm_count * 32;
    window_naive_kernel<<<(unsigned)blocks, 256, 0, st>>>(p);
    PMX_LAUNCH_CHECK();
    return PMX_OK;
  }
  return dispatch_window<false>(ctx, p, st);
}

int pmx_rolling(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                int64_t window_steps, int64_t min_periods, double* d_min, double* d_max,
                double* d_mean, int32_t* d_count, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0, "bad sizes");
  PMX_REQUIRE(window_steps >= 1 && window_steps <= 0x7fffffff, "window_steps must be >= 1");
  PMX_REQUIRE(min_periods >= 0 && min_periods <= 0x7fffffff, "bad min_periods");
  if (n_series == 0 || n_steps == 0) return PMX_OK;
  PMX_REQUIRE(d_x && (d_min || d_max || d_mean || d_count), "NULL buffer");
  WinParams p{};
  p.x = d_x; p.n_series = n_series; p.n_steps = n_steps; p.step_ms = 1;
  p.w = (int32_t)window_steps; p.min_periods = (int32_t)min_periods;
  p.o_min = d_min; p.o_max = d_max; p.o_mean = d_mean; p.o_cnt = d_count;
  return window_dense(ctx, p, pmx_st(stream));
}

int pmx_rate(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
             int64_t window_steps, int64_t step_ms, int mode, double* d_increase, double* d_rate,
             int32_t* d_n_resets, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(mode == 0 || mode == 1, "mode must be 0 (first..last) or 1 (extrapolated)");
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0 && step_ms > 0, "bad sizes");
  PMX_REQUIRE(window_steps >= 1 && window_steps <= 0x7fffffff, "window_steps must be >= 1");
  if (n_series == 0 || n_steps == 0) return PMX_OK;
  PMX_REQUIRE(d_x && (d_increase || d_rate || d_n_resets), "NULL buffer");
  WinParams p{};
  p.x = d_x; p.n_series = n_series; p.n_steps = n_steps; p.step_ms = step_ms;
  p.w = (int32_t)window_steps; p.min_periods = 1; p.rate_mode = mode;
  p.o_inc = d_increase; p.o_rate = d_rate; p.o_nres = d_n_resets;
  return window_dense(ctx, p, pmx_st(stream));
}

int pmx_fused_align_rolling_rate(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val,
                                 const int64_t* d_off, int64_t n_series, int64_t start_ms,
                                 int64_t step_ms, int64_t n_steps, int64_t lookback_ms,
                                 int64_t window_steps, int64_t min_periods, double* d_aligned,
                                 double* d_min, double* d_max, double* d_mean, double* d_rate,
                                 pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0 && step_ms > 0 && lookback_ms > 0, "bad sizes");
  PMX_REQUIRE(window_steps >= 1, "window_steps must be >= 1");
  PMX_REQUIRE(min_periods >= 0 && min_periods <= 0x7fffffff, "bad min_periods");
  if (n_series == 0 || n_steps == 0) return PMX_OK;
  PMX_REQUIRE(d_off, "NULL buffer");
  const int nt = window_block_threads();
  if (window_steps - 1 > nt * E / 2)
    return pmx_fail(PMX_E_LIMIT, "pmx_fused_align_rolling_rate: window_steps %lld exceeds the fused limit %d; "
                    "use pmx_align + pmx_rolling + pmx_rate", (long long)window_steps, nt * E / 2 + 1);
  WinParams p{};
  p.ts = d_ts; p.val = d_val; p.off = d_off; p.n_series = n_series; p.n_steps = n_steps;
  p.start_ms = start_ms; p.step_ms = step_ms; p.lookback_ms = lookback_ms;
  p.w = (int32_t)window_steps; p.min_periods = (int32_t)min_periods;
  p.o_aligned = d_aligned; p.o_min = d_min; p.o_max = d_max; p.o_mean = d_mean; p.o_rate = d_rate;
  return dispatch_window<true>(ctx, p, pmx_st(stream));
}

}  // extern "C"
