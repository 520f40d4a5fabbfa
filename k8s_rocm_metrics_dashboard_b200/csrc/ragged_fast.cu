// ragged_fast.cu -- rolling min/max/mean/count and increase/rate at EVERY SAMPLE of ragged series
// (config 5: high-cardinality series, irregular timestamps, window (t_i - W, t_i]).
//
// Same window algebra as window_fast.cu, with a window start that differs per output:
//   * the flat CSR sample arrays are cut into tiles of 1024 samples (128 halo + 896 outputs); a
//     tile may hold pieces of many series.  Every thread owns 8 consecutive samples, which it
//     loads straight from global memory (full sectors) and publishes their timestamps in shared
//     memory; the series starts that fall into the tile are staged next to them;
//   * per output: series start (walk over the staged boundaries) and window start l (two-pointer
//     over the staged timestamps, clipped to the series start);
//   * window [l, e] = suffix of l's 8-chunk + whole chunks in between + prefix of the own chunk;
//     nothing is subtracted.  Windows never cross a series start, so every aggregate that is
//     used lies inside one series.  The chunk totals shared by the thread's 8 outputs are folded
//     once, the few chunks that differ per output are added on the fly;
//   * outputs whose window reaches before the halo, starts in an all-NaN chunk part, or whose tile
//     holds more series starts than fit, are folded directly from global memory (out of line).
// n_resets is not produced here (ragged.cu's direct kernel serves that request).
#include <math.h>
#include "ragged_common.cuh"

namespace pmxr {
namespace {

constexpr int E = 8;
constexpr int NT = 128;
constexpr int NE = NT * E;          // samples per tile
constexpr int HALO = 128;
constexpr int OUT = NE - HALO;
constexpr int NSP = NE + NE / 8 + 2;
constexpr int BMAX = NE + 128;      // series starts kept per tile

__device__ __forceinline__ int pad8(int i) { return i + (i >> 3); }
__device__ __forceinline__ double dmin(double a, double b) { return b < a ? b : a; }
__device__ __forceinline__ double dmax(double a, double b) { return b > a ? b : a; }
__device__ __forceinline__ double delta_of(double cur, double prev) { return cur < prev ? cur : cur - prev; }
__device__ __forceinline__ void st4(double* p, double a, double b, double c, double d) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}

// The arrays of the statistics that a launch does not produce shrink to one element: 52.7 KB for
// both, 40.8 KB for the rolling statistics alone, 25.7 KB for increase / rate alone.  Measured on
// 400 000 series: rolling alone keeps four CTAs per SM (126 registers; five at 96 registers spill and
// are slower, 28.7 against 34.5 G samples/s) and gains from the code it no longer carries;
// increase / rate alone runs six CTAs per SM at 80 registers (36.2; 33.3 at five, 34.3 at seven).
template <bool ROLL, bool RATE>
struct Smem {
  int64_t ts[NSP];
  double S_sum[ROLL ? NE : 1], S_mn[ROLL ? NE : 1], S_mx[ROLL ? NE : 1], S_d[RATE ? NE : 1];
  double T_sum[ROLL ? NT : 1], T_mn[ROLL ? NT : 1], T_mx[ROLL ? NT : 1];
  double T_fv[RATE ? NT : 1], T_lv[RATE ? NT : 1], T_di[RATE ? NT : 1], T_tj[RATE ? NT : 1];
  uint32_t T_meta[NT];
  uint16_t B[BMAX];                  // tile-relative (0 .. NE): 16 bits do; with the direct folds returning in
                                     // registers the CTA needs 52.7 KB instead of 59 and four fit an SM
  uint16_t S_meta[NE];
  int nb;
  int overflow;
  int64_t next_sc;
  int64_t next_off_sc;
};

// Direct fold of one output from global memory.  Cold.  Returns min, max, mean, count (roll)
// or increase, rate (rate); `what` 0 = rolling, 1 = rate.
struct Folded { double v[4]; };
__device__ __noinline__ Folded fold_direct(const int64_t* __restrict__ ts, const double* __restrict__ val,
                                           int64_t sstart, int64_t f, int64_t window_ms, int need, int what) {
  Folded res;
  double* out = res.v;
  out[0] = out[1] = out[2] = out[3] = 0.0;
  const int64_t edge = ts[f] - window_ms;
  int64_t lo = sstart, hi = f;
  while (lo < hi) { const int64_t m = (lo + hi) >> 1; if (ts[m] <= edge) lo = m + 1; else hi = m; }
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  if (what == 0) {
    double ws = 0.0, wmn = INF, wmx = -INF; int wcf = 0, wcn = 0;
    for (int64_t k = lo; k <= f; ++k) {
      const double u = val[k];
      if (u != u) continue;
      ++wcn;
      if (pmx_isfinite(u)) { ws += u; wmn = dmin(wmn, u); wmx = dmax(wmx, u); ++wcf; }
    }
    const bool ok = wcf >= need;
    out[0] = ok ? wmn : pmx_nan(); out[1] = ok ? wmx : pmx_nan();
    out[2] = ok ? ws / (double)wcf : pmx_nan(); out[3] = (double)wcn;
  } else {
    int cnt = 0; double lv = 0.0, acc = 0.0; int64_t ft = 0, lt = 0;
    for (int64_t k = lo; k <= f; ++k) {
      const double u = val[k];
      if (u != u) continue;
      if (cnt > 0) acc += delta_of(u, lv); else ft = ts[k];
      lv = u; lt = ts[k]; ++cnt;
    }
    const bool have = cnt >= 2 && lt > ft;
    out[0] = cnt >= 2 ? acc : pmx_nan();
    out[1] = have ? acc * 1000.0 / (double)(lt - ft) : pmx_nan();
  }
  return res;
}

template <bool ROLL, bool RATE>
__global__ void __launch_bounds__(NT, ROLL ? 4 : 6)
ragged_fast_kernel(const RParams p, int64_t n_tiles, int tiles_per_cta) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem<ROLL, RATE>& sm = *reinterpret_cast<Smem<ROLL, RATE>*>(smem_raw);
  const int t = threadIdx.x;
  const int e0 = t * E;
  const int64_t N = p.n_samples;
  const int64_t tile_a = (int64_t)blockIdx.x * tiles_per_cta;
  const int64_t tile_b = tile_a + tiles_per_cta < n_tiles ? tile_a + tiles_per_cta : n_tiles;
  if (tile_a >= tile_b) return;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const int need = p.min_periods > 1 ? p.min_periods : 1;
  constexpr bool do_roll = ROLL, do_rate = RATE;

  // series that contains the first flat index of the CTA's first tile
  int64_t sc, off_sc;
  {
    int64_t f = tile_a * OUT - HALO;
    if (f < 0) f = 0;
    int64_t lo = 0, hi = p.n_series + 1;                   // first index with off > f
    while (lo < hi) { const int64_t m = (lo + hi) >> 1; if (p.off[m] <= f) lo = m + 1; else hi = m; }
    sc = lo - 1;
    if (sc < 0) sc = 0;
    off_sc = p.off[sc];
  }

  for (int64_t tile = tile_a; tile < tile_b; ++tile) {
    const int64_t tile_lo = tile * OUT - HALO;             // flat index of tile element 0 (multiple of 4)
    const int64_t n_here = N - tile_lo < NE ? N - tile_lo : NE;   // elements of the tile that exist
    const int out_hi = (int)(n_here < NE ? n_here : NE);   // outputs are elements [HALO, out_hi)

    // ---- own samples, timestamps to shared memory
    double a[E]; int64_t tk[E];
#pragma unroll
    for (int r = 0; r < E; ++r) {
      const int64_t f = tile_lo + e0 + r;
      const bool in = f >= 0 && f < N;
      tk[r] = in ? p.ts[f] : 0x7fffffffffffffffLL;
      a[r] = in ? p.val[f] : pmx_nan();
      sm.ts[pad8(e0 + r)] = tk[r];
    }
    // ---- series starts inside the tile: B[i] = off[sc+1+i] - tile_lo, until one is >= NE
    {
      int nb = 0;
      for (;;) {
        const int i = nb + t;
        const int64_t s = sc + 1 + i;
        int64_t v = s <= p.n_series ? p.off[s] - tile_lo : (int64_t)NE;
        if (v > NE) v = NE;
        if (v < 0) v = 0;                                    // empty series before the tile start
        sm.B[i] = (uint16_t)v;
        __syncthreads();
        const bool done = sm.B[nb + NT - 1] >= NE;
        nb += NT;
        if (done || nb + NT > BMAX) { if (t == 0) { sm.nb = nb; sm.overflow = done ? 0 : 1; } break; }
      }
    }
    __syncthreads();
    const int nb = sm.nb;
    const bool tile_slow = sm.overflow != 0;                 // more series starts than fit: direct folds
    if (t == 0) {
      // series containing the next tile's first element (tile_lo + OUT): last start <= OUT
      if (tile_slow) {                                       // the staged list is incomplete: ask global memory
        const int64_t f = tile_lo + OUT;
        int64_t lo = sc, hi = p.n_series + 1;
        while (lo < hi) { const int64_t m = (lo + hi) >> 1; if (p.off[m] <= f) lo = m + 1; else hi = m; }
        sm.next_sc = lo - 1;
        sm.next_off_sc = p.off[lo - 1];
      } else {
        int lo = 0, hi = nb;
        while (lo < hi) { const int m = (lo + hi) >> 1; if (sm.B[m] <= OUT) lo = m + 1; else hi = m; }
        sm.next_sc = sc + lo;
        sm.next_off_sc = lo > 0 ? tile_lo + sm.B[lo - 1] : off_sc;
      }
    }

    // ---- per output: series start and window start
    int wl[E];                                             // window start (tile element index), -1 = take the direct fold
    int sr[E];                                             // series start, tile-relative (may be negative)
    {
      int j = 0;                                           // boundaries B[0..j) are <= current element
      { int lo = 0, hi = nb; while (lo < hi) { const int m = (lo + hi) >> 1; if (sm.B[m] <= e0) lo = m + 1; else hi = m; } j = lo; }
      int srel = j > 0 ? sm.B[j - 1] : (int)(off_sc - tile_lo < -0x40000000LL ? -0x40000000LL : off_sc - tile_lo);
      int l;
      {                                                    // first element: binary search inside the series
        const int64_t edge = tk[0] - p.window_ms;
        int lo = srel > 0 ? srel : 0, hi = e0;
        while (lo < hi) { const int m = (lo + hi) >> 1; if (sm.ts[pad8(m)] <= edge) lo = m + 1; else hi = m; }
        l = lo;
      }
#pragma unroll
      for (int r = 0; r < E; ++r) {
        const int e = e0 + r;
        while (j < nb && sm.B[j] <= e) { srel = sm.B[j]; ++j; }
        const int lob = srel > 0 ? srel : 0;
        if (l < lob) l = lob;
        const int64_t edge = tk[r] - p.window_ms;
        if (l < e && sm.ts[pad8(l)] <= edge) {               // usually one step; after a scrape gap or a
          ++l;                                               // series start it jumps: binary search then
          if (l < e && sm.ts[pad8(l)] <= edge) {
            int lo = l + 1, hi = e;
            while (lo < hi) { const int m = (lo + hi) >> 1; if (sm.ts[pad8(m)] <= edge) lo = m + 1; else hi = m; }
            l = lo;
          }
        }
        sr[r] = srel;
        // the series continues before the tile and the window still covers element 0: the halo
        // is too short for this output
        const bool short_halo = l == 0 && srel < 0 && tile_lo > 0 && sm.ts[0] > edge;
        wl[r] = (short_halo || tile_slow) ? -1 : l;
      }
    }

    // ---- in-chunk suffix scans
    double c_fv; int c_cn;
    {
      double ss = 0.0, smn = INF, smx = -INF, sd = 0.0, sfv = 0.0, lv = 0.0;
      int scf = 0, scn = 0, foff = 0, lpos = 0;
#pragma unroll
      for (int r = E - 1; r >= 0; --r) {
        const double v = a[r];
        if (pmx_isfinite(v)) { ss += v; smn = dmin(smn, v); smx = dmax(smx, v); ++scf; }
        if (v == v) {
          if (scn > 0) sd += delta_of(sfv, v); else { lv = v; lpos = r; }
          sfv = v; foff = r; ++scn;
        }
        const int si = r * NT + t;
        if (ROLL) { sm.S_sum[si] = ss; sm.S_mn[si] = smn; sm.S_mx[si] = smx; }
        if (RATE) sm.S_d[si] = sd;
        sm.S_meta[si] = (uint16_t)(scf | (scn << 4) | (((foff - r) & 7) << 8));
      }
      if (ROLL) { sm.T_sum[t] = ss; sm.T_mn[t] = smn; sm.T_mx[t] = smx; }
      if (RATE) { sm.T_fv[t] = sfv; sm.T_lv[t] = lv; sm.T_di[t] = sd; }
      sm.T_meta[t] = (uint32_t)(scf | (scn << 4) | (foff << 8) | (lpos << 12));
      c_fv = sfv; c_cn = scn;
    }
    __syncthreads();
    // ---- junction delta of the own chunk with the last valid sample before it (any chunk back to
    // the halo); published so that chunk ranges can be summed without order bookkeeping
    double clv = 0.0; int clpos = -1;
    double Jc = 0.0;
    if (RATE) {
      for (int k = t - 1; k >= 0; --k) {
        const uint32_t m = sm.T_meta[k];
        if ((m >> 4) & 0xf) { clv = sm.T_lv[k]; clpos = k * E + (int)((m >> 12) & 7); break; }
      }
      Jc = (c_cn > 0 && clpos >= 0) ? delta_of(c_fv, clv) : 0.0;
      sm.T_tj[t] = sm.T_di[t] + Jc;
      __syncthreads();
    }

    const int clb = wl[E - 1] >= 0 ? (wl[E - 1] >> 3) : t;   // chunk of the last output's window start
    const int64_t obase = tile_lo + e0;                    // flat index of the thread's first output
    // ---------------------------------------------------------------- rolling statistics
    if (do_roll && e0 + E > HALO && e0 < out_hi) {
      double ms = 0.0, mmn = INF, mmx = -INF; int mcf = 0, mcn = 0;      // chunks clb+1 .. t-1
      for (int k = clb + 1; k < t; ++k) {
        const uint32_t m = sm.T_meta[k];
        ms += sm.T_sum[k]; mmn = dmin(mmn, sm.T_mn[k]); mmx = dmax(mmx, sm.T_mx[k]);
        mcf += m & 0xf; mcn += (m >> 4) & 0xf;
      }
      double ps = 0.0, pmn = INF, pmx = -INF; int pcf = 0, pcn = 0;
      double omn[4], omx[4], omean[4]; int ocnt[4];
#pragma unroll
      for (int r = 0; r < E; ++r) {
        const double v = a[r];
        if (pmx_isfinite(v)) { ps += v; pmn = dmin(pmn, v); pmx = dmax(pmx, v); ++pcf; }
        pcn += (v == v) ? 1 : 0;
        const int l = wl[r];
        double ws, wmn, wmx; int wcf, wcn;
        if (l < 0) {                                       // direct fold from global memory
          const Folded fd = fold_direct(p.ts, p.val, tile_lo + sr[r], obase + r, p.window_ms, need, 0);
          omn[r & 3] = fd.v[0]; omx[r & 3] = fd.v[1]; omean[r & 3] = fd.v[2];
          ocnt[r & 3] = (int)fd.v[3];
        } else {
        const int cl = l >> 3;
        if (cl == t) {                                     // window inside the own chunk
          ws = 0.0; wmn = INF; wmx = -INF; wcf = 0; wcn = 0;
#pragma unroll
          for (int rr = 0; rr < E; ++rr) {
            if (rr <= r && rr >= (l & 7)) {
              const double u = a[rr];
              if (pmx_isfinite(u)) { ws += u; wmn = dmin(wmn, u); wmx = dmax(wmx, u); ++wcf; }
              wcn += (u == u) ? 1 : 0;
            }
          }
        } else {
          const int si = (l & 7) * NT + cl;
          const uint32_t m0 = sm.S_meta[si];
          ws = sm.S_sum[si]; wmn = sm.S_mn[si]; wmx = sm.S_mx[si]; wcf = m0 & 0xf; wcn = (m0 >> 4) & 0xf;
          for (int k = cl + 1; k <= clb && k < t; ++k) {    // chunks this output has on top of the shared fold
            const uint32_t m = sm.T_meta[k];
            ws += sm.T_sum[k]; wmn = dmin(wmn, sm.T_mn[k]); wmx = dmax(wmx, sm.T_mx[k]);
            wcf += m & 0xf; wcn += (m >> 4) & 0xf;
          }
          ws += ms; wmn = dmin(wmn, mmn); wmx = dmax(wmx, mmx); wcf += mcf; wcn += mcn;
          ws += ps; wmn = dmin(wmn, pmn); wmx = dmax(wmx, pmx); wcf += pcf; wcn += pcn;
        }
        const bool ok = wcf >= need;
        omean[r & 3] = ok ? ws / (double)wcf : pmx_nan();
        omn[r & 3] = ok ? wmn : pmx_nan();
        omx[r & 3] = ok ? wmx : pmx_nan();
        ocnt[r & 3] = wcn;
        }
        if ((r & 3) == 3) {                                // four outputs ready: one 256-bit store per array
          const int e = e0 + r - 3;
          const int64_t o = obase + r - 3;
          if (e >= HALO && e + 3 < out_hi) {
            if (p.o_min) st4(p.o_min + o, omn[0], omn[1], omn[2], omn[3]);
            if (p.o_max) st4(p.o_max + o, omx[0], omx[1], omx[2], omx[3]);
            if (p.o_mean) st4(p.o_mean + o, omean[0], omean[1], omean[2], omean[3]);
            if (p.o_cnt) *reinterpret_cast<int4*>(p.o_cnt + o) = make_int4(ocnt[0], ocnt[1], ocnt[2], ocnt[3]);
          } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              if (e + q >= HALO && e + q < out_hi) {
                if (p.o_min) p.o_min[o + q] = omn[q];
                if (p.o_max) p.o_max[o + q] = omx[q];
                if (p.o_mean) p.o_mean[o + q] = omean[q];
                if (p.o_cnt) p.o_cnt[o + q] = ocnt[q];
              }
            }
          }
        }
      }
    }
    // ---------------------------------------------------------------- increase / rate
    if (do_rate && e0 + E > HALO && e0 < out_hi) {
      double md = 0.0;                                     // chunks clb+1 .. t-1, junctions included
      for (int k = clb + 1; k < t; ++k) md += sm.T_tj[k];
      double pd = 0.0, plv = 0.0; int pcn = 0, plpos = 0;
      double oinc[4], orate[4];
#pragma unroll
      for (int r = 0; r < E; ++r) {
        const double v = a[r];
        if (v == v) {
          if (pcn > 0) pd += delta_of(v, plv);
          plv = v; plpos = e0 + r; ++pcn;
        }
        const int l = wl[r];
        const int cl = l >> 3;
        int fpos = 0x7fffffff, lpos = -1; double inc = 0.0;
        bool direct = l < 0;
        if (!direct && cl == t) {                          // window inside the own chunk: ordered fold
          int cnt = 0; double lv = 0.0;
#pragma unroll
          for (int rr = 0; rr < E; ++rr) {
            if (rr <= r && rr >= (l & 7)) {
              const double u = a[rr];
              if (u == u) {
                if (cnt > 0) inc += delta_of(u, lv); else fpos = e0 + rr;
                lv = u; lpos = e0 + rr; ++cnt;
              }
            }
          }
        } else if (!direct) {
          const int si = (l & 7) * NT + cl;
          const uint32_t m0 = sm.S_meta[si];
          if ((m0 >> 4) & 0xf) {                           // first chunk part holds a sample
            fpos = l + (int)((m0 >> 8) & 7);
            inc = sm.S_d[si];
            for (int k = cl + 1; k <= clb && k < t; ++k) inc += sm.T_tj[k];
            inc += md + pd + (pcn > 0 ? Jc : 0.0);
            lpos = pcn > 0 ? plpos : clpos;
          } else direct = true;                            // NaN run at the window start
        }
        if (direct) {
          const Folded fd = fold_direct(p.ts, p.val, tile_lo + sr[r], obase + r, p.window_ms, need, 1);
          oinc[r & 3] = fd.v[0]; orate[r & 3] = fd.v[1];
        } else {
          const bool have = lpos > fpos;
          double rate = pmx_nan();
          if (have) {
            const int64_t dt = sm.ts[pad8(lpos)] - sm.ts[pad8(fpos)];
            rate = dt > 0 ? inc * 1000.0 / (double)dt : pmx_nan();
          }
          oinc[r & 3] = have ? inc : pmx_nan();
          orate[r & 3] = rate;
        }
        if ((r & 3) == 3) {
          const int e = e0 + r - 3;
          const int64_t o = obase + r - 3;
          if (e >= HALO && e + 3 < out_hi) {
            if (p.o_inc) st4(p.o_inc + o, oinc[0], oinc[1], oinc[2], oinc[3]);
            if (p.o_rate) st4(p.o_rate + o, orate[0], orate[1], orate[2], orate[3]);
          } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              if (e + q >= HALO && e + q < out_hi) {
                if (p.o_inc) p.o_inc[o + q] = oinc[q];
                if (p.o_rate) p.o_rate[o + q] = orate[q];
              }
            }
          }
        }
      }
    }
    __syncthreads();                                         // shared arrays free for the next tile
    sc = sm.next_sc;
    off_sc = sm.next_off_sc;
    __syncthreads();
  }
}

template <bool ROLL, bool RATE>
int launch_as(const RParams& p, int64_t n_tiles, int64_t grid, int64_t tpc, cudaStream_t st) {
  PMX_CUDA(cudaFuncSetAttribute(ragged_fast_kernel<ROLL, RATE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)sizeof(Smem<ROLL, RATE>)));
  ragged_fast_kernel<ROLL, RATE><<<(unsigned)grid, NT, sizeof(Smem<ROLL, RATE>), st>>>(p, n_tiles, (int)tpc);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; }

}  // namespace

int launch_ragged_fast(const pmx_ctx* ctx, const RParams& p, cudaStream_t st) {
  if (p.o_nres || p.rate_mode != 0) return 1;
  if (!aligned32(p.o_min) || !aligned32(p.o_max) || !aligned32(p.o_mean) || !aligned32(p.o_inc) ||
      !aligned32(p.o_rate) || (reinterpret_cast<uintptr_t>(p.o_cnt) & 15))
    return 1;
  if (p.n_samples < 1) return PMX_OK;
  const int64_t n_tiles = (p.n_samples + OUT - 1) / OUT;
  int64_t ctas = (int64_t)ctx->sm_count * 16;
  if (ctas > n_tiles) ctas = n_tiles;
  const int64_t tpc = (n_tiles + ctas - 1) / ctas;
  const int64_t grid = (n_tiles + tpc - 1) / tpc;
  const bool roll = p.o_min || p.o_max || p.o_mean || p.o_cnt, rate = p.o_inc || p.o_rate;
  if (roll && rate) return launch_as<true, true>(p, n_tiles, grid, tpc, st);
  if (roll) return launch_as<true, false>(p, n_tiles, grid, tpc, st);
  if (rate) return launch_as<false, true>(p, n_tiles, grid, tpc, st);
  return PMX_OK;
}

}  // namespace pmxr
