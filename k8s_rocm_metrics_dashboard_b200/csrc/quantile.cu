// quantile.cu -- K6: exact quantiles by radix select on order-preserving keys.
//
// float64 -> uint64 key that sorts like the number (NaN excluded), then PMX_QUANTILE_PASSES
// passes of 11-bit digits, most significant first.  A pass is
//   hist : per-CTA shared-memory histograms (one per target still being refined) of the digit of
//          every key that matches the target's prefix, flushed with integer atomics -- so the
//          global histogram is deterministic and can be all-reduced across ranks as is;
//   pick : prefix-scan the (reduced) histogram on the device, choose the bin that holds the
//          target rank, append it to the prefix, subtract the ranks below it.
// After the last pass every target's prefix IS the key of the wanted order statistic.  Targets
// per segment: for each q the ranks lo = floor((n-1)q) and hi = min(lo+1, n-1), then numpy's
// _lerp.  No sample ever moves between GPUs; a pass exchanges 16 KiB per target.
// The passes before the list pass (2, or 3 for very long segments) histogram every value (with
// pmx_quantile_init there is no pass 0: keys are taken relative to the segment's minimum and the
// bits its range leaves zero are skipped).  The list pass reads the data once more, only to copy the
// keys that still match a target prefix into a per-segment candidate list (collect_kernel: no
// histogram, the speed of the read); its digit and all later ones come from that list -- the rest at
// once in pmx_quantile_tail (one GPU) or after an all-gather of the few keys left
// (pmx_quantile_tail_filter / _merge).  A segment whose list overflows (many repeated values)
// keeps reading the full data in the per-pass kernels.
//
// State per segment (uint64 words): [0] n, [1] how the selection began and whether it is finished,
// [2] base key, then per target t: [3+2t] prefix, [4+2t] remaining rank inside the prefix bucket.
#include <stdlib.h>
#include "pmx_common.cuh"
#include "quantile_keys.cuh"

namespace {

constexpr int BINS = PMX_QUANTILE_BINS;
constexpr int MAX_TARGETS = 16;

__host__ __device__ __forceinline__ int digit_shift(int pass) { return pass < 5 ? 53 - 11 * pass : 0; }
__host__ __device__ __forceinline__ int digit_bits(int pass) { return pass < 5 ? 11 : 9; }

// State of one segment: [0] number of values, [1] 0 for a selection that began with pass 0, or
// (1 << 32) | Z for one that began from the segment's count / min / max (pmx_quantile_init),
// [2] base key (the key of the minimum; 0 otherwise), then (prefix, remaining rank) per target.
// A selection from count/min/max works on key - base, whose leading Z bits are zero for every
// value of the segment, and starts behind them: pass p >= 1 takes bits [Z + 11(p-1), +11).
constexpr int ST_HDR = 3;
constexpr unsigned TAIL_KEYS = 8192;     // keys pmx_quantile_tail keeps in shared memory per segment
constexpr unsigned GATHER_ROW = PMX_QUANTILE_GATHER_ROW;   // per segment and rank: count, then up to ROW-1 keys
constexpr unsigned GATHER_KEYS = GATHER_ROW - 1 - 2 * MAX_TARGETS;   // ... and, at the row's end, min and max key per target
constexpr int ST_RESOLVED_BIT = 40;      // in word [1]: every target's key is complete
#define ST_PF(tg) (ST_HDR + 2 * (tg))
#define ST_RK(tg) (ST_HDR + 2 * (tg) + 1)
__host__ __device__ __forceinline__ int state_words(int64_t n_q) { return ST_HDR + 4 * (int)n_q; }

// all_match: every value of the segment carries every target's prefix (no prefix yet, or only the
// leading zero bits of a selection started from count/min/max), so nothing has to be compared.
struct Geom { int shift; int bits; int pshift; bool has_prefix; bool all_match; uint64_t base; };
__device__ __forceinline__ Geom seg_geom(const uint64_t* st, int pass) {
  Geom g;
  const uint64_t w1 = st[1];
  if ((w1 >> 32) & 1) {
    const int start = (int)(w1 & 0xff) + 11 * (pass - 1);         // key bits known before this pass
    int bits = 64 - start;
    if (bits > 11) bits = 11;
    if (bits < 0 || pass < 1) bits = 0;                           // 0: nothing left to resolve
    g.bits = bits; g.pshift = 64 - start; g.shift = 64 - start - bits;
    if (g.shift < 0) g.shift = 0;
    g.has_prefix = start > 0; g.base = st[2]; g.all_match = pass == 1;
  } else {
    g.bits = pass < PMX_QUANTILE_PASSES ? digit_bits(pass) : 0;
    g.shift = digit_shift(pass < 5 ? pass : 5);
    g.pshift = 64 - 11 * pass; g.has_prefix = pass > 0; g.base = 0; g.all_match = pass == 0;
  }
  if ((w1 >> ST_RESOLVED_BIT) & 1) g.bits = 0;                  // pmx_quantile_tail finished this segment
  return g;
}

// Adds one key per lane to a CTA-private histogram: a plain shared-memory atomic.  The hardware
// folds the lanes of a warp that hit one address (ATOMS.POPC.INC), and on B200 that beats every
// warp-level folding scheme (leader rounds, match_any, run merging) on spread, repeated and
// constant digits alike -- tools/micro/hist_variants.cu: 59 / 47 / 47 us for 35 M values against
// 59 / 70 / 70 us for two leader rounds.
__device__ __forceinline__ void hist_add(unsigned int* h, bool hit, unsigned d) {
  if (hit) atomicAdd(&h[d], 1u);
}

// distinct prefixes of a segment's targets (lo/hi ranks and nearby quantiles usually share one)
// s_urep[u] = first target that carries distinct prefix u: the kernels accumulate into that
// target's global histogram only; pick_kernel reads that one for every target of u.
__device__ __forceinline__ void load_prefixes(const uint64_t* __restrict__ state, int sw, int sg, int n_t,
                                              int pass, uint64_t* s_uprefix, int* s_umap, int* s_urep,
                                              int* s_nu) {
  int nu = 0;
  if (pass == 0) { nu = 1; s_urep[0] = 0; }
  else for (int k = 0; k < n_t; ++k) {
    const uint64_t pf = state[(size_t)sg * sw + ST_PF(k)];
    int u = 0;
    while (u < nu && s_uprefix[u] != pf) ++u;
    if (u == nu) { s_uprefix[nu] = pf; s_urep[nu] = k; ++nu; }
    s_umap[k] = u;
  }
  *s_nu = nu;
}

// mode 0: histogram of every row; mode 2 (the list pass and later): only segments whose candidate
// list overflowed -- cand_hist_kernel serves the others from their lists.
template <int MODE>
__global__ void __launch_bounds__(256)
hist_kernel(const double* __restrict__ x, int64_t n_steps, const int32_t* __restrict__ seg,
            int n_t, int pass, const uint64_t* __restrict__ state, int sw,
            unsigned long long* __restrict__ hist, int rows_per_cta, int64_t n_rows,
            const int32_t* __restrict__ order,
            unsigned long long* __restrict__ cand, unsigned int* __restrict__ cand_cnt, unsigned cap,
            unsigned int* __restrict__ cand_overflow) {
  constexpr int mode = MODE;
  extern __shared__ unsigned int sh[];                     // [n_uniq][BINS]
  __shared__ uint64_t s_uprefix[MAX_TARGETS];
  __shared__ int s_umap[MAX_TARGETS];
  __shared__ int s_urep[MAX_TARGETS];
  __shared__ int s_nu;
  __shared__ unsigned int s_bitmap[BINS / 32];             // previous digits carried by some target prefix
  __shared__ Geom s_geom;                                  // where this pass looks in the current segment's keys
  const int64_t s0 = (int64_t)blockIdx.x * rows_per_cta;
  const int n_h = pass == 0 ? 1 : n_t;                     // histograms per segment in global memory
  int cur_seg = -1;
  if (mode == 2 && *cand_overflow == 0) return;            // every candidate list is complete (the usual case)

  auto flush = [&](int sg) {
    __syncthreads();
    for (int u = 0; u < s_nu; ++u) {
      const unsigned int* src = sh + u * BINS;
      unsigned long long* dst = hist + ((size_t)sg * n_h + s_urep[u]) * BINS;
      for (int i = threadIdx.x; i < BINS; i += blockDim.x) {
        const unsigned c = src[i];
        if (c) atomicAdd(&dst[i], (unsigned long long)c);
      }
    }
  };

  for (int r = 0; r < rows_per_cta; ++r) {
    if (s0 + r >= n_rows) break;
    const int64_t s = order ? (int64_t)order[s0 + r] : s0 + r;   // rows grouped by segment when given
    const int sg = seg[s];
    if (sg < 0) continue;
    if (mode == 2 && cand_cnt[sg] <= cap) continue;        // complete candidate list: not our job
    if (sg != cur_seg) {                                   // flush, switch segment
      if (cur_seg >= 0) flush(cur_seg);
      __syncthreads();
      if (threadIdx.x < BINS / 32) s_bitmap[threadIdx.x] = 0;
      if (threadIdx.x == 0) {
        load_prefixes(state, sw, sg, n_t, pass, s_uprefix, s_umap, s_urep, &s_nu);
        s_geom = seg_geom(state + (size_t)sg * sw, pass);
      }
      __syncthreads();
      if (pass > 0 && threadIdx.x < s_nu) {
        const unsigned pd = (unsigned)s_uprefix[threadIdx.x] & (BINS - 1);
        atomicOr(&s_bitmap[pd >> 5], 1u << (pd & 31));
      }
      for (int i = threadIdx.x; i < s_nu * BINS; i += blockDim.x) sh[i] = 0;
      cur_seg = sg;
      __syncthreads();
    }
    const int nu = s_nu;
    const Geom gm = s_geom;
    if (gm.bits == 0) continue;                            // this segment's keys are fully resolved
    const int shift = gm.shift, pshift = gm.pshift, bits = gm.bits;
    const unsigned mask = (1u << bits) - 1u;
    const uint64_t kbase = gm.base;
    const bool all_match = gm.all_match;
    const bool use_bitmap = pass >= 2;                     // the previous digit exists and is 11 bits wide
    // previous digits of up to four distinct prefixes in registers (the usual case: lo/hi ranks and
    // neighbouring quantiles share prefixes); more go through the shared bitmap
    unsigned pd0 = 0xffffffffu, pd1 = 0xffffffffu, pd2 = 0xffffffffu, pd3 = 0xffffffffu;
    if (nu > 0) pd0 = (unsigned)s_uprefix[0] & (BINS - 1);
    if (nu > 1) pd1 = (unsigned)s_uprefix[1] & (BINS - 1);
    if (nu > 2) pd2 = (unsigned)s_uprefix[2] & (BINS - 1);
    if (nu > 3) pd3 = (unsigned)s_uprefix[3] & (BINS - 1);
    const bool few = nu <= 4;
    const unsigned kb_lo = (unsigned)kbase, kb_hi = (unsigned)(kbase >> 32);
    const double* xr = x + s * n_steps;
    const bool vec = (reinterpret_cast<uintptr_t>(xr) & 15) == 0;
    // Trips are warp-uniform: every lane works on four elements per trip (two 16-byte loads, the
    // second half a CTA-width further) while the next trip's four are already on their way.  Full
    // trips need no bounds checks; the last, partial one is loaded element by element.
    const int64_t per_trip = (int64_t)blockDim.x * 4;
    const int64_t n_full = vec ? n_steps / per_trip : 0;
    const int64_t n_trips = (n_steps + per_trip - 1) / per_trip;
    const double qnan = __longlong_as_double(0x7FF8000000000000LL);
    auto load_tail = [&](int64_t j0, double* v) {
      const int64_t ja = j0 + 2 * threadIdx.x, jb = ja + 2 * blockDim.x;
      v[0] = ja < n_steps ? xr[ja] : qnan; v[1] = ja + 1 < n_steps ? xr[ja + 1] : qnan;
      v[2] = jb < n_steps ? xr[jb] : qnan; v[3] = jb + 1 < n_steps ? xr[jb + 1] : qnan;
    };
    const double2* pa = reinterpret_cast<const double2*>(xr) + threadIdx.x;     // first half of this lane's trip
    const int64_t half = blockDim.x, stride = 2 * (int64_t)blockDim.x;          // in double2 units
    double vn[4];
    if (n_full > 0) { const double2 a = pa[0], b = pa[half]; vn[0] = a.x; vn[1] = a.y; vn[2] = b.x; vn[3] = b.y; }
    else load_tail(0, vn);
    for (int64_t trip = 0; trip < n_trips; ++trip) {
      double v[4] = {vn[0], vn[1], vn[2], vn[3]};
      if (trip + 1 < n_full) {
        pa += stride;
        const double2 a = pa[0], b = pa[half];
        vn[0] = a.x; vn[1] = a.y; vn[2] = b.x; vn[3] = b.y;
      } else if (trip + 1 < n_trips) {
        load_tail((trip + 1) * per_trip, vn);
      }
      // order-preserving key relative to the segment's base key, in 32-bit halves:
      // key = bits ^ (sign ? ~0 : msb), then key - kbase with borrow; w = 32-bit view of the key
      // from bit `shift` up: the digit and, above it, the previous digit
      unsigned r_lo[4], r_hi[4], w[4];
      bool valid[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        valid[i] = v[i] == v[i];
        const unsigned b_lo = (unsigned)__double2loint(v[i]), b_hi = (unsigned)__double2hiint(v[i]);
        const unsigned sm = (unsigned)((int)b_hi >> 31);
        const unsigned k_lo = b_lo ^ sm, k_hi = b_hi ^ (sm | 0x80000000u);
        r_lo[i] = k_lo - kb_lo; r_hi[i] = k_hi - kb_hi - (k_lo < kb_lo ? 1u : 0u);
        w[i] = shift < 32 ? __funnelshift_r(r_lo[i], r_hi[i], shift) : r_hi[i] >> (shift - 32);
      }
      if (all_match) {
#pragma unroll
        for (int i = 0; i < 4; ++i) hist_add(sh, valid[i], w[i] & mask);
        continue;
      }
      // cheap filter first, for the four elements at once: does any target prefix end in the key's
      // previous digit?  From pass 2 on hardly any key passes and the whole trip is skipped.
      bool maybe[4];
      bool some = false;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        maybe[i] = valid[i];
        if (use_bitmap) {
          const unsigned pdig = (w[i] >> bits) & (BINS - 1);
          if (few) maybe[i] = valid[i] & ((pdig == pd0) | (pdig == pd1) | (pdig == pd2) | (pdig == pd3));
          else { const unsigned bm = s_bitmap[pdig >> 5]; maybe[i] = valid[i] & ((bm >> (pdig & 31)) & 1u); }
        }
        some |= maybe[i];
      }
      if (__ballot_sync(0xffffffffu, some) == 0) continue;
#pragma unroll 1
      for (int i = 0; i < 4; ++i) {
        const bool mb = i == 0 ? maybe[0] : i == 1 ? maybe[1] : i == 2 ? maybe[2] : maybe[3];
        if (__ballot_sync(0xffffffffu, mb) == 0) continue;
        const unsigned lo = i == 0 ? r_lo[0] : i == 1 ? r_lo[1] : i == 2 ? r_lo[2] : r_lo[3];
        const unsigned hi = i == 0 ? r_hi[0] : i == 1 ? r_hi[1] : i == 2 ? r_hi[2] : r_hi[3];
        const unsigned wi = i == 0 ? w[0] : i == 1 ? w[1] : i == 2 ? w[2] : w[3];
        const unsigned d = wi & mask;
        const uint64_t key = ((uint64_t)hi << 32) | lo;
        const uint64_t top = key >> pshift;                // has_prefix holds here: pshift < 64
        unsigned hits = 0;                                 // bit u: this lane's key carries distinct prefix u
        for (int u = 0; u < nu; ++u) hits |= (mb && top == s_uprefix[u]) ? (1u << u) : 0u;
        const unsigned any = __ballot_sync(0xffffffffu, hits != 0);
        if (any == 0) continue;
        for (int u = 0; u < nu; ++u) hist_add(sh + u * BINS, (hits >> u) & 1, d);
      }
    }
  }
  if (cur_seg >= 0) flush(cur_seg);
}

// The list pass without a histogram: every key that carries a target's prefix (as the passes before
// `pass` left it) is appended to its segment's candidate list, and nothing else happens -- no
// shared-memory histograms, so twice the CTAs fit an SM and the sweep runs at the speed of the
// read.  cand_hist_kernel then takes this pass's digit from the lists.  Hits are staged per warp in
// shared memory and leave as one global atomic + coalesced stores per ~200 keys.
constexpr int COLLECT_WBUF = 256;                            // staged keys per warp; a trip adds at most 128
__global__ void __launch_bounds__(256)
collect_kernel(const double* __restrict__ x, int64_t n_steps, const int32_t* __restrict__ seg,
               int n_t, int pass, const uint64_t* __restrict__ state, int sw, int rows_per_cta, int64_t n_rows,
               const int32_t* __restrict__ order, unsigned long long* __restrict__ cand,
               unsigned int* __restrict__ cand_cnt, unsigned cap, unsigned int* __restrict__ cand_overflow) {
  __shared__ uint64_t s_uprefix[MAX_TARGETS];
  __shared__ int s_umap[MAX_TARGETS];
  __shared__ int s_urep[MAX_TARGETS];
  __shared__ int s_nu;
  __shared__ Geom s_geom;
  __shared__ unsigned long long s_buf[8][COLLECT_WBUF];
  const int64_t s0 = (int64_t)blockIdx.x * rows_per_cta;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long* wbuf = s_buf[warp];
  unsigned wcnt = 0;                                         // keys staged by this warp (warp-uniform)
  int cur_seg = -1;
  auto flush = [&](int sg) {
    if (wcnt == 0) return;
    unsigned base = 0;
    if (lane == 0) {
      base = atomicAdd(&cand_cnt[sg], wcnt);
      if (base + wcnt > cap) *cand_overflow = 1u;
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    __syncwarp();
    for (unsigned i = lane; i < wcnt; i += 32)
      if (base + i < cap) cand[(size_t)sg * cap + base + i] = wbuf[i];
    __syncwarp();
    wcnt = 0;
  };
  for (int r = 0; r < rows_per_cta; ++r) {
    if (s0 + r >= n_rows) break;
    const int64_t s = order ? (int64_t)order[s0 + r] : s0 + r;
    const int sg = seg[s];
    if (sg < 0) continue;
    if (sg != cur_seg) {
      if (cur_seg >= 0) flush(cur_seg);
      __syncthreads();
      if (threadIdx.x == 0) {
        load_prefixes(state, sw, sg, n_t, pass, s_uprefix, s_umap, s_urep, &s_nu);
        s_geom = seg_geom(state + (size_t)sg * sw, pass);
      }
      cur_seg = sg;
      __syncthreads();
    }
    const int nu = s_nu;
    const Geom gm = s_geom;
    if (gm.bits == 0 || !gm.has_prefix) continue;            // nothing left to resolve / no prefix to match yet
    const int pshift = gm.pshift;
    const unsigned kb_lo = (unsigned)gm.base, kb_hi = (unsigned)(gm.base >> 32);
    // up to four distinct prefixes in registers (lo/hi ranks and neighbouring quantiles share them);
    // ~0 never equals the top bits of a key
    const bool few = nu <= 4;
    const uint64_t p0 = nu > 0 ? s_uprefix[0] : ~0ULL, p1 = nu > 1 ? s_uprefix[1] : ~0ULL,
                   p2 = nu > 2 ? s_uprefix[2] : ~0ULL, p3 = nu > 3 ? s_uprefix[3] : ~0ULL;
    const bool top32 = pshift >= 32;                         // the known bits sit in the key's high word
    const unsigned q0 = (unsigned)p0, q1 = (unsigned)p1, q2 = (unsigned)p2, q3 = (unsigned)p3;
    const double* xr = x + s * n_steps;
    const bool vec = (reinterpret_cast<uintptr_t>(xr) & 15) == 0;
    const int64_t per_trip = (int64_t)blockDim.x * 4;
    const int64_t n_full = vec ? n_steps / per_trip : 0;
    const int64_t n_trips = (n_steps + per_trip - 1) / per_trip;
    const double qnan = __longlong_as_double(0x7FF8000000000000LL);
    auto load_tail = [&](int64_t j0, double* v) {
      const int64_t ja = j0 + 2 * threadIdx.x, jb = ja + 2 * blockDim.x;
      v[0] = ja < n_steps ? xr[ja] : qnan; v[1] = ja + 1 < n_steps ? xr[ja + 1] : qnan;
      v[2] = jb < n_steps ? xr[jb] : qnan; v[3] = jb + 1 < n_steps ? xr[jb + 1] : qnan;
    };
    const double2* pa = reinterpret_cast<const double2*>(xr) + threadIdx.x;
    const int64_t half = blockDim.x, stride = 2 * (int64_t)blockDim.x;
    double vn[4];
    if (n_full > 0) { const double2 a = pa[0], b = pa[half]; vn[0] = a.x; vn[1] = a.y; vn[2] = b.x; vn[3] = b.y; }
    else load_tail(0, vn);
    for (int64_t trip = 0; trip < n_trips; ++trip) {
      double v[4] = {vn[0], vn[1], vn[2], vn[3]};
      if (trip + 1 < n_full) {
        pa += stride;
        const double2 a = pa[0], b = pa[half];
        vn[0] = a.x; vn[1] = a.y; vn[2] = b.x; vn[3] = b.y;
      } else if (trip + 1 < n_trips) {
        load_tail((trip + 1) * per_trip, vn);
      }
      unsigned r_lo[4], r_hi[4];
      bool hit[4];
      bool some = false;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const unsigned b_lo = (unsigned)__double2loint(v[i]), b_hi = (unsigned)__double2hiint(v[i]);
        const unsigned sm = (unsigned)((int)b_hi >> 31);
        const unsigned k_lo = b_lo ^ sm, k_hi = b_hi ^ (sm | 0x80000000u);
        r_lo[i] = k_lo - kb_lo; r_hi[i] = k_hi - kb_hi - (k_lo < kb_lo ? 1u : 0u);
        bool h;
        if (top32 && few) {
          const unsigned top = r_hi[i] >> (pshift - 32);
          h = (top == q0) | (top == q1) | (top == q2) | (top == q3);
        } else {
          const uint64_t top = (((uint64_t)r_hi[i] << 32) | r_lo[i]) >> pshift;
          if (few) h = (top == p0) | (top == p1) | (top == p2) | (top == p3);
          else { h = false; for (int u = 0; u < nu; ++u) h |= top == s_uprefix[u]; }
        }
        hit[i] = h & (v[i] == v[i]);
        some |= hit[i];
      }
      if (__ballot_sync(0xffffffffu, some) == 0) continue;   // the usual trip: no key of this warp matches
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const unsigned any = __ballot_sync(0xffffffffu, hit[i]);
        if (hit[i]) wbuf[wcnt + (unsigned)__popc(any & ((1u << lane) - 1u))] = ((uint64_t)r_hi[i] << 32) | r_lo[i];
        wcnt += (unsigned)__popc(any);
      }
      if (wcnt > COLLECT_WBUF - 128) flush(sg);
    }
  }
  if (cur_seg >= 0) flush(cur_seg);
}

// passes >= 3 over the candidate lists: grid (blocks, n_seg)
__global__ void __launch_bounds__(256)
cand_hist_kernel(const unsigned long long* __restrict__ cand, const unsigned int* __restrict__ cand_cnt,
                 unsigned cap, int n_t, int pass, const uint64_t* __restrict__ state, int sw,
                 unsigned long long* __restrict__ hist) {
  extern __shared__ unsigned int sh[];
  __shared__ uint64_t s_uprefix[MAX_TARGETS];
  __shared__ int s_umap[MAX_TARGETS];
  __shared__ int s_urep[MAX_TARGETS];
  __shared__ int s_nu;
  const int sg = blockIdx.y;
  const unsigned n = cand_cnt[sg];
  if (n > cap || n == 0) return;
  const unsigned per = (n + gridDim.x - 1) / gridDim.x;
  const unsigned a = blockIdx.x * per, b = a + per < n ? a + per : n;
  if (a >= b) return;
  const Geom gm = seg_geom(state + (size_t)blockIdx.y * sw, pass);
  if (gm.bits == 0) return;
  const int shift = gm.shift, pshift = gm.pshift;
  const unsigned mask = (1u << gm.bits) - 1u;
  if (threadIdx.x == 0) load_prefixes(state, sw, sg, n_t, pass, s_uprefix, s_umap, s_urep, &s_nu);
  __syncthreads();
  const int nu = s_nu;
  for (int i = threadIdx.x; i < nu * BINS; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const unsigned long long* keys = cand + (size_t)sg * cap;
  for (unsigned i0 = a; i0 < b; i0 += blockDim.x) {
    const unsigned i = i0 + threadIdx.x;
    const bool in = i < b;
    const uint64_t key = in ? keys[i] : 0;
    const unsigned d = (unsigned)(key >> shift) & mask;
    const uint64_t top = gm.has_prefix ? key >> pshift : 0;
    for (int u = 0; u < nu; ++u) hist_add(sh + u * BINS, in && top == s_uprefix[u], d);
  }
  __syncthreads();
  for (int u = 0; u < nu; ++u) {
    const unsigned int* src = sh + u * BINS;
    unsigned long long* dst = hist + ((size_t)sg * n_t + s_urep[u]) * BINS;
    for (int i = threadIdx.x; i < BINS; i += blockDim.x) {
      const unsigned c = src[i];
      if (c) atomicAdd(&dst[i], (unsigned long long)c);
    }
  }
}

// All remaining passes of one segment in one CTA, from its candidate list alone (single GPU: the
// list holds every key that still matches a target).  Segments whose list overflowed are left to
// the per-pass kernels; the others are marked resolved, and those kernels return at once for them.
//
// GATHERED = true: the keys come from the rows pmx_quantile_tail_filter wrote on every rank, after an
// all-gather ([n_ranks][n_seg][ROW], row = count then keys): each rank settles every segment from the
// union, identically.  A segment that does not fit (a rank kept more than GATHER_KEYS keys, or the
// union exceeds TAIL_KEYS) is counted in *unresolved and left to the per-pass protocol.
template <bool GATHERED>
__global__ void __launch_bounds__(256)
tail_kernel(const unsigned long long* __restrict__ cand, const unsigned int* __restrict__ cand_cnt,
            unsigned cap, int n_t, int from_pass, uint64_t* __restrict__ state, int sw,
            const unsigned long long* __restrict__ rows, int n_ranks, unsigned int* __restrict__ unresolved,
            const double* __restrict__ x, int64_t n_series, int64_t n_steps, const int32_t* __restrict__ seg) {
  extern __shared__ unsigned long long skeys[];            // TAIL_KEYS keys that still match a target
  __shared__ unsigned int sh[BINS];
  __shared__ uint64_t s_old[MAX_TARGETS], s_pf[MAX_TARGETS], s_k[MAX_TARGETS], s_nk[MAX_TARGETS];
  __shared__ unsigned int s_warp[8];
  __shared__ unsigned int s_cnt;
  // smallest and largest key that matches each target's prefix: equal means every remaining key of
  // that target is the same value (repeated readings), and the target is settled whatever its rank
  __shared__ unsigned long long s_tmin[MAX_TARGETS], s_tmax[MAX_TARGETS];
  __shared__ int s_done[MAX_TARGETS];
  const int sg = blockIdx.x;
  uint64_t* st = state + (size_t)sg * sw;
  const unsigned n_list = GATHERED ? 0u : cand_cnt[sg];
  if (st[0] == 0) return;                                   // empty segment
  // One GPU: a segment whose list overflowed (or whose surviving keys do not fit shared memory:
  // many repeated values) is finished from the data itself, by this one CTA -- slow, rare, and it
  // spares every selection the per-pass kernels.
  bool from_data = !GATHERED && n_list > cap;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < n_t) {
    s_pf[threadIdx.x] = st[ST_PF(threadIdx.x)];
    s_k[threadIdx.x] = s_nk[threadIdx.x] = st[ST_RK(threadIdx.x)];
    s_tmin[threadIdx.x] = ~0ULL; s_tmax[threadIdx.x] = 0; s_done[threadIdx.x] = 0;
  }
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  // The list was cut with the prefixes known during pass 2; pass 2's pick has since added a digit.
  // One sweep (eight loads in flight per thread) keeps the keys that match a target's prefix as
  // it is now -- a few per target unless values repeat -- in shared memory; the passes run on those.
  if (GATHERED) {
    unsigned total = 0; bool fits = true;
    for (int r = 0; r < n_ranks; ++r) {
      const unsigned long long c = rows[((size_t)r * gridDim.x + sg) * GATHER_ROW];
      if (c > GATHER_KEYS) fits = false; else total += (unsigned)c;
    }
    if (threadIdx.x < n_t) {
      unsigned long long lo = ~0ULL, hi = 0;
      for (int r = 0; r < n_ranks; ++r) {
        const unsigned long long* row = rows + ((size_t)r * gridDim.x + sg) * GATHER_ROW;
        const unsigned long long a = row[GATHER_ROW - 2 * MAX_TARGETS + threadIdx.x], b = row[GATHER_ROW - MAX_TARGETS + threadIdx.x];
        lo = a < lo ? a : lo; hi = b > hi ? b : hi;
      }
      s_tmin[threadIdx.x] = lo; s_tmax[threadIdx.x] = hi;
    }
    __syncthreads();
    if (!fits || total > TAIL_KEYS) {
      // the keys do not fit -- but targets whose remaining keys are all one value need no keys
      bool all = true;
      for (int tg = 0; tg < n_t; ++tg) all = all && s_tmin[tg] == s_tmax[tg];
      if (!all) { if (threadIdx.x == 0) atomicAdd(unresolved, 1u); return; }
      if (threadIdx.x < n_t) { st[ST_PF(threadIdx.x)] = s_tmin[threadIdx.x]; st[ST_RK(threadIdx.x)] = 0; }
      if (threadIdx.x == 0) st[1] |= 1ULL << ST_RESOLVED_BIT;
      return;
    }
    unsigned off = 0;
    for (int r = 0; r < n_ranks; ++r) {
      const unsigned long long* row = rows + ((size_t)r * gridDim.x + sg) * GATHER_ROW;
      const unsigned c = (unsigned)row[0];
      for (unsigned i = threadIdx.x; i < c; i += blockDim.x) skeys[off + i] = row[1 + i];
      off += c;
    }
    if (threadIdx.x == 0) s_cnt = total;
  } else if (!from_data) {
    const Geom g0 = seg_geom(st, from_pass);
    const unsigned long long* glist = cand + (size_t)sg * cap;
    for (unsigned i0 = 0; i0 < n_list; i0 += blockDim.x * 8) {
      unsigned long long kk[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const unsigned i = i0 + u * blockDim.x + threadIdx.x;
        kk[u] = i < n_list ? glist[i] : ~0ULL;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const unsigned i = i0 + u * blockDim.x + threadIdx.x;
        bool hit = false;
        if (i < n_list && g0.bits > 0) {
          const uint64_t top = kk[u] >> g0.pshift;
          for (int tg = 0; tg < n_t; ++tg) hit |= top == s_pf[tg];
        }
        if (hit) {
          const unsigned pos = atomicAdd(&s_cnt, 1u); if (pos < TAIL_KEYS) skeys[pos] = kk[u];
          const uint64_t top = kk[u] >> g0.pshift;
          for (int tg = 0; tg < n_t; ++tg)
            if (top == s_pf[tg]) {                           // plain reads first: both only ever move one way
              if (kk[u] < *(volatile unsigned long long*)&s_tmin[tg]) atomicMin(&s_tmin[tg], kk[u]);
              if (kk[u] > *(volatile unsigned long long*)&s_tmax[tg]) atomicMax(&s_tmax[tg], kk[u]);
            }
        }
      }
    }
  }
  __syncthreads();
  const unsigned n = s_cnt;
  // Too many keys still match for shared memory (repeated values, or a narrow value range that
  // two digits did not split): targets whose keys are all equal are settled at once; for the
  // others the remaining passes read the candidate list itself (or, if that overflowed, the data).
  bool from_list = false;
  if (n > TAIL_KEYS) {
    if (GATHERED) return;                                   // (not reached: the merge checked the counts)
    if (threadIdx.x < n_t && s_tmin[threadIdx.x] == s_tmax[threadIdx.x]) {
      s_done[threadIdx.x] = 1; s_pf[threadIdx.x] = s_tmin[threadIdx.x]; s_k[threadIdx.x] = s_nk[threadIdx.x] = 0;
    }
    __syncthreads();
    from_list = true;
  }
  if (from_data) {                                          // overflowed list: the same shortcut needs one sweep of the data
    from_list = false;
  }
  const unsigned long long* keys = skeys;
  if (!from_data && !from_list && n <= 512) {
    // A short list is settled directly: among the keys that carry the same prefix, key i has
    // r keys below it and e equal to it, and it is target t's order statistic iff r <= k_t < r + e.
    const Geom g0 = seg_geom(st, from_pass);
    if (g0.bits > 0) {
      for (unsigned i = threadIdx.x; i < n; i += blockDim.x) {
        const uint64_t ki = keys[i], top = ki >> g0.pshift;
        unsigned r = 0, e = 0;
        for (unsigned j = 0; j < n; ++j) {
          const uint64_t kj = keys[j];
          if ((kj >> g0.pshift) == top) { r += kj < ki; e += kj == ki; }
        }
        for (int tg = 0; tg < n_t; ++tg)
          if (s_pf[tg] == top && s_k[tg] >= r && s_k[tg] < (unsigned long long)r + e) { s_nk[tg] = ki; }
      }
      __syncthreads();
      // s_nk now holds the complete keys; they become the prefixes, the ranks are used up
      if (threadIdx.x < n_t) { st[ST_PF(threadIdx.x)] = s_nk[threadIdx.x]; st[ST_RK(threadIdx.x)] = 0; }
    }
    if (threadIdx.x == 0) st[1] |= 1ULL << ST_RESOLVED_BIT;
    return;
  }
  constexpr int PER = BINS / 256;
  for (int pass = from_pass; pass <= PMX_QUANTILE_PASSES; ++pass) {
    const Geom gm = seg_geom(st, pass);
    if (gm.bits == 0) break;                                // uniform: nothing left to resolve
    const unsigned mask = (1u << gm.bits) - 1u;
    __syncthreads();
    if (threadIdx.x < n_t) s_old[threadIdx.x] = s_pf[threadIdx.x];
    __syncthreads();
    for (int tg = 0; tg < n_t; ++tg) {
      if (s_done[tg]) continue;                             // settled by equal keys
      bool is_rep = true;
      for (int j = 0; j < tg; ++j) if (!s_done[j] && s_old[j] == s_old[tg]) { is_rep = false; break; }
      if (!is_rep) continue;                                // uniform
      for (int i = threadIdx.x; i < BINS; i += blockDim.x) sh[i] = 0;
      __syncthreads();
      const uint64_t want = s_old[tg];
      if (from_list) {
        const unsigned long long* glist = cand + (size_t)sg * cap;
        for (unsigned i0 = 0; i0 < n_list; i0 += blockDim.x * 4) {
          unsigned long long kk[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) { const unsigned i = i0 + u * blockDim.x + threadIdx.x; kk[u] = i < n_list ? glist[i] : ~0ULL; }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const unsigned i = i0 + u * blockDim.x + threadIdx.x;
            hist_add(sh, i < n_list && (kk[u] >> gm.pshift) == want, (unsigned)(kk[u] >> gm.shift) & mask);
          }
        }
      } else if (!from_data) {
        for (unsigned i0 = 0; i0 < n; i0 += blockDim.x) {
          const unsigned i = i0 + threadIdx.x;
          const bool in = i < n;
          const uint64_t key = in ? keys[i] : 0;
          hist_add(sh, in && (key >> gm.pshift) == want, (unsigned)(key >> gm.shift) & mask);
        }
      } else {
        for (int64_t r = 0; r < n_series; ++r) {
          if (seg[r] != sg) continue;                       // uniform
          const double* xr = x + r * n_steps;
          for (int64_t j0 = 0; j0 < n_steps; j0 += blockDim.x) {
            const int64_t j = j0 + threadIdx.x;
            const double v = j < n_steps ? xr[j] : pmx_nan();
            const uint64_t key = f64_key(v) - gm.base;
            hist_add(sh, v == v && (key >> gm.pshift) == want, (unsigned)(key >> gm.shift) & mask);
          }
        }
      }
      __syncthreads();
      unsigned int loc[PER]; unsigned int sum = 0;
#pragma unroll
      for (int i = 0; i < PER; ++i) { loc[i] = sh[threadIdx.x * PER + i]; sum += loc[i]; }
      unsigned int inc = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const unsigned v = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += v; }
      if (lane == 31) s_warp[warp] = inc;
      __syncthreads();
      unsigned int base = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) if (w < warp) base += s_warp[w];
      for (int j = tg; j < n_t; ++j) {                      // every target that shares this prefix
        if (s_done[j] || s_old[j] != want) continue;
        const unsigned long long k = s_k[j];                // ranks are replaced only after the pass (s_nk)
        unsigned long long acc = base + inc - sum;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
          if (k >= acc && k < acc + loc[i]) {
            s_pf[j] = (want << gm.bits) | (uint64_t)(threadIdx.x * PER + i);   // prefixes are compared through s_old
            s_nk[j] = k - acc;
          }
          acc += loc[i];
        }
      }
      __syncthreads();                                      // sh and s_warp are reused by the next prefix
    }
    if (threadIdx.x < n_t) s_k[threadIdx.x] = s_nk[threadIdx.x];
  }
  __syncthreads();
  if (threadIdx.x < n_t) { st[ST_PF(threadIdx.x)] = s_pf[threadIdx.x]; st[ST_RK(threadIdx.x)] = s_k[threadIdx.x]; }
  if (threadIdx.x == 0) st[1] |= 1ULL << ST_RESOLVED_BIT;
}

// Multi-GPU form of the tail, step 1: every rank writes, per segment, the keys of its candidate
// list that match a target's prefix as it is now (row = count, keys).  The count is ~0 when the
// list itself overflowed and may exceed GATHER_KEYS (then only the first GATHER_KEYS keys are there).
__global__ void __launch_bounds__(256)
tail_filter_kernel(const unsigned long long* __restrict__ cand, const unsigned int* __restrict__ cand_cnt,
                   unsigned cap, int n_t, int from_pass, const uint64_t* __restrict__ state, int sw,
                   unsigned long long* __restrict__ rows) {
  __shared__ uint64_t s_pf[MAX_TARGETS];
  __shared__ unsigned long long s_tmin[MAX_TARGETS], s_tmax[MAX_TARGETS];
  __shared__ unsigned int s_cnt;
  const int sg = blockIdx.x;
  const uint64_t* st = state + (size_t)sg * sw;
  unsigned long long* row = rows + (size_t)sg * GATHER_ROW;
  const unsigned n_list = cand_cnt[sg];
  if (threadIdx.x < MAX_TARGETS) {                          // no key of this rank matches: neutral for min / max
    row[GATHER_ROW - 2 * MAX_TARGETS + threadIdx.x] = ~0ULL; row[GATHER_ROW - MAX_TARGETS + threadIdx.x] = 0;
    s_tmin[threadIdx.x] = ~0ULL; s_tmax[threadIdx.x] = 0;
  }
  if (st[0] == 0) { if (threadIdx.x == 0) row[0] = 0; return; }
  if (n_list > cap) {                                       // an overflowed list says nothing about equal keys either
    if (threadIdx.x == 0) row[0] = ~0ULL;
    if (threadIdx.x < MAX_TARGETS) { row[GATHER_ROW - 2 * MAX_TARGETS + threadIdx.x] = 0; row[GATHER_ROW - MAX_TARGETS + threadIdx.x] = ~0ULL; }
    return;
  }
  if (threadIdx.x < n_t) s_pf[threadIdx.x] = st[ST_PF(threadIdx.x)];
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  const Geom g0 = seg_geom(st, from_pass);
  const unsigned long long* glist = cand + (size_t)sg * cap;
  for (unsigned i0 = 0; i0 < n_list; i0 += blockDim.x * 8) {
    unsigned long long kk[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const unsigned i = i0 + u * blockDim.x + threadIdx.x;
      kk[u] = i < n_list ? glist[i] : ~0ULL;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const unsigned i = i0 + u * blockDim.x + threadIdx.x;
      bool hit = false;
      if (i < n_list && g0.bits > 0) {
        const uint64_t top = kk[u] >> g0.pshift;
        for (int tg = 0; tg < n_t; ++tg) hit |= top == s_pf[tg];
      }
      if (hit) {
        const unsigned pos = atomicAdd(&s_cnt, 1u); if (pos < GATHER_KEYS) row[1 + pos] = kk[u];
        const uint64_t top = kk[u] >> g0.pshift;
        for (int tg = 0; tg < n_t; ++tg)
          if (top == s_pf[tg]) {
            if (kk[u] < *(volatile unsigned long long*)&s_tmin[tg]) atomicMin(&s_tmin[tg], kk[u]);
            if (kk[u] > *(volatile unsigned long long*)&s_tmax[tg]) atomicMax(&s_tmax[tg], kk[u]);
          }
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) row[0] = s_cnt;
  if (threadIdx.x < n_t) {
    row[GATHER_ROW - 2 * MAX_TARGETS + threadIdx.x] = s_tmin[threadIdx.x];
    row[GATHER_ROW - MAX_TARGETS + threadIdx.x] = s_tmax[threadIdx.x];
  }
}

// One CTA per segment walks the segment's targets.  Targets that share a prefix share the
// histogram of the first of them (the hist kernels fill only that one), so all old prefixes are
// read before any is replaced.
__global__ void __launch_bounds__(256)
pick_kernel(const unsigned long long* __restrict__ hist, const double* __restrict__ q, int n_q,
            int pass, uint64_t* __restrict__ state, int sw) {
  const int sg = blockIdx.x;
  const int n_t = 2 * n_q;
  const int n_h = pass == 0 ? 1 : n_t;
  uint64_t* st = state + (size_t)sg * sw;
  __shared__ uint64_t s_old[MAX_TARGETS];
  __shared__ unsigned long long s_warp[8];
  __shared__ uint64_t s_pf[MAX_TARGETS], s_k[MAX_TARGETS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < n_t) {
    const uint64_t old = pass == 0 ? 0 : st[ST_PF(threadIdx.x)];
    s_old[threadIdx.x] = old;
    s_pf[threadIdx.x] = old;                                // kept if no bin holds the rank (inconsistent counts)
    s_k[threadIdx.x] = pass == 0 ? 0 : st[ST_RK(threadIdx.x)];
  }
  const unsigned long long n_known = pass == 0 ? 0 : st[0];
  const Geom gm = seg_geom(st, pass);
  __syncthreads();
  if (pass > 0 && (n_known == 0 || gm.bits == 0)) return;   // empty segment, or nothing left to resolve
  constexpr int PER = BINS / 256;                          // 8 bins per thread
  for (int tg = 0; tg < n_t; ++tg) {
    // one scan per distinct prefix: tg is a representative if no earlier target shares its prefix
    bool is_rep = true;
    if (pass == 0) is_rep = tg == 0;
    else for (int j = 0; j < tg; ++j) if (s_old[j] == s_old[tg]) { is_rep = false; break; }
    if (!is_rep) continue;
    const unsigned long long* h = hist + ((size_t)sg * n_h + tg) * BINS;
    unsigned long long loc[PER]; unsigned long long sum = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) { loc[i] = h[threadIdx.x * PER + i]; sum += loc[i]; }
    // block-wide exclusive scan of the 256 partial sums
    unsigned long long inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += v;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    unsigned long long base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { const unsigned long long v = s_warp[w]; if (w < warp) base += v; total += v; }
    if (pass == 0) {
      if (threadIdx.x == 0) { st[0] = total; st[1] = 0; st[2] = 0; }
      if (total == 0) { if (threadIdx.x < n_t) { st[ST_PF(threadIdx.x)] = 0; st[ST_RK(threadIdx.x)] = 0; } return; }
    }
    for (int j = tg; j < n_t; ++j) {                        // every target that shares this histogram
      if (pass > 0 && s_old[j] != s_old[tg]) continue;
      unsigned long long k;
      if (pass == 0) {
        const unsigned long long n = total;
        const double pos = (double)(n - 1) * q[j >> 1];
        unsigned long long lo = (unsigned long long)floor(pos);
        if (lo > n - 1) lo = n - 1;
        const unsigned long long hi = lo + 1 < n ? lo + 1 : n - 1;
        k = (j & 1) ? hi : lo;
      } else {
        k = st[ST_RK(j)];                                  // written by an earlier pass only
      }
      // the bin b with cum[b] <= k < cum[b+1]
      unsigned long long acc = base + inc - sum;
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        if (k >= acc && k < acc + loc[i]) {
          s_pf[j] = (s_old[j] << gm.bits) | (uint64_t)(threadIdx.x * PER + i);
          s_k[j] = k - acc;
        }
        acc += loc[i];
      }
    }
    __syncthreads();                                        // s_warp is reused by the next scan
  }
  if (threadIdx.x < n_t) { st[ST_PF(threadIdx.x)] = s_pf[threadIdx.x]; st[ST_RK(threadIdx.x)] = s_k[threadIdx.x]; }
}

// pmx_quantile_init: begin from what a group reduction already knows.  Every key of a segment lies
// in [key(min), key(max)]; relative to key(min) the keys have Z = clz(key(max) - key(min)) leading
// zero bits, so the passes start behind them and pass 0 is not needed.
__global__ void init_kernel(const long long* __restrict__ count, const double* __restrict__ mn,
                            const double* __restrict__ mx, const double* __restrict__ q, int n_q,
                            uint64_t* __restrict__ state, int sw, int64_t n_seg) {
  const int64_t sg = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (sg >= n_seg) return;
  uint64_t* st = state + (size_t)sg * sw;
  const long long n = count[sg];
  int Z = 64; uint64_t base = 0;
  if (n > 0) {
    // a zero minimum may stand for -0.0 and a zero maximum for +0.0 (min/max do not order the two)
    const double a = mn[sg], b = mx[sg];
    const uint64_t ka = f64_key(a == 0.0 ? -0.0 : a), kb = f64_key(b == 0.0 ? 0.0 : b);
    base = ka;
    Z = kb > ka ? __clzll((long long)(kb - ka)) : 64;
  }
  st[0] = n > 0 ? (uint64_t)n : 0;
  st[1] = (1ULL << 32) | (uint64_t)Z;
  st[2] = base;
  for (int tg = 0; tg < 2 * n_q; ++tg) {
    unsigned long long k = 0;
    if (n > 0) {
      const double pos = (double)(n - 1) * q[tg >> 1];
      unsigned long long lo = (unsigned long long)floor(pos);
      if (lo > (unsigned long long)n - 1) lo = (unsigned long long)n - 1;
      const unsigned long long hi = lo + 1 < (unsigned long long)n ? lo + 1 : (unsigned long long)n - 1;
      k = (tg & 1) ? hi : lo;
    }
    st[ST_PF(tg)] = 0;
    st[ST_RK(tg)] = k;
  }
}

__global__ void finish_kernel(const uint64_t* __restrict__ state, int sw, int64_t n_seg,
                              const double* __restrict__ q, int n_q, int64_t* o_n, int64_t* o_lo,
                              int64_t* o_hi, double* o_xlo, double* o_xhi, double* o_val) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_seg * n_q) return;
  const int64_t sg = i / n_q; const int qi = (int)(i % n_q);
  const uint64_t* st = state + (size_t)sg * sw;
  const unsigned long long n = st[0];
  if (qi == 0 && o_n) o_n[sg] = (int64_t)n;
  if (n == 0) {
    if (o_lo) o_lo[i] = -1;
    if (o_hi) o_hi[i] = -1;
    if (o_xlo) o_xlo[i] = pmx_nan();
    if (o_xhi) o_xhi[i] = pmx_nan();
    if (o_val) o_val[i] = pmx_nan();
    return;
  }
  const double pos = (double)(n - 1) * q[qi];
  unsigned long long lo = (unsigned long long)floor(pos);
  if (lo > n - 1) lo = n - 1;
  const unsigned long long hi = lo + 1 < n ? lo + 1 : n - 1;
  const double t = pos - (double)lo;
  const uint64_t base = ((st[1] >> 32) & 1) ? st[2] : 0;
  const double a = key_f64(st[ST_PF(2 * qi)] + base);
  const double b = key_f64(st[ST_PF(2 * qi + 1)] + base);
  // numpy _lerp: a + (b-a)*t, and b - (b-a)*(1-t) when t >= 0.5 (no FMA contraction)
  const double d = __dsub_rn(b, a);
  double v = __dadd_rn(a, __dmul_rn(d, t));
  if (t >= 0.5) v = __dsub_rn(b, __dmul_rn(d, __dsub_rn(1.0, t)));
  if (o_lo) o_lo[i] = (int64_t)lo;
  if (o_hi) o_hi[i] = (int64_t)hi;
  if (o_xlo) o_xlo[i] = a;
  if (o_xhi) o_xhi[i] = b;
  if (o_val) o_val[i] = v;
}

}  // namespace

// candidate lists in ctx scratch: [n_seg counts, overflow flag | padding][n_seg * cap keys]
static int cand_layout(pmx_ctx* ctx, int64_t n_seg, unsigned* cap, unsigned int** cand_cnt, unsigned long long** cand,
                       size_t* hdr_bytes) {
  int64_t c = ((int64_t)256 << 20) / 8 / n_seg;
  if (c > (1 << 20)) c = 1 << 20;
  if (c < 4096) c = 4096;
  *cap = (unsigned)c;
  const size_t hdr = ((size_t)(n_seg + 1) * sizeof(unsigned int) + 255) & ~size_t(255);
  unsigned char* scr = (unsigned char*)pmx_scratch(ctx, PMX_SCR_QCAND, hdr + (size_t)n_seg * c * 8);
  if (!scr) return PMX_E_CUDA;
  *cand_cnt = (unsigned int*)scr;
  *cand = (unsigned long long*)(scr + hdr);
  if (hdr_bytes) *hdr_bytes = hdr;
  return PMX_OK;
}

extern "C" {

int pmx_quantile_hist(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                      const int32_t* d_seg, int64_t n_seg, const int32_t* d_row_order, int64_t n_rows,
                      int64_t n_q, int pass, const uint64_t* d_state, uint64_t* d_hist, pmx_stream stream) {
  return pmx_quantile_hist_lp(ctx, d_x, n_series, n_steps, d_seg, n_seg, d_row_order, n_rows, n_q, pass, 2, d_state,
                              d_hist, stream);
}

int pmx_quantile_list_pass(int64_t n_series, int64_t n_steps, int64_t n_seg) {
  // candidates after pass p number about n * (share of a 2^(11 p) th of the key range) per target; a
  // list holds 2^20 keys.  Counter-like values (half of them inside one binary exponent) reach
  // ~n/40 per distinct prefix after one digit, so segments beyond a few million values wait a digit.
  if (n_seg <= 0) return 2;
  return (n_series * n_steps) / n_seg > ((int64_t)6 << 20) ? 3 : 2;
}

int pmx_quantile_hist_lp(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                         const int32_t* d_seg, int64_t n_seg, const int32_t* d_row_order, int64_t n_rows,
                         int64_t n_q, int pass, int list_pass, const uint64_t* d_state, uint64_t* d_hist,
                         pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(list_pass == 2 || list_pass == 3, "list_pass must be 2 or 3");
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0 && n_seg >= 0 && n_seg <= 65535, "bad sizes (n_seg <= 65535)");
  PMX_REQUIRE(n_q >= 1 && 2 * n_q <= MAX_TARGETS, "n_q must be in [1, 8]");
  PMX_REQUIRE(pass >= 0 && pass <= PMX_QUANTILE_PASSES, "pass out of range");
  if (n_seg == 0) return PMX_OK;
  PMX_REQUIRE(d_hist && d_state, "NULL buffer");
  const int n_t = 2 * (int)n_q;
  const int n_h = pass == 0 ? 1 : n_t;
  cudaStream_t st = pmx_st(stream);
  PMX_CUDA(cudaMemsetAsync(d_hist, 0, (size_t)n_seg * n_h * BINS * sizeof(uint64_t), st));
  if (n_series == 0 || n_steps == 0) return PMX_OK;
  PMX_REQUIRE(d_x && d_seg, "NULL buffer");
  // candidate lists (ctx scratch; they live from pass 2 to pass 5 of one selection)
  unsigned cap = 0; unsigned int* cand_cnt = nullptr; unsigned long long* cand = nullptr;
  int mode = 0;
  if (pass >= list_pass) {
    size_t hdr;
    if (int rc = cand_layout(ctx, n_seg, &cap, &cand_cnt, &cand, &hdr)) return rc;
    mode = 2;
    if (pass == list_pass) PMX_CUDA(cudaMemsetAsync(cand_cnt, 0, hdr, st));
  }
  const size_t smem = (size_t)n_h * BINS * sizeof(unsigned int);
  PMX_CUDA(cudaFuncSetAttribute(hist_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(MAX_TARGETS * BINS * sizeof(unsigned int))));
  PMX_CUDA(cudaFuncSetAttribute(hist_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(MAX_TARGETS * BINS * sizeof(unsigned int))));
  // a CTA sweeps consecutive rows so that a segment's private histogram is flushed rarely
  if (!d_row_order) n_rows = n_series;
  PMX_REQUIRE(n_rows >= 0 && n_rows <= n_series, "n_rows out of range");
  if (n_rows == 0) return PMX_OK;
  int64_t want_ctas = (int64_t)ctx->sm_count * 16;         // measured: fewer, longer CTAs are slower
  int64_t rows = (n_rows + want_ctas - 1) / want_ctas;
  if (rows < 1) rows = 1;
  const int64_t grid = (n_rows + rows - 1) / rows;
  unsigned int* overflow = cand_cnt ? cand_cnt + n_seg : nullptr;
  if (pass == list_pass) {
    // the list pass reads the data once, only to cut the lists; its digit then comes from the lists
    collect_kernel<<<(unsigned)grid, 256, 0, st>>>(d_x, n_steps, d_seg, n_t, pass, d_state, state_words(n_q), (int)rows,
                                                   n_rows, d_row_order, cand, cand_cnt, cap, overflow);
    PMX_LAUNCH_CHECK();
  }
  if (mode == 2) {
    PMX_CUDA(cudaFuncSetAttribute(cand_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(MAX_TARGETS * BINS * sizeof(unsigned int))));
    dim3 grid(32, (unsigned)n_seg);
    cand_hist_kernel<<<grid, 256, smem, st>>>(cand, cand_cnt, cap, n_t, pass, d_state, state_words(n_q),
                                            (unsigned long long*)d_hist);
    PMX_LAUNCH_CHECK();
  }
  if (mode == 0)
    hist_kernel<0><<<(unsigned)grid, 256, smem, st>>>(d_x, n_steps, d_seg, n_t, pass, d_state, state_words(n_q),
                                                     (unsigned long long*)d_hist, (int)rows, n_rows, d_row_order,
                                                     cand, cand_cnt, cap, overflow);
  else
    hist_kernel<2><<<(unsigned)grid, 256, smem, st>>>(d_x, n_steps, d_seg, n_t, pass, d_state, state_words(n_q),
                                                     (unsigned long long*)d_hist, (int)rows, n_rows, d_row_order,
                                                     cand, cand_cnt, cap, overflow);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_quantile_tail(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg,
                      int64_t n_seg, int64_t n_q, int from_pass, uint64_t* d_state, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_seg >= 0 && n_seg <= 65535, "n_seg must be in [0, 65535]");
  PMX_REQUIRE(n_q >= 1 && 2 * n_q <= MAX_TARGETS, "n_q must be in [1, 8]");
  PMX_REQUIRE(from_pass >= 3 && from_pass <= PMX_QUANTILE_PASSES, "from_pass must follow pass 2 (which builds the candidate lists)");
  if (n_seg == 0) return PMX_OK;
  PMX_REQUIRE(d_state && d_x && d_seg && n_series > 0 && n_steps > 0, "NULL buffer or empty data");
  unsigned cap; unsigned int* cand_cnt; unsigned long long* cand;
  if (int rc = cand_layout(ctx, n_seg, &cap, &cand_cnt, &cand, nullptr)) return rc;
  const size_t smem = (size_t)TAIL_KEYS * sizeof(unsigned long long);
  PMX_CUDA(cudaFuncSetAttribute(tail_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tail_kernel<false><<<(unsigned)n_seg, 256, smem, pmx_st(stream)>>>(cand, cand_cnt, cap, 2 * (int)n_q, from_pass, d_state,
                                                                   state_words(n_q), nullptr, 1, nullptr, d_x, n_series,
                                                                   n_steps, d_seg);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_quantile_tail_filter(pmx_ctx* ctx, int64_t n_seg, int64_t n_q, int from_pass, const uint64_t* d_state,
                             uint64_t* d_rows, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_seg >= 0 && n_seg <= 65535, "n_seg must be in [0, 65535]");
  PMX_REQUIRE(n_q >= 1 && 2 * n_q <= MAX_TARGETS, "n_q must be in [1, 8]");
  PMX_REQUIRE(from_pass >= 3 && from_pass <= PMX_QUANTILE_PASSES, "from_pass must follow pass 2 (which builds the candidate lists)");
  if (n_seg == 0) return PMX_OK;
  PMX_REQUIRE(d_state && d_rows, "NULL buffer");
  unsigned cap; unsigned int* cand_cnt; unsigned long long* cand;
  if (int rc = cand_layout(ctx, n_seg, &cap, &cand_cnt, &cand, nullptr)) return rc;
  tail_filter_kernel<<<(unsigned)n_seg, 256, 0, pmx_st(stream)>>>(cand, cand_cnt, cap, 2 * (int)n_q, from_pass, d_state,
                                                                state_words(n_q), (unsigned long long*)d_rows);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_quantile_tail_merge(pmx_ctx* ctx, const uint64_t* d_rows_all, int n_ranks, int64_t n_seg, int64_t n_q,
                            int from_pass, uint64_t* d_state, int32_t* d_unresolved, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_seg >= 0 && n_seg <= 65535, "n_seg must be in [0, 65535]");
  PMX_REQUIRE(n_q >= 1 && 2 * n_q <= MAX_TARGETS, "n_q must be in [1, 8]");
  PMX_REQUIRE(n_ranks >= 1, "n_ranks must be positive");
  PMX_REQUIRE(from_pass >= 3 && from_pass <= PMX_QUANTILE_PASSES, "from_pass must follow pass 2");
  PMX_REQUIRE(d_unresolved, "NULL buffer");
  cudaStream_t st = pmx_st(stream);
  PMX_CUDA(cudaMemsetAsync(d_unresolved, 0, sizeof(int32_t), st));
  if (n_seg == 0) return PMX_OK;
  PMX_REQUIRE(d_state && d_rows_all, "NULL buffer");
  const size_t smem = (size_t)TAIL_KEYS * sizeof(unsigned long long);
  PMX_CUDA(cudaFuncSetAttribute(tail_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tail_kernel<true><<<(unsigned)n_seg, 256, smem, st>>>(nullptr, nullptr, 0xffffffffu, 2 * (int)n_q, from_pass, d_state,
                                                      state_words(n_q), (const unsigned long long*)d_rows_all, n_ranks,
                                                      (unsigned int*)d_unresolved, nullptr, 0, 0, nullptr);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_quantile_init(pmx_ctx* ctx, const int64_t* d_count, const double* d_min, const double* d_max,
                      int64_t n_seg, const double* d_q, int64_t n_q, uint64_t* d_state, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_seg >= 0 && n_seg <= 65535, "n_seg must be in [0, 65535]");
  PMX_REQUIRE(n_q >= 1 && 2 * n_q <= MAX_TARGETS, "n_q must be in [1, 8]");
  if (n_seg == 0) return PMX_OK;
  PMX_REQUIRE(d_count && d_min && d_max && d_q && d_state, "NULL buffer");
  init_kernel<<<(unsigned)((n_seg + 127) / 128), 128, 0, pmx_st(stream)>>>((const long long*)d_count, d_min, d_max, d_q,
                                                                         (int)n_q, d_state, state_words(n_q), n_seg);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_quantile_pick(pmx_ctx* ctx, const uint64_t* d_hist, int64_t n_seg, const double* d_q,
                      int64_t n_q, int pass, uint64_t* d_state, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_seg >= 0 && n_seg <= 65535, "n_seg must be in [0, 65535]");
  PMX_REQUIRE(n_q >= 1 && 2 * n_q <= MAX_TARGETS, "n_q must be in [1, 8]");
  PMX_REQUIRE(pass >= 0 && pass <= PMX_QUANTILE_PASSES, "pass out of range");
  if (n_seg == 0) return PMX_OK;
  PMX_REQUIRE(d_hist && d_q && d_state, "NULL buffer");
  pick_kernel<<<(unsigned)n_seg, 256, 0, pmx_st(stream)>>>((const unsigned long long*)d_hist, d_q, (int)n_q, pass, d_state,
                                               state_words(n_q));
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_quantile_finish(pmx_ctx* ctx, const uint64_t* d_state, int64_t n_seg, const double* d_q,
                        int64_t n_q, int64_t* d_n, int64_t* d_lo, int64_t* d_hi, double* d_x_lo,
                        double* d_x_hi, double* d_value, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_seg >= 0, "bad sizes");
  PMX_REQUIRE(n_q >= 1 && 2 * n_q <= MAX_TARGETS, "n_q must be in [1, 8]");
  if (n_seg == 0) return PMX_OK;
  PMX_REQUIRE(d_state && d_q, "NULL buffer");
  const int64_t n = n_seg * n_q;
  finish_kernel<<<(unsigned)((n + 127) / 128), 128, 0, pmx_st(stream)>>>(d_state, state_words(n_q), n_seg, d_q, (int)n_q,
                                                                        d_n, d_lo, d_hi, d_x_lo, d_x_hi, d_value);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_quantile_select(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                        const int32_t* d_seg, int64_t n_seg, const int32_t* d_row_order, int64_t n_rows,
                        const double* d_q, int64_t n_q,
                        int64_t* d_n, int64_t* d_lo, int64_t* d_hi, double* d_x_lo, double* d_x_hi,
                        double* d_value, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_q >= 1 && 2 * n_q <= MAX_TARGETS, "n_q must be in [1, 8]");
  PMX_REQUIRE(n_seg >= 0 && n_seg <= 65535, "n_seg must be in [0, 65535]");
  if (n_seg == 0) return PMX_OK;
  const size_t state_bytes = ((size_t)n_seg * state_words(n_q) * 8 + 255) & ~size_t(255);
  const size_t hist_bytes = (size_t)n_seg * 2 * n_q * BINS * 8;
  unsigned char* scr = (unsigned char*)pmx_scratch(ctx, PMX_SCR_QSELECT, state_bytes + hist_bytes);
  if (!scr) return PMX_E_CUDA;
  uint64_t* state = (uint64_t*)scr;
  uint64_t* hist = (uint64_t*)(scr + state_bytes);
  PMX_CUDA(cudaMemsetAsync(state, 0, state_bytes, pmx_st(stream)));
  // passes 0..2 over the data; the candidate lists of pass 2 then settle every segment in one launch
  const bool data = n_series > 0 && n_steps > 0;
  for (int pass = 0; pass < (data ? 3 : PMX_QUANTILE_PASSES); ++pass) {
    int rc = pmx_quantile_hist(ctx, d_x, n_series, n_steps, d_seg, n_seg, d_row_order, n_rows, n_q, pass, state, hist,
                               stream);
    if (rc) return rc;
    rc = pmx_quantile_pick(ctx, hist, n_seg, d_q, n_q, pass, state, stream);
    if (rc) return rc;
  }
  if (data) {
    const int rc = pmx_quantile_tail(ctx, d_x, n_series, n_steps, d_seg, n_seg, n_q, 3, state, stream);
    if (rc) return rc;
  }
  return pmx_quantile_finish(ctx, state, n_seg, d_q, n_q, d_n, d_lo, d_hi, d_x_lo, d_x_hi, d_value, stream);
}

}  // extern "C"
