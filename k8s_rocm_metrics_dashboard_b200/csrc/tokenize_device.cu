// tokenize_device.cu -- X1 on the GPU: the response body is already in device
// memory (one pinned H2D copy of the bytes that came off the socket) and is cut into sample tokens
// there, so the host never walks it.  Replaces json.loads at app.py:163/178 for bodies in
// Prometheus' own compact encoding; anything else is reported, not guessed (the host tokenizer in
// tokenize.cu accepts general JSON).
//
// The body is a flat array of bytes; what a byte means depends on whether it lies inside a JSON
// string, i.e. on the parity of the unescaped quotes before it.  Three kernels:
//   count : warp w walks chunk w (8 KiB) in rounds of 512 bytes, 16 bytes per lane per round, and
//           reports the chunk's quote parity and its number of sample starts ('[' followed by a
//           digit or '-'), series starts ('{"metric":{') and '[' -- each counted twice, once for
//           "the chunk begins outside a string" and once for "inside";
//   scan  : one CTA turns the parities into every chunk's true starting state and the counts that
//           belong to it into output offsets, and checks the envelope and the bracket balance;
//   emit  : the same walk with the state known; every sample start writes its two token offsets
//           (the ',' behind the timestamp is found in the lane's and its neighbour's comma masks),
//           every series start its CSR offset and the span of its label object.
// A lane never looks at its 16 bytes one by one: each byte class ('"', '[' or '{', ',') becomes a
// 16-bit mask by an exact SWAR comparison of the four words, the state inside the segment is the
// prefix parity of the quote mask, and only set bits are visited.  A chunk belongs to one warp, so
// the running state passes from round to round by ballot and shuffle: no shared memory, no barrier.
// Token offsets have the meaning pmx_tokenize gives them, so pmx_parse_f64 / pmx_parse_ts_ms take
// the body and these arrays as they are.
#include <stdint.h>
#include "pmx_common.cuh"
#include "byte_masks.h"

namespace {

using namespace pmx;

constexpr int TT = 256;                 // threads per CTA
constexpr int WARPS = TT / 32;
constexpr int SEG = 16;                 // bytes per lane per round
constexpr int ROUND = 32 * SEG;         // 512 bytes per warp and round
constexpr int ROUNDS = 16;              // rounds per chunk
constexpr int64_t CHUNK = (int64_t)ROUND * ROUNDS;            // 8 KiB, one warp
constexpr unsigned FULL = 0xffffffffu;

struct ChunkSum { uint32_t parity, samples[2], series[2], brackets[2]; };
struct ChunkBase { uint32_t in_string; uint32_t pad; int64_t samples, series; };

// status words written by the scan / emit kernels
enum { ST_SAMPLES = 0, ST_SERIES = 1, ST_ERROR = 2, ST_WHERE = 3 };
enum { ERR_ENVELOPE = 1, ERR_BRACKETS = 2, ERR_CAPACITY = 3, ERR_SAMPLE = 4, ERR_LABELS = 5 };

__device__ __forceinline__ unsigned char byte_at(const uint8_t* __restrict__ b, uint64_t len, uint64_t i) {
  return i < len ? b[i] : (unsigned char)0;
}

// The 16 bytes of this lane's segment (zeros past the end), as four words.
__device__ __forceinline__ void load_words(const uint8_t* __restrict__ body, uint64_t len, uint64_t pos, uint32_t* w) {
  w[0] = w[1] = w[2] = w[3] = 0;
  if (pos + SEG <= len) {
    const uint4 v = *reinterpret_cast<const uint4*>(body + pos);
    w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
  } else {
    for (int k = 0; k < SEG; ++k)
      if (pos + k < len) w[k >> 2] |= (uint32_t)body[pos + k] << (8 * (k & 3));
  }
}

// Mask of the segment's unescaped quotes.  A quote is escaped iff an odd run of backslashes ends
// right before it; backslashes are rare, so only a segment that may hold one, or that begins with a
// quote behind a backslash, looks back through global memory.  All lanes of the warp call this.
__device__ __forceinline__ unsigned quote_mask(const uint8_t* __restrict__ body, uint64_t len, uint64_t pos,
                                               const uint32_t* w) {
  unsigned q = eq16(w, 0x22222222u);
  const int lane = threadIdx.x & 31;
  unsigned prev = __shfl_up_sync(FULL, w[3] >> 24, 1);
  if (lane == 0) prev = (pos > 0 && pos <= len) ? body[pos - 1] : 0u;
  const uint32_t bs = maybe4(w[0], 0x5c5c5c5cu) | maybe4(w[1], 0x5c5c5c5cu) | maybe4(w[2], 0x5c5c5c5cu) |
                      maybe4(w[3], 0x5c5c5c5cu);
  if (q && (bs || (prev == '\\' && (q & 1u)))) {
    unsigned m = q;
    while (m) {
      const int k = __ffs(m) - 1;
      m &= m - 1;
      uint64_t i = pos + k;
      int run = 0;
      while (i > 0 && body[i - 1] == '\\') { ++run; --i; }
      if (run & 1) q &= ~(1u << k);
    }
  }
  return q;
}

__device__ __forceinline__ bool starts_number(unsigned c) { return (c >= '0' && c <= '9') || c == '-'; }

// '{' at p opens a series object iff it is followed by "metric":{
__device__ __forceinline__ bool series_at(const uint8_t* __restrict__ body, uint64_t len, uint64_t p) {
  const char pat[10] = {'"', 'm', 'e', 't', 'r', 'i', 'c', '"', ':', '{'};
  if (p + 10 >= len) return false;
#pragma unroll
  for (int k = 0; k < 10; ++k) if (body[p + 1 + k] != (unsigned char)pat[k]) return false;
  return true;
}

__device__ __forceinline__ unsigned warp_sum(unsigned v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
  return v;
}

__global__ void __launch_bounds__(TT)
tokd_count_kernel(const uint8_t* __restrict__ body, uint64_t len, ChunkSum* __restrict__ sums, int64_t n_chunks) {
  const int lane = threadIdx.x & 31;
  const int64_t chunk = (int64_t)blockIdx.x * WARPS + (threadIdx.x >> 5);
  if (chunk >= n_chunks) return;                            // whole warps leave
  const uint64_t c0 = (uint64_t)chunk * CHUNK;
  unsigned run = 0;                                         // quote parity of the chunk so far
  unsigned n_smp[2] = {0, 0}, n_ser[2] = {0, 0}, n_br[2] = {0, 0};
  for (int r = 0; r < ROUNDS; ++r) {
    const uint64_t base = c0 + (uint64_t)r * ROUND;
    if (base >= len) break;                                 // uniform
    const uint64_t pos = base + (uint64_t)lane * SEG;
    uint32_t w[4];
    load_words(body, len, pos, w);
    const unsigned q = quote_mask(body, len, pos, w);
    const unsigned in0 = inside_mask(q);
    const unsigned ballot = __ballot_sync(FULL, (in0 >> 16) & 1u);
    const unsigned before = run ^ (__popc(ballot & ((1u << lane) - 1u)) & 1u);
    run ^= __popc(ballot) & 1u;
    const unsigned ins = (in0 ^ (before ? 0xffffu : 0u)) & 0xffffu;     // inside a string if the chunk began outside
    const unsigned open = open_mask(w);
    unsigned next0 = __shfl_down_sync(FULL, w[0] & 0xffu, 1);
    if (lane == 31) next0 = byte_at(body, len, pos + SEG);
    if (open) {
      const unsigned brace = open & bit5_16(w), square = open & ~brace;
      n_br[0] += __popc(square & ~ins);
      n_br[1] += __popc(square & ins);
      unsigned m = square;
      while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        if (starts_number(k + 1 < SEG ? seg_byte(w, k + 1) : next0)) ++n_smp[(ins >> k) & 1u];
      }
      m = brace;
      while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        if (series_at(body, len, pos + k)) ++n_ser[(ins >> k) & 1u];
      }
    }
  }
  ChunkSum s;
  s.parity = run;
  s.samples[0] = warp_sum(n_smp[0]); s.samples[1] = warp_sum(n_smp[1]);
  s.series[0] = warp_sum(n_ser[0]); s.series[1] = warp_sum(n_ser[1]);
  s.brackets[0] = warp_sum(n_br[0]); s.brackets[1] = warp_sum(n_br[1]);
  if (lane == 0) sums[chunk] = s;
}

// exclusive scans over a CTA of 1024 threads: XOR of one bit, sum of a 64-bit word.  *tot: CTA totals.
__device__ __forceinline__ unsigned cta_xor_excl(unsigned bit, unsigned* s_w, unsigned* tot) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned ballot = __ballot_sync(FULL, bit);
  if (lane == 0) s_w[warp] = __popc(ballot) & 1u;
  __syncthreads();
  const unsigned wb = __ballot_sync(FULL, s_w[lane]);       // 32 warps: one bit each
  __syncthreads();
  *tot = __popc(wb) & 1u;
  return (__popc(ballot & ((1u << lane) - 1u)) ^ __popc(wb & ((1u << warp) - 1u))) & 1u;
}
__device__ __forceinline__ unsigned long long cta_sum_excl(unsigned long long v, unsigned long long* s_w,
                                                           unsigned long long* tot) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const unsigned long long u = __shfl_up_sync(FULL, inc, d); if (lane >= d) inc += u; }
  if (lane == 31) s_w[warp] = inc;
  __syncthreads();
  unsigned long long wi = s_w[lane], wt = wi;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const unsigned long long u = __shfl_up_sync(FULL, wi, d); if (lane >= d) wi += u; }
  const unsigned long long below = __shfl_sync(FULL, wi - wt, warp);
  *tot = __shfl_sync(FULL, wi, 31);
  __syncthreads();
  return below + inc - v;
}

// One CTA: chunk states and output bases; envelope and balance checks; totals.  The chunk summaries
// are taken 1024 at a time, one per thread (coalesced); the state and the counts before each tile
// are carried in registers.  Inside a tile the sample and series counts share one 64-bit scan
// (at most 2^23 each).
__global__ void __launch_bounds__(1024)
tokd_scan_kernel(const uint8_t* __restrict__ body, uint64_t len, const ChunkSum* __restrict__ sums,
                 ChunkBase* __restrict__ bases, int64_t n_chunks, int64_t cap_samples, int64_t cap_series,
                 int64_t* __restrict__ series_off, int64_t* __restrict__ status) {
  __shared__ unsigned s_par[32];
  __shared__ unsigned long long s_sum[32];
  const int t = threadIdx.x;
  unsigned state = 0;                                       // inside a string at the start of the tile?
  unsigned long long x = 0, y = 0;                          // samples, series before the tile
  unsigned long long my_br = 0;                             // this thread's share of the '[' count
  ChunkSum zero;
  zero.parity = 0; zero.samples[0] = zero.samples[1] = zero.series[0] = zero.series[1] = zero.brackets[0] = zero.brackets[1] = 0;
  ChunkSum nxt = t < n_chunks ? sums[t] : zero;
  for (int64_t base = 0; base < n_chunks; base += 1024) {
    const int64_t c = base + t;
    const ChunkSum cs = nxt;
    nxt = c + 1024 < n_chunks ? sums[c + 1024] : zero;      // the next tile's load flies during this tile's scans
    unsigned tile_par;
    const unsigned h = state ^ cta_xor_excl(cs.parity, s_par, &tile_par);      // state at the start of chunk c
    unsigned long long tot;
    const unsigned long long before =
        cta_sum_excl((unsigned long long)(h ? cs.samples[1] : cs.samples[0]) |
                     ((unsigned long long)(h ? cs.series[1] : cs.series[0]) << 32), s_sum, &tot);
    my_br += h ? cs.brackets[1] : cs.brackets[0];
    if (c < n_chunks) {
      ChunkBase o;
      o.in_string = h; o.pad = 0;
      o.samples = (long long)(x + (before & 0xffffffffULL)); o.series = (long long)(y + (before >> 32));
      bases[c] = o;
    }
    state ^= tile_par;
    x += tot & 0xffffffffULL; y += tot >> 32;
  }
  unsigned long long z;
  cta_sum_excl(my_br, s_sum, &z);
  if (t == 0) {
    // envelope: Prometheus' own encoding of a successful matrix / vector answer
    const char pre[] = "{\"status\":\"success\",\"data\":{\"resultType\":\"";
    const int np = (int)sizeof(pre) - 1;
    int err = 0; long long where = 0;
    bool okp = len >= (uint64_t)np + 8;
    for (int i = 0; okp && i < np; ++i) okp = body[i] == (unsigned char)pre[i];
    int kind = -1;                                          // 0 matrix, 1 vector
    if (okp) {
      if (body[np] == 'm' && body[np + 1] == 'a' && body[np + 5] == 'x') kind = 0;
      else if (body[np] == 'v' && body[np + 1] == 'e' && body[np + 5] == 'r') kind = 1;
    }
    if (kind < 0) err = ERR_ENVELOPE;
    // every '[' outside a string is the result array, a "values" array or a sample.  Any other
    // count means samples this scan did not see (white space after '[', a shape other than
    // [t,"v"]) or further arrays ("warnings", "infos"): such bodies go to the host tokenizer.
    if (!err) {
      const unsigned long long expect = kind == 0 ? x + y + 1 : x + 1;
      if (z != expect) { err = ERR_BRACKETS; where = (long long)z; }
    }
    if (!err && ((long long)x > cap_samples || (long long)y > cap_series)) err = ERR_CAPACITY;
    status[ST_SAMPLES] = (long long)x; status[ST_SERIES] = (long long)y; status[ST_ERROR] = err; status[ST_WHERE] = where;
    if (!err) series_off[y] = (long long)x;
  }
}

__device__ __forceinline__ void report(int64_t* status, int err, uint64_t where) {
  if (atomicCAS((unsigned long long*)&status[ST_ERROR], 0ULL, (unsigned long long)err) == 0ULL)
    status[ST_WHERE] = (long long)where;
}

__global__ void __launch_bounds__(TT)
tokd_emit_kernel(const uint8_t* __restrict__ body, uint64_t len, const ChunkBase* __restrict__ bases, int64_t n_chunks,
                 uint32_t* __restrict__ ts_off, uint32_t* __restrict__ val_off,
                 int64_t* __restrict__ series_off, uint32_t* __restrict__ metric_off,
                 uint32_t* __restrict__ metric_len, int64_t cap_samples, int64_t cap_series,
                 int64_t* __restrict__ status) {
  const int lane = threadIdx.x & 31;
  const int64_t chunk = (int64_t)blockIdx.x * WARPS + (threadIdx.x >> 5);
  if (chunk >= n_chunks) return;                            // whole warps leave
  // The scan may have refused the body, and other warps may report an error while this one starts:
  // one lane looks, the warp leaves together (the ballots below need all lanes).
  int refused = lane == 0 ? (int)(status[ST_ERROR] != 0) : 0;
  refused = __shfl_sync(FULL, refused, 0);
  if (refused) return;
  const uint64_t c0 = (uint64_t)chunk * CHUNK;
  const ChunkBase cb = bases[chunk];
  unsigned run = cb.in_string;
  long long smp_base = cb.samples, ser_base = cb.series;
  for (int r = 0; r < ROUNDS; ++r) {
    const uint64_t base = c0 + (uint64_t)r * ROUND;
    if (base >= len) break;                                 // uniform
    const uint64_t pos = base + (uint64_t)lane * SEG;
    uint32_t w[4];
    load_words(body, len, pos, w);
    const unsigned q = quote_mask(body, len, pos, w);
    const unsigned in0 = inside_mask(q);
    const unsigned ballot = __ballot_sync(FULL, (in0 >> 16) & 1u);
    const unsigned before = run ^ (__popc(ballot & ((1u << lane) - 1u)) & 1u);
    run ^= __popc(ballot) & 1u;
    const unsigned ins = (in0 ^ (before ? 0xffffu : 0u)) & 0xffffu;     // inside a string
    const unsigned open = open_mask(w) & ~ins;
    unsigned next0 = __shfl_down_sync(FULL, w[0] & 0xffu, 1);
    if (lane == 31) next0 = byte_at(body, len, pos + SEG);
    // commas of this segment and the next one: the ',' behind a timestamp lies there (a timestamp
    // longer than that is looked for byte by byte)
    unsigned commas = eq16(w, 0x2c2c2c2cu);
    {
      unsigned nxt = __shfl_down_sync(FULL, commas, 1);
      if (lane == 31) {
        uint32_t v[4];
        load_words(body, len, pos + SEG, v);
        nxt = eq16(v, 0x2c2c2c2cu);
      }
      commas |= nxt << 16;
    }
    unsigned smask = 0, rmask = 0;                          // sample starts / series starts in this segment
    if (open) {
      const unsigned brace = open & bit5_16(w);
      unsigned m = open & ~brace;
      while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        if (starts_number(k + 1 < SEG ? seg_byte(w, k + 1) : next0)) smask |= 1u << k;
      }
      m = brace;
      while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        if (series_at(body, len, pos + k)) rmask |= 1u << k;
      }
    }
    // exclusive sums over the warp of both counts, packed 16 + 16
    const unsigned mine = (unsigned)__popc(smask) | ((unsigned)__popc(rmask) << 16);
    unsigned inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned u = __shfl_up_sync(FULL, inc, d); if (lane >= d) inc += u; }
    const unsigned tot = __shfl_sync(FULL, inc, 31), packed = inc - mine;
    long long si = smp_base + (packed & 0xffffu), ri = ser_base + (packed >> 16);
    unsigned m = smask | rmask;
    while (m) {
      const int k = __ffs(m) - 1;
      m &= m - 1;
      const uint64_t p = pos + k;
      if ((smask >> k) & 1u) {
        // [<number>,"<value>"] : the timestamp token ends at the ',', the value token begins
        // behind the quote that follows it
        const unsigned ahead = commas >> (k + 2);           // k + 2 <= 17
        uint64_t e;
        if (ahead) e = p + 2 + (uint64_t)(__ffs(ahead) - 1);
        else {
          e = pos + 2 * SEG;
          if (e < p + 2) e = p + 2;
          while (e < len && e < p + 48 && body[e] != ',') ++e;
        }
        if (e + 1 >= len || body[e] != ',' || body[e + 1] != '"') { report(status, ERR_SAMPLE, p); }
        else if (si >= 0 && si < cap_samples) { ts_off[si] = (uint32_t)(p + 1); val_off[si] = (uint32_t)(e + 2); }
        ++si;
      } else {
        // {"metric":{...},...} : the label object begins 10 bytes in and ends at the first '}'
        // outside a string (label values are strings, the object does not nest)
        const uint64_t m0 = p + 10;
        uint64_t e = m0 + 1;
        bool in_str = false;
        while (e < len) {
          const unsigned char ch = body[e];
          if (in_str) { if (ch == '\\') ++e; else if (ch == '"') in_str = false; }
          else if (ch == '"') in_str = true;
          else if (ch == '}') break;
          ++e;
        }
        if (e >= len) report(status, ERR_LABELS, p);
        if (ri >= 0 && ri < cap_series) {                   // the counts were checked against the capacities; a
          series_off[ri] = si;                              // body that changes under the kernel must still not
          metric_off[ri] = (uint32_t)m0;                    // write past the arrays
          metric_len[ri] = (uint32_t)(e - m0 + 1);
        }
        ++ri;
      }
    }
    smp_base += tot & 0xffffu;
    ser_base += tot >> 16;
  }
}

}  // namespace

extern "C" {

int pmx_tokenize_device(pmx_ctx* ctx, const uint8_t* d_body, uint64_t len,
                        uint32_t* d_ts_off, uint32_t* d_val_off, int64_t cap_samples,
                        int64_t* d_series_off, uint32_t* d_metric_off, uint32_t* d_metric_len,
                        int64_t cap_series, int64_t* d_status, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(d_body && d_ts_off && d_val_off && d_series_off && d_metric_off && d_metric_len && d_status, "NULL buffer");
  PMX_REQUIRE(len > 0 && len <= 0xffffffffULL, "body must hold 1 byte .. 4 GiB (uint32 offsets); split the query range");
  PMX_REQUIRE((reinterpret_cast<uintptr_t>(d_body) & 15) == 0, "d_body must be 16-byte aligned");
  PMX_REQUIRE(cap_samples >= 0 && cap_series >= 0, "negative capacity");
  cudaStream_t st = pmx_st(stream);
  const int64_t n_chunks = (int64_t)((len + CHUNK - 1) / CHUNK);
  const size_t sums_bytes = ((size_t)n_chunks * sizeof(ChunkSum) + 255) & ~size_t(255);
  unsigned char* scr = (unsigned char*)pmx_scratch(ctx, PMX_SCR_TOKENIZE, sums_bytes + (size_t)n_chunks * sizeof(ChunkBase));
  if (!scr) return PMX_E_CUDA;
  ChunkSum* sums = (ChunkSum*)scr;
  ChunkBase* bases = (ChunkBase*)(scr + sums_bytes);
  const unsigned grid = (unsigned)((n_chunks + WARPS - 1) / WARPS);
  tokd_count_kernel<<<grid, TT, 0, st>>>(d_body, len, sums, n_chunks);
  PMX_LAUNCH_CHECK();
  tokd_scan_kernel<<<1, 1024, 0, st>>>(d_body, len, sums, bases, n_chunks, cap_samples, cap_series, d_series_off, d_status);
  PMX_LAUNCH_CHECK();
  tokd_emit_kernel<<<grid, TT, 0, st>>>(d_body, len, bases, n_chunks, d_ts_off, d_val_off, d_series_off, d_metric_off,
                                                    d_metric_len, cap_samples, cap_series, d_status);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
