// tokenize_device.cu -- X1 on the GPU: the response body is already in device
// memory (one pinned H2D copy of the bytes that came off the socket) and is cut into sample tokens
// there, so the host never walks it.  Replaces json.loads at app.py:163/178 for bodies in
// Prometheus' own compact encoding; anything else is reported, not guessed (the host tokenizer in
// tokenize.cu accepts general JSON).
//
// The body is a flat array of bytes; what a byte means depends on whether it lies inside a JSON
// string, i.e. on the parity of the unescaped quotes before it.  Three kernels:
//   count : CTA c walks chunk c (64 KiB) in rounds of 4 KiB, 16 bytes per thread per round, and
//           reports the chunk's quote parity and its number of sample starts ('[' followed by a
//           digit or '-'), series starts ('{"metric":{') and '[' -- each counted twice, once for
//           "the chunk begins outside a string" and once for "inside";
//   scan  : one CTA turns the parities into every chunk's true starting state and the counts that
//           belong to it into output offsets, and checks the envelope and the bracket balance;
//   emit  : the same walk with the state known; every sample start writes its two token offsets
//           (the thread scans forward to the ',' and checks the '"' behind it), every series start
//           its CSR offset and the span of its label object.
// Token offsets have the meaning pmx_tokenize gives them, so pmx_parse_f64 / pmx_parse_ts_ms take
// the body and these arrays as they are.
#include <stdint.h>
#include "pmx_common.cuh"

namespace {

constexpr int TT = 256;                 // threads per CTA
constexpr int SEG = 16;                 // bytes per thread per round
constexpr int ROUND = TT * SEG;         // 4 KiB
constexpr int ROUNDS = 16;              // rounds per chunk
constexpr int64_t CHUNK = (int64_t)ROUND * ROUNDS;

struct ChunkSum { uint32_t parity, samples[2], series[2], brackets[2]; };
struct ChunkBase { uint32_t in_string; uint32_t pad; int64_t samples, series; };

// status words written by the scan / emit kernels
enum { ST_SAMPLES = 0, ST_SERIES = 1, ST_ERROR = 2, ST_WHERE = 3 };
enum { ERR_ENVELOPE = 1, ERR_BRACKETS = 2, ERR_CAPACITY = 3, ERR_SAMPLE = 4, ERR_LABELS = 5 };

__device__ __forceinline__ unsigned char byte_at(const uint8_t* __restrict__ b, uint64_t len, uint64_t i) {
  return i < len ? b[i] : (unsigned char)0;
}

// The 16 bytes of this thread's segment (zeros past the end) and the mask of its unescaped quotes.
// A quote is escaped iff an odd run of backslashes ends right before it; runs are rare and short,
// so the look-back reads global memory.
__device__ __forceinline__ unsigned load_segment(const uint8_t* __restrict__ body, uint64_t len, uint64_t pos,
                                                 unsigned char* c) {
  uint32_t w[4] = {0, 0, 0, 0};
  if (pos + SEG <= len) {
    const uint4 v = *reinterpret_cast<const uint4*>(body + pos);
    w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
  } else {
    for (int k = 0; k < SEG; ++k)
      if (pos + k < len) w[k >> 2] |= (uint32_t)body[pos + k] << (8 * (k & 3));
  }
  unsigned qmask = 0;
#pragma unroll
  for (int k = 0; k < SEG; ++k) {
    c[k] = (unsigned char)(w[k >> 2] >> (8 * (k & 3)));
    if (c[k] == '"') qmask |= 1u << k;
  }
  unsigned m = qmask;
  while (m) {                                               // drop escaped quotes
    const int k = __ffs(m) - 1;
    m &= m - 1;
    uint64_t i = pos + k;
    int run = 0;
    while (i > 0 && body[i - 1] == '\\') { ++run; --i; }
    if (run & 1) qmask &= ~(1u << k);
  }
  return qmask;
}

// exclusive XOR over the CTA of one bit per thread; *total receives the CTA's XOR
__device__ __forceinline__ unsigned cta_xor_scan(unsigned bit, unsigned* s_warp, unsigned* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned ballot = __ballot_sync(0xffffffffu, bit);
  const unsigned before = __popc(ballot & ((1u << lane) - 1u)) & 1u;
  __syncthreads();                                          // s_warp may still be read from the last call
  if (lane == 0) s_warp[warp] = __popc(ballot) & 1u;
  __syncthreads();
  unsigned acc = 0, all = 0;
#pragma unroll
  for (int w = 0; w < TT / 32; ++w) { const unsigned v = s_warp[w]; if (w < warp) acc ^= v; all ^= v; }
  *total = all;
  return acc ^ before;
}

// exclusive sum over the CTA; *total receives the CTA's sum
__device__ __forceinline__ unsigned cta_sum_scan(unsigned v, unsigned* s_warp, unsigned* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const unsigned u = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += u; }
  __syncthreads();
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  unsigned acc = 0, all = 0;
#pragma unroll
  for (int w = 0; w < TT / 32; ++w) { const unsigned u = s_warp[w]; if (w < warp) acc += u; all += u; }
  *total = all;
  return acc + inc - v;
}

__device__ __forceinline__ bool starts_number(unsigned char c) { return (c >= '0' && c <= '9') || c == '-'; }

// '{' at p opens a series object iff it is followed by "metric":{
__device__ __forceinline__ bool series_at(const uint8_t* __restrict__ body, uint64_t len, uint64_t p) {
  const char pat[10] = {'"', 'm', 'e', 't', 'r', 'i', 'c', '"', ':', '{'};
  if (p + 10 >= len) return false;
#pragma unroll
  for (int k = 0; k < 10; ++k) if (body[p + 1 + k] != (unsigned char)pat[k]) return false;
  return true;
}

__global__ void __launch_bounds__(TT)
tokd_count_kernel(const uint8_t* __restrict__ body, uint64_t len, ChunkSum* __restrict__ sums) {
  __shared__ unsigned s_warp[TT / 32];
  const uint64_t c0 = (uint64_t)blockIdx.x * CHUNK;
  unsigned run = 0;                                         // quote parity of the chunk so far
  unsigned n_smp[2] = {0, 0}, n_ser[2] = {0, 0}, n_br[2] = {0, 0};
  for (int r = 0; r < ROUNDS; ++r) {
    const uint64_t pos = c0 + (uint64_t)r * ROUND + (uint64_t)threadIdx.x * SEG;
    if (c0 + (uint64_t)r * ROUND >= len) break;             // uniform
    unsigned char c[SEG];
    const unsigned qmask = pos < len ? load_segment(body, len, pos, c) : 0u;
    unsigned total;
    const unsigned before = run ^ cta_xor_scan(__popc(qmask) & 1u, s_warp, &total);
    if (pos < len) {
#pragma unroll
      for (int k = 0; k < SEG; ++k) {
        if (c[k] != '[' && c[k] != '{') continue;
        const unsigned par = before ^ (__popc(qmask & ((1u << k) - 1u)) & 1u);
        if (c[k] == '[') {
          ++n_br[par];
          const unsigned char nx = k + 1 < SEG ? c[k + 1] : byte_at(body, len, pos + SEG);
          if (starts_number(nx)) ++n_smp[par];
        } else if (series_at(body, len, pos + k)) {
          ++n_ser[par];
        }
      }
    }
    run ^= total;
  }
  // CTA totals of the six counters
  unsigned tot[6];
  unsigned vals[6] = {n_smp[0], n_smp[1], n_ser[0], n_ser[1], n_br[0], n_br[1]};
  for (int i = 0; i < 6; ++i) cta_sum_scan(vals[i], s_warp, &tot[i]);
  if (threadIdx.x == 0) {
    ChunkSum s;
    s.parity = run; s.samples[0] = tot[0]; s.samples[1] = tot[1]; s.series[0] = tot[2]; s.series[1] = tot[3];
    s.brackets[0] = tot[4]; s.brackets[1] = tot[5];
    sums[blockIdx.x] = s;
  }
}

// One CTA: chunk states and output bases; envelope and balance checks; totals.
__global__ void __launch_bounds__(1024)
tokd_scan_kernel(const uint8_t* __restrict__ body, uint64_t len, const ChunkSum* __restrict__ sums,
                 ChunkBase* __restrict__ bases, int64_t n_chunks, int64_t cap_samples, int64_t cap_series,
                 int64_t* __restrict__ series_off, int64_t* __restrict__ status) {
  __shared__ unsigned s_par[1024];
  __shared__ unsigned long long s_smp[1024], s_ser[1024], s_br[1024];
  const int t = threadIdx.x;
  const int64_t per = (n_chunks + 1023) / 1024;
  const int64_t a = (int64_t)t * per, b = a + per < n_chunks ? a + per : n_chunks;
  unsigned par = 0;
  for (int64_t c = a; c < b; ++c) par ^= sums[c].parity;
  s_par[t] = par;
  __syncthreads();
  if (t == 0) { unsigned acc = 0; for (int i = 0; i < 1024; ++i) { const unsigned v = s_par[i]; s_par[i] = acc; acc ^= v; } }
  __syncthreads();
  unsigned h = s_par[t];                                    // state at the start of this thread's first chunk
  unsigned long long smp = 0, ser = 0, br = 0;
  for (int64_t c = a; c < b; ++c) { smp += sums[c].samples[h]; ser += sums[c].series[h]; br += sums[c].brackets[h]; h ^= sums[c].parity; }
  s_smp[t] = smp; s_ser[t] = ser; s_br[t] = br;
  __syncthreads();
  if (t == 0) {
    unsigned long long x = 0, y = 0, z = 0;
    for (int i = 0; i < 1024; ++i) {
      const unsigned long long u = s_smp[i], v = s_ser[i], w = s_br[i];
      s_smp[i] = x; s_ser[i] = y; s_br[i] = z; x += u; y += v; z += w;
    }
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
  __syncthreads();
  h = s_par[t];
  smp = s_smp[t]; ser = s_ser[t];
  for (int64_t c = a; c < b; ++c) {
    ChunkBase o; o.in_string = h; o.pad = 0; o.samples = (long long)smp; o.series = (long long)ser;
    bases[c] = o;
    smp += sums[c].samples[h]; ser += sums[c].series[h]; h ^= sums[c].parity;
  }
}

__device__ __forceinline__ void report(int64_t* status, int err, uint64_t where) {
  if (atomicCAS((unsigned long long*)&status[ST_ERROR], 0ULL, (unsigned long long)err) == 0ULL)
    status[ST_WHERE] = (long long)where;
}

__global__ void __launch_bounds__(TT)
tokd_emit_kernel(const uint8_t* __restrict__ body, uint64_t len, const ChunkBase* __restrict__ bases,
                 uint32_t* __restrict__ ts_off, uint32_t* __restrict__ val_off,
                 int64_t* __restrict__ series_off, uint32_t* __restrict__ metric_off,
                 uint32_t* __restrict__ metric_len, int64_t cap_samples, int64_t cap_series,
                 int64_t* __restrict__ status) {
  __shared__ unsigned s_warp[TT / 32];
  __shared__ int s_refused;
  // The scan may have refused the body, and other CTAs of this kernel may report an error while
  // this one starts: one thread looks, everybody leaves together (the scans below need all warps).
  if (threadIdx.x == 0) s_refused = status[ST_ERROR] != 0;
  __syncthreads();
  if (s_refused) return;
  const uint64_t c0 = (uint64_t)blockIdx.x * CHUNK;
  const ChunkBase cb = bases[blockIdx.x];
  unsigned run = cb.in_string;
  long long smp_base = cb.samples, ser_base = cb.series;
  for (int r = 0; r < ROUNDS; ++r) {
    const uint64_t pos = c0 + (uint64_t)r * ROUND + (uint64_t)threadIdx.x * SEG;
    if (c0 + (uint64_t)r * ROUND >= len) break;             // uniform
    unsigned char c[SEG];
    const unsigned qmask = pos < len ? load_segment(body, len, pos, c) : 0u;
    unsigned total;
    const unsigned before = run ^ cta_xor_scan(__popc(qmask) & 1u, s_warp, &total);
    unsigned smask = 0, rmask = 0;                          // sample starts / series starts in this segment
    if (pos < len) {
#pragma unroll
      for (int k = 0; k < SEG; ++k) {
        if (c[k] != '[' && c[k] != '{') continue;
        if (before ^ (__popc(qmask & ((1u << k) - 1u)) & 1u)) continue;      // inside a string
        if (c[k] == '[') {
          const unsigned char nx = k + 1 < SEG ? c[k + 1] : byte_at(body, len, pos + SEG);
          if (starts_number(nx)) smask |= 1u << k;
        } else if (series_at(body, len, pos + k)) {
          rmask |= 1u << k;
        }
      }
    }
    unsigned tot;
    const unsigned packed = cta_sum_scan((unsigned)__popc(smask) | ((unsigned)__popc(rmask) << 16), s_warp, &tot);
    long long si = smp_base + (packed & 0xffffu), ri = ser_base + (packed >> 16);
    unsigned m = smask | rmask;
    while (m) {
      const int k = __ffs(m) - 1;
      m &= m - 1;
      const uint64_t p = pos + k;
      if ((smask >> k) & 1u) {
        // [<number>,"<value>"] : the timestamp token ends at the ',', the value token begins
        // behind the quote that follows it
        uint64_t e = p + 2;
        while (e < len && e < p + 48 && body[e] != ',') ++e;
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
    run ^= total;
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
  tokd_count_kernel<<<(unsigned)n_chunks, TT, 0, st>>>(d_body, len, sums);
  PMX_LAUNCH_CHECK();
  tokd_scan_kernel<<<1, 1024, 0, st>>>(d_body, len, sums, bases, n_chunks, cap_samples, cap_series, d_series_off, d_status);
  PMX_LAUNCH_CHECK();
  tokd_emit_kernel<<<(unsigned)n_chunks, TT, 0, st>>>(d_body, len, bases, d_ts_off, d_val_off, d_series_off, d_metric_off,
                                                    d_metric_len, cap_samples, cap_series, d_status);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
