/* pmx.h -- C ABI of libpmx.so, the B200-native Prometheus aggregate engine.
 *
 * Drop-in boundary for the data path of k8s-rocm-metrics-dashboard.  The reference has no FFI
 * of its own (it is one Python file); every entry point below names the reference lines
 * (/root/reference/app.py) or the SURVEY.md section-8 row whose work it takes over.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no exceptions across the boundary.
 *   - every function returns 0 (PMX_OK) or a negative pmx_status; pmx_last_error() returns the
 *     message of the calling thread's last failure.
 *   - pointers named d_* are DEVICE pointers owned by the caller (e.g. torch tensors); the
 *     library never frees or retains them.  Pointers named h_* are host pointers.
 *   - every device entry point is asynchronous on `stream` (a cudaStream_t passed as void*);
 *     ctx scratch is stream-ordered, so one ctx must be driven from one stream at a time.
 *   - ts is int64 milliseconds; values are IEEE binary64; ragged series are CSR:
 *     d_off[s]..d_off[s+1] index the flat d_ts/d_val arrays (int64, n_series+1 entries).
 *   - [S,T] outputs are row-major, one row per series.  Output pointers may be NULL to skip
 *     that output unless stated otherwise.
 */
#ifndef PMX_H
#define PMX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PMX_VERSION 100                      /* 0.1.0 */
#define PMX_STALE_NAN_BITS 0x7FF0000000000002ULL   /* Prometheus staleness marker            */
#define PMX_QUANTILE_DIGIT_BITS 11           /* radix-select digit: 2048-bin histograms      */
#define PMX_QUANTILE_BINS 2048
#define PMX_QUANTILE_PASSES 6                /* 11+11+11+11+11+9 = 64 key bits               */
#define PMX_QUANTILE_GATHER_ROW 1024   /* uint64 words per segment and rank in pmx_quantile_tail_filter rows */

typedef enum pmx_status {
  PMX_OK = 0,
  PMX_E_INVALID = -1,      /* bad argument (NULL, negative size, misaligned pointer)          */
  PMX_E_CUDA = -2,         /* a CUDA runtime call failed; message holds cudaGetErrorString    */
  PMX_E_VALUE = -3,        /* malformed number  -> adapter raises ValueError (app.py:188)     */
  PMX_E_KEY = -4,          /* missing label     -> adapter raises KeyError  (app.py:186-187)  */
  PMX_E_CAPACITY = -5,     /* caller-provided output capacity too small                       */
  PMX_E_JSON = -6,         /* body is not a Prometheus success response (json.loads failure)  */
  PMX_E_LIMIT = -7         /* size beyond a documented limit of an entry point                */
} pmx_status;

typedef struct pmx_ctx pmx_ctx;             /* per (process, GPU): scratch + constant tables   */
typedef void* pmx_stream;                   /* cudaStream_t                                   */

int pmx_version(void);
const char* pmx_last_error(void);
/* Binds to CUDA device `device`; allocates nothing until an entry point needs scratch. */
int pmx_ctx_create(int device, pmx_ctx** out);
void pmx_ctx_destroy(pmx_ctx* ctx);
/* Bytes of device scratch the ctx currently holds (grow-only, reused across calls). */
size_t pmx_ctx_scratch_bytes(const pmx_ctx* ctx);

/* ---- X1  Prometheus JSON ingestion (HOST).  Replaces json.loads at app.py:163,178 for the
 * "result" array: finds every sample's timestamp / value token without building objects.
 * Works for resultType "matrix" ("values":[[t,"v"],...]) and "vector" ("value":[t,"v"]).
 *   h_ts_off[i], h_val_off[i] : byte offset of sample i's timestamp token (ends at ',') and of
 *                               its value token (first byte after the opening quote; ends at '"')
 *                               in the FLAT buffer = body followed by h_extra.
 *   h_series_off[s]           : CSR sample offsets (n_series+1 entries).
 *   h_metric_off/len[s]       : byte span of series s' "metric" object ({...}) in body.
 *   h_extra                   : value tokens that contained JSON escapes, decoded; the caller
 *                               appends h_extra[0..extra_len) to body to form the flat buffer.
 * Capacities: cap_samples >= len/8+1 and cap_series >= len/24+1 and cap_extra >= len always
 * suffice.  Returns PMX_E_JSON unless "status" is "success" and the shape is as above. */
typedef struct pmx_tokenize_result {
  int64_t n_samples, n_series, extra_len;
  int32_t result_type;                      /* 0 = matrix, 1 = vector                          */
} pmx_tokenize_result;
int pmx_tokenize(const char* h_body, size_t len,
                 uint32_t* h_ts_off, uint32_t* h_val_off, int64_t cap_samples,
                 int64_t* h_series_off, uint32_t* h_metric_off, uint32_t* h_metric_len,
                 int64_t cap_series, char* h_extra, size_t cap_extra,
                 pmx_tokenize_result* res);
/* Same result with the result array cut at series boundaries and walked by n_threads host threads
 * (bodies >= 1 MiB in Prometheus' canonical envelope; everything else is passed to pmx_tokenize).
 * Decoded escaped tokens keep their raw length in h_extra, so extra_len may be larger. */
int pmx_tokenize_mt(const char* h_body, size_t len,
                    uint32_t* h_ts_off, uint32_t* h_val_off, int64_t cap_samples,
                    int64_t* h_series_off, uint32_t* h_metric_off, uint32_t* h_metric_len,
                    int64_t cap_series, char* h_extra, size_t cap_extra, int n_threads,
                    pmx_tokenize_result* res);

/* The interned label table (replaces the dicts json.loads builds at app.py:178 and
 * app.py:186-194 reads).  For the n_series "metric" objects whose byte spans the tokenizers return,
 * every distinct string -- label name or value, JSON escapes decoded -- gets an id in order of first
 * appearance: string i is h_strings[h_str_off[i] .. h_str_off[i+1]) (UTF-8).  Series s has the pairs
 * h_pair_off[s] .. h_pair_off[s+1] of (h_key_id, h_val_id), in the order of its object.  A label
 * column (all series' gpu_id, say) is then one pass over int32 pairs.  HOST code.
 * Capacities: cap_pairs >= total labels, cap_strings >= bytes of distinct strings (len always
 * suffices), cap_distinct + 1 entries in h_str_off; PMX_E_CAPACITY names what was too small. */
typedef struct pmx_labels_result { int64_t n_pairs, n_strings, strings_len; } pmx_labels_result;
int pmx_labels_intern(const char* h_body, size_t len, const uint32_t* h_metric_off,
                      const uint32_t* h_metric_len, int64_t n_series, int64_t* h_pair_off,
                      int32_t* h_key_id, int32_t* h_val_id, int64_t cap_pairs, char* h_strings,
                      size_t cap_strings, int64_t* h_str_off, int64_t cap_distinct,
                      pmx_labels_result* res);

/* The same cut on the GPU, for a body that is already in device memory (the
 * host then never walks the text).  Accepts Prometheus' own compact encoding of a successful
 * "matrix" or "vector" answer: no white space between tokens, every sample written [<number>,"<text>"],
 * every series object beginning {"metric":{ .  Offsets have the meaning given above and index
 * d_body itself (value tokens with JSON escapes are not decoded: pmx_parse_f64 then reports them).
 * d_body must be 16-byte aligned and at most 4 GiB.  d_status (int64[4], device) receives
 * [0] n_samples, [1] n_series, [2] 0 or an error (1 envelope not recognised, 2 '[' count does not
 * match samples + series: white space, another shape, or "warnings"/"infos" arrays -- use
 * pmx_tokenize for those, 3 capacity, 4 malformed sample, 5 unterminated label object), [3] a byte
 * position or count for the error; outputs are valid only if [2] == 0.
 * d_series_off needs cap_series + 1 entries.  Uses ctx scratch (36 bytes per 64 KiB of body). */
int pmx_tokenize_device(pmx_ctx* ctx, const uint8_t* d_body, uint64_t len,
                        uint32_t* d_ts_off, uint32_t* d_val_off, int64_t cap_samples,
                        int64_t* d_series_off, uint32_t* d_metric_off, uint32_t* d_metric_len,
                        int64_t cap_series, int64_t* d_status, pmx_stream stream);

/* ---- X2 / H2  value decode (K1).  Replaces float(metric["value"][1]) at app.py:188: decimal
 * text -> binary64, correctly rounded, bit-exact with CPython float() for ASCII input (sign,
 * digits with single underscores, '.', exponent, inf/infinity/nan in any case, surrounding
 * ASCII whitespace).  d_first_bad (device int64, required) receives the smallest index of a
 * malformed token, or a number >= n (0x7f7f7f7f7f7f7f7f) when there is none; the adapter turns an
 * index into ValueError.  d_text should be 16-byte aligned (then the common token shapes are
 * decoded from 16-byte loads; any alignment is accepted). */
int pmx_parse_f64(pmx_ctx* ctx, const uint8_t* d_text, uint64_t text_len,
                  const uint32_t* d_val_off, int64_t n, double* d_out,
                  int64_t* d_first_bad, pmx_stream stream);
/* Timestamp tokens (JSON numbers, seconds, at most 3 decimals) -> exact int64 ms. */
int pmx_parse_ts_ms(pmx_ctx* ctx, const uint8_t* d_text, uint64_t text_len,
                    const uint32_t* d_ts_off, int64_t n, int64_t* d_out,
                    int64_t* d_first_bad, pmx_stream stream);
/* Both tokens of every sample in one launch (what a matrix body needs: app.py:188 reads the value,
 * the range path also the timestamp).  d_first_bad: int64[2], [0] for values, [1] for timestamps,
 * as above.  Either output (with its offsets) may be NULL. */
int pmx_parse_samples(pmx_ctx* ctx, const uint8_t* d_text, uint64_t text_len,
                      const uint32_t* d_ts_off, const uint32_t* d_val_off, int64_t n,
                      int64_t* d_out_ts, double* d_out_val, int64_t* d_first_bad, pmx_stream stream);

/* ---- X3  step-aligned resample (K2).  grid g_j = start + j*step.  For each series and j:
 * idx = (last sample with ts <= g_j); valid iff it exists, g_j - ts[idx] < lookback and the
 * sample is not the stale marker; invalid -> NaN / -1.  d_out_idx is relative to the series. */
int pmx_align(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_off,
              int64_t n_series, int64_t start_ms, int64_t step_ms, int64_t n_steps,
              int64_t lookback_ms, double* d_out_val, int32_t* d_out_idx, pmx_stream stream);

/* ---- X3, second form: bucket down-sampling (the reference never reads a
 * timestamp -- app.py:188 takes value[1] only -- so this is the survey's spec, pinned to
 * Series.resample(...).last()/.mean()/.count() in tests/golden/range_golden.npz).
 * Bin j covers [start_ms + j*step_ms, start_ms + (j+1)*step_ms), labelled by its left edge.  Per bin:
 * d_last = the last non-NaN sample (NaN if none), d_mean = mean of the non-NaN samples summed in
 * time order, d_count = their number.  Stale markers are NaNs here.  Any output may be NULL.
 * At most 65535 series per call. */
int pmx_bucket(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_off,
               int64_t n_series, int64_t start_ms, int64_t step_ms, int64_t n_steps,
               double* d_last, double* d_mean, int32_t* d_count, pmx_stream stream);

/* ---- X4  rolling min/max/mean/count (K3) over the trailing `window_steps` points of a dense
 * [S,T] matrix (pandas rolling('<W>') on a regular index: window (t-W, t], NaN skipped,
 * +-inf excluded from min/max/mean but counted, stat = NaN below min_periods finite points). */
int pmx_rolling(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                int64_t window_steps, int64_t min_periods,
                double* d_min, double* d_max, double* d_mean, int32_t* d_count,
                pmx_stream stream);
/* Ragged form: window (t_i - window_ms, t_i] evaluated at each sample of each series. */
int pmx_rolling_ragged(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val,
                       const int64_t* d_off, int64_t n_series, int64_t n_samples,
                       int64_t window_ms, int64_t min_periods,
                       double* d_min, double* d_max, double* d_mean, int32_t* d_count,
                       pmx_stream stream);

/* ---- X5  counter increase / rate with reset handling (K4).  Window as in pmx_rolling; NaN
 * skipped; >=2 samples needed.  increase = (x_last - x_first) + sum of the values that
 * preceded each drop inside the window.  mode 0: rate = increase / elapsed seconds first..last.
 * mode 1: Prometheus' extrapolatedRate -- the increase is stretched to the window edges when they
 * are closer than 1.1 mean sample spacings (else by half a spacing), not beyond the counter's zero
 * crossing, and rate = extrapolated increase / window seconds (d_increase is the extrapolated one). */
int pmx_rate(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
             int64_t window_steps, int64_t step_ms, int mode,
             double* d_increase, double* d_rate, int32_t* d_n_resets, pmx_stream stream);
int pmx_rate_ragged(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val,
                    const int64_t* d_off, int64_t n_series, int64_t n_samples,
                    int64_t window_ms, int mode,
                    double* d_increase, double* d_rate, int32_t* d_n_resets, pmx_stream stream);

/* ---- X3+X4+X5 fused (the roofline kernel): reads (ts,val) once, writes the aligned value,
 * rolling min/max/mean and rate of every grid point: 16 B in, 40 B out per sample. */
int pmx_fused_align_rolling_rate(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val,
                                 const int64_t* d_off, int64_t n_series,
                                 int64_t start_ms, int64_t step_ms, int64_t n_steps,
                                 int64_t lookback_ms, int64_t window_steps, int64_t min_periods,
                                 double* d_aligned, double* d_min, double* d_max,
                                 double* d_mean, double* d_rate, pmx_stream stream);

/* ---- X6 / H7-H9  group-by reductions (K5).  Groups are given as a CSR over series:
 * d_perm[d_goff[g] .. d_goff[g+1]) are the series of group g in ascending series order
 * (pmx_group_index builds both on the host from a group id per series; id < 0 = no group).
 * per_step = 0: reduce over all steps -> [G];  per_step = 1: reduce over series only -> [G,T].
 * NaN skipped; count is int64; empty group -> sum 0, count 0, min/max/mean NaN.  Sums visit
 * series in group order (per_step: strictly sequential, the order pandas uses at app.py:218). */
int pmx_group_index(const int32_t* h_group_id, int64_t n_series, int64_t n_groups,
                    int32_t* h_perm, int64_t* h_goff);
int pmx_group_reduce(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                     const int32_t* d_perm, const int64_t* d_goff, int64_t n_groups,
                     int per_step, double* d_sum, int64_t* d_count, double* d_min,
                     double* d_max, double* d_mean, pmx_stream stream);

/* ---- pmx_group_reduce(per_step = 0) without any device allocation.  No counterpart in the
 * reference: it exists for callers that cannot allocate when they call -- the GPU's memory is
 * already handed out in full to another allocator (a late cudaMalloc may fail), or the implicit
 * device-wide synchronisation of cudaMalloc/cudaFree does not fit a refresh loop.  The per-series
 * partials live in a fixed 2 MiB array inside the library (held by every GPU that loads it):
 * at most 65536 series (PMX_E_LIMIT beyond); no ctx; the first call on a device looks the array's
 * address up once and should be made outside stream capture; all calls share the array, so calls
 * that overlap on the GPU race.  The kernels run on the CURRENT device: make the device that owns
 * `stream` and the buffers current first (a stream of another device is a CUDA error here, not a
 * diagnosis).  Results equal pmx_group_reduce's. */
int pmx_group_reduce_noalloc(const double* d_x, int64_t n_series, int64_t n_steps,
                             const int32_t* d_perm, const int64_t* d_goff, int64_t n_groups,
                             double* d_sum, int64_t* d_count, double* d_min, double* d_max,
                             double* d_mean, pmx_stream stream);

/* ---- group partials across ranks.  Per-card groups are rank-local; per-metric groups
 * span ranks (the reference's own per-column statistics, app.py:216-221, seen cluster-wide).
 * pmx_group_pack writes one rank's partials as 4 x n doubles (sum, count, min or +inf, max or
 * -inf); after ONE all-gather of those buffers pmx_group_combine folds d_gathered[n_ranks][4][n] in
 * rank order into the outputs of pmx_group_reduce (empty group: sum 0, count 0, NaN).  Every rank
 * folds the same data in the same order, so all ranks hold identical bits. */
int pmx_group_pack(const double* d_sum, const int64_t* d_count, const double* d_min,
                   const double* d_max, int64_t n, double* d_packed, pmx_stream stream);
int pmx_group_combine(const double* d_gathered, int n_ranks, int64_t n, double* d_sum,
                      int64_t* d_count, double* d_min, double* d_max, double* d_mean,
                      pmx_stream stream);

/* ---- H5 / H6 / H9  the per-card steps of the instant path.
 * pmx_pivot_gather: long -> wide (df.pivot at app.py:207): out[cell] = val[src[cell]] or NaN when
 *   src[cell] < 0 (missing cell).  Duplicate (gpu_id, metric) keys are detected by the adapter
 *   from the label table and raised as ValueError before this call.
 * pmx_ratio_pct: out = num/den*100 with the two roundings of Python's (used/total)*100
 *   (app.py:210-212); *d_zero_flag becomes 1 if any denominator is 0 so the adapter can raise
 *   ZeroDivisionError.  Strides are in elements.
 * pmx_where_gt: out = x > threshold ? x : NaN  (app.py:343 power_values[power_values > 0]). */
int pmx_pivot_gather(pmx_ctx* ctx, const double* d_val, const int32_t* d_cell_src, int64_t n_cells,
                     double* d_out, pmx_stream stream);
int pmx_ratio_pct(pmx_ctx* ctx, const double* d_num, int64_t num_stride, const double* d_den,
                  int64_t den_stride, int64_t n, double* d_out, int64_t out_stride,
                  int32_t* d_zero_flag, pmx_stream stream);
int pmx_where_gt(pmx_ctx* ctx, const double* d_x, int64_t n, double threshold, double* d_out,
                 pmx_stream stream);

/* ---- Series cache (no counterpart in the reference, which re-processes the
 * whole range on every refresh, app.py:24/:486).  Samples of series s live ascending in row s of a
 * [n_series][cap] slab, d_count[s] of them held.  pmx_slab_tail_lengths: d_first[s] = number of held
 * samples with ts <= threshold_ms, d_len[s] = the rest.  With d_off = exclusive cumulative sum of
 * d_len (n_series + 1 entries, made by the caller), pmx_slab_tail_gather copies every series' tail
 * to d_out_*[d_off[s] ..): a CSR that pmx_fused_align_rolling_rate takes as it is. */
int pmx_slab_tail_lengths(pmx_ctx* ctx, const int64_t* d_ts, const int64_t* d_count, int64_t n_series, int64_t cap,
                          int64_t threshold_ms, int64_t* d_first, int64_t* d_len, pmx_stream stream);
/* pmx_slab_tail_lengths and the cumulative sum in one launch: d_first[s] as above, d_off[0 .. n_series] =
 * exclusive sums of the tail lengths (the CSR offsets pmx_slab_tail_gather and the fused kernel take).
 * d_work: one uint32 that is zero before the first call; the call leaves it zero. */
int pmx_slab_tail_offsets(pmx_ctx* ctx, const int64_t* d_ts, const int64_t* d_count, int64_t n_series, int64_t cap,
                          int64_t threshold_ms, int64_t* d_first, int64_t* d_off, uint32_t* d_work, pmx_stream stream);
int pmx_slab_tail_gather(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_first,
                         const int64_t* d_off, int64_t n_series, int64_t cap, int64_t* d_out_ts, double* d_out_val,
                         pmx_stream stream);

/* pmx_slab_append: the decoded samples of a new answer (CSR d_new_ts / d_new_val / d_in_off over n_in
 * series, d_in_row[i] = slab row of series i or -1 to ignore it) are appended to their rows: samples
 * not newer than the row's newest are skipped (answers may overlap), a full row drops its oldest.
 * If a dropped sample is newer than needed_from_ms (a grid point not yet produced still needs it)
 * NOTHING is appended.  d_status (int64[2]): [0] rows that would have lost such a sample (0 = the
 * append happened), [1] samples stored.  d_work: 2 * n_in int64 of scratch owned by the caller.
 * pmx_csr_copy: series s of (d_ts, d_val, d_off) is copied to d_out_*[d_dst[s] ..) -- how the range
 * client concatenates the answers to consecutive sub-ranges per label set. */
int pmx_slab_append(pmx_ctx* ctx, int64_t* d_ts, double* d_val, int64_t* d_count, int64_t cap,
                    const int64_t* d_new_ts, const double* d_new_val, const int64_t* d_in_off,
                    const int64_t* d_in_row, int64_t n_in, int64_t needed_from_ms, int64_t* d_work,
                    int64_t* d_status, pmx_stream stream);
int pmx_csr_copy(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_off,
                 const int64_t* d_dst, int64_t n_series, int64_t* d_out_ts, double* d_out_val,
                 pmx_stream stream);

/* ---- X7  exact quantiles (K6) by radix select over order-preserving 64-bit keys.
 * Values of series s belong to segment d_seg[s] (or none if < 0).  State per segment:
 *   d_state: opaque uint64 [n_seg * (3 + 4*n_q)] owned by the caller, zeroed before pass 0 or
 *   filled by pmx_quantile_init.
 * One pass p = 0..PMX_QUANTILE_PASSES-1 is: pmx_quantile_hist (local histograms of the p-th
 * digit of the keys that match each target's prefix) -> [all-reduce d_hist across ranks, sum,
 * uint64] -> pmx_quantile_pick (device-side prefix scan: choose the bin holding each target
 * rank, extend the prefix).  After the last pass pmx_quantile_finish converts the keys back
 * and interpolates like numpy's _lerp (Series.quantile / np.percentile 'linear').
 * d_hist is uint64 [n_seg * 2*n_q * PMX_QUANTILE_BINS] (pass 0 uses the first n_seg*BINS).
 * d_row_order (optional, n_rows entries): the rows to visit, grouped by segment -- the d_perm of
 * pmx_group_index serves; a CTA then switches its private histograms rarely.  NULL = all rows in
 * storage order.
 * Pass 2 (the list pass) reads the data only to keep the keys that still match a target in ctx
 * scratch (<= 2^20 per segment); its digit and those of passes 3-5 are read from those lists; one
 * selection at a time per ctx, passes in order 0..5. */
int pmx_quantile_hist(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                      const int32_t* d_seg, int64_t n_seg,
                      const int32_t* d_row_order, int64_t n_rows, int64_t n_q, int pass,
                      const uint64_t* d_state, uint64_t* d_hist, pmx_stream stream);
/* The pass that builds the candidate lists is 2 by default; pmx_quantile_hist_lp lets the caller move
 * it to 3 (one more full pass, far fewer candidates), which pmx_quantile_list_pass recommends for
 * segments of more than ~6 M values.  All ranks of a selection must use the same list_pass, and the
 * tail calls (from_pass = list_pass + 1 or later) follow it. */
int pmx_quantile_list_pass(int64_t n_series, int64_t n_steps, int64_t n_seg);
int pmx_quantile_hist_lp(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                         const int32_t* d_seg, int64_t n_seg,
                         const int32_t* d_row_order, int64_t n_rows, int64_t n_q, int pass, int list_pass,
                         const uint64_t* d_state, uint64_t* d_hist, pmx_stream stream);
/* Optional start from a group reduction.  With the count, minimum and maximum of every segment over
 * ALL ranks (pmx_group_reduce on the same values, combined across ranks) the keys are taken
 * relative to key(min) and the leading bits that key(max) - key(min) leaves zero are skipped, so
 * pass 0 is not needed: after this call run passes 1..PMX_QUANTILE_PASSES (hist -> all-reduce ->
 * pick) instead of 0..PASSES-1.  A pass does nothing for a segment whose keys are already resolved
 * (min == max resolves at once; a range within a factor of two of one sign needs passes 1..5).
 * d_min / d_max of an empty segment are ignored. */
int pmx_quantile_init(pmx_ctx* ctx, const int64_t* d_count, const double* d_min, const double* d_max,
                      int64_t n_seg, const double* d_q, int64_t n_q, uint64_t* d_state, pmx_stream stream);
/* Single GPU only (the candidate lists of pass 2 are local): finish every segment in one launch --
 * all passes from_pass..PMX_QUANTILE_PASSES, one CTA per segment, from the segment's candidate list;
 * a segment whose list overflowed (very many repeated values) is finished by its CTA from d_x
 * itself, slowly.  Call it after pass 2's pick with from_pass = 3; no further pass is needed. */
int pmx_quantile_tail(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg,
                      int64_t n_seg, int64_t n_q, int from_pass, uint64_t* d_state, pmx_stream stream);
/* The same for several GPUs, after pass 2's pick: every rank writes the keys of its lists that match
 * a target as it stands now (d_rows: uint64 [n_seg * PMX_QUANTILE_GATHER_ROW], row = count, keys),
 * the rows are all-gathered ([n_ranks][n_seg][ROW], rank-major), and pmx_quantile_tail_merge settles
 * every segment from the union on every rank alike.  *d_unresolved (device int32) counts the
 * segments that did not fit (many repeated values); if it is 0 -- the same on all ranks -- passes
 * from_pass.. and their all-reduces can be left out, otherwise they run as usual. */
int pmx_quantile_tail_filter(pmx_ctx* ctx, int64_t n_seg, int64_t n_q, int from_pass, const uint64_t* d_state,
                             uint64_t* d_rows, pmx_stream stream);
int pmx_quantile_tail_merge(pmx_ctx* ctx, const uint64_t* d_rows_all, int n_ranks, int64_t n_seg, int64_t n_q,
                            int from_pass, uint64_t* d_state, int32_t* d_unresolved, pmx_stream stream);
int pmx_quantile_pick(pmx_ctx* ctx, const uint64_t* d_hist, int64_t n_seg, const double* d_q,
                      int64_t n_q, int pass, uint64_t* d_state, pmx_stream stream);
int pmx_quantile_finish(pmx_ctx* ctx, const uint64_t* d_state, int64_t n_seg, const double* d_q,
                        int64_t n_q, int64_t* d_n, int64_t* d_lo, int64_t* d_hi,
                        double* d_x_lo, double* d_x_hi, double* d_value, pmx_stream stream);
/* One GPU: the per-segment summary of pmx_group_reduce(per_step = 0) AND the exact quantiles of
 * pmx_quantile_select from ONE sweep of d_x (plus a sample of 1/64 of it).  Replaces, for the range
 * matrices, what app.py:216-221 and :335-345 compute for the instant vector (mean / max / min per
 * metric) together with SURVEY X7's Series.quantile -- the reference has no counterpart of the
 * quantiles.  A sample fixes a bracket of two data values around each quantile; the sweep forms
 * the per-series partial sums (same order as pmx_group_reduce: d_sum .. d_mean agree bit for bit),
 * counts what lies below / on the bounds and copies the keys strictly inside; the order statistics
 * are then selected from those few keys.  The result is exact whenever the counts confirm the
 * brackets; *d_failed (device int32) is the number of (segment, quantile) pairs for which they did
 * not -- a sample more than five standard deviations off, or more than 256 MiB / n_seg of keys
 * inside one segment's brackets -- and those pairs hold NaN: when it is not zero, use
 * pmx_group_reduce + the pmx_quantile_* passes instead.  d_perm / d_goff: the CSR over series of
 * pmx_group_reduce; d_seg[s] = group of series s (or -1); n_q in [1, 4] (else PMX_E_LIMIT).
 * Outputs as in pmx_group_reduce and pmx_quantile_finish (each may be NULL except d_count). */
int pmx_summary_quantiles(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                          const int32_t* d_seg, int64_t n_seg, const int32_t* d_perm, const int64_t* d_goff,
                          const double* d_q, int64_t n_q, double* d_sum, int64_t* d_count, double* d_min,
                          double* d_max, double* d_mean, int64_t* d_n, int64_t* d_lo, int64_t* d_hi,
                          double* d_x_lo, double* d_x_hi, double* d_value, int32_t* d_failed, pmx_stream stream);
/* pmx_summary_quantiles in two halves, for callers that have the series' raw samples before the
 * matrix exists (the aggregate step: the fused kernel is still producing the aligned matrix).
 * pmx_summary_brackets_csr takes the brackets from the raw samples (d_val / d_off: the CSR of
 * pmx_fused_align_rolling_rate; one sample per 64 grid steps at the same relative position) -- the
 * brackets are hints that the sweep verifies, so they need not come from the matrix itself; it can
 * run on another stream beside the fused kernel.  pmx_summary_sweep then does the rest on the
 * matrix (same arguments and results as pmx_summary_quantiles; same n_series, n_steps, n_seg, d_q,
 * n_q as the brackets call, which must have completed in stream order). */
int pmx_summary_brackets_csr(pmx_ctx* ctx, const double* d_val, const int64_t* d_off, int64_t n_series, int64_t n_steps,
                             int64_t n_seg, const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q,
                             pmx_stream stream);
int pmx_summary_sweep(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg, int64_t n_seg,
                      const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q, double* d_sum,
                      int64_t* d_count, double* d_min, double* d_max, double* d_mean, int64_t* d_n, int64_t* d_lo,
                      int64_t* d_hi, double* d_x_lo, double* d_x_hi, double* d_value, int32_t* d_failed, pmx_stream stream);
/* The same for several ranks (one process per GPU, series sharded; no counterpart in the
 * reference, which is one process): five calls with four small collectives between them, one sweep
 * of the data on every rank.  All ranks pass the same n_seg, d_q and n_q, and each its own rows
 * (n_series may be 0).  State between the calls lives in ctx scratch, so the five calls of one
 * selection must not interleave with another selection on the same ctx.
 *   pmx_bsel_begin   sample + this rank's brackets -> d_bounds [n_seg * bounds_w] f64
 *                    all-gather d_bounds -> [n_ranks][n_seg * bounds_w]
 *   pmx_bsel_sweep   hull of the brackets, the sweep, this rank's partial summary and counters and
 *                    the first digit of its keys inside every bracket
 *                    -> d_packed [n_seg * (pack_w + n_q * PMX_QUANTILE_BINS / 2)] 64-bit words
 *                    all-gather d_packed -> [n_ranks][...]
 *   pmx_bsel_hist    summary of all ranks in rank order (d_sum .. d_mean, as pmx_group_combine),
 *                    counters and first digits summed, every wanted rank placed, second digit of
 *                    this rank's keys in the buckets that hold them
 *                    -> d_hist2 [n_seg * n_q * 2 * PMX_QUANTILE_BINS] u32; all-reduce (sum), or
 *                    all-gather and let pmx_bsel_rows add the n_hist = n_ranks copies
 *   pmx_bsel_rows    second digit of every bucket; this rank's keys of the buckets (22 bits)
 *                    -> d_rows [n_seg * n_q * row_w] u64; all-gather -> [n_ranks][...]
 *   pmx_bsel_finish  outputs of pmx_quantile_finish; *d_failed as in pmx_summary_quantiles -- it
 *                    depends on combined data only, so every rank sees the same value.
 * pmx_bsel_widths returns bounds_w, pack_w, row_w. */
int pmx_bsel_widths(int* bounds_w, int* pack_w, int* row_w);
int pmx_bsel_begin(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg, int64_t n_seg,
                   const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q, double* d_bounds,
                   pmx_stream stream);
/* pmx_bsel_begin from the series' raw samples (as pmx_summary_brackets_csr): may run beside the fused kernel */
int pmx_bsel_begin_csr(pmx_ctx* ctx, const double* d_val, const int64_t* d_off, int64_t n_series, int64_t n_steps, int64_t n_seg,
                       const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q, double* d_bounds,
                       pmx_stream stream);
int pmx_bsel_sweep(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg, int64_t n_seg,
                   const int32_t* d_perm, const int64_t* d_goff, int64_t n_q, const double* d_bounds_all, int n_ranks,
                   uint64_t* d_packed, pmx_stream stream);
int pmx_bsel_hist(pmx_ctx* ctx, int64_t n_series, int64_t n_steps, int64_t n_seg, const double* d_q, int64_t n_q,
                  const uint64_t* d_packed_all, int n_ranks, double* d_sum, int64_t* d_count, double* d_min, double* d_max,
                  double* d_mean, uint32_t* d_hist2, pmx_stream stream);
int pmx_bsel_rows(pmx_ctx* ctx, int64_t n_series, int64_t n_steps, int64_t n_seg, int64_t n_q, const uint32_t* d_hist2,
                  int n_hist, uint64_t* d_rows, pmx_stream stream);
int pmx_bsel_finish(pmx_ctx* ctx, int64_t n_series, int64_t n_steps, int64_t n_seg, int64_t n_q, const uint64_t* d_rows_all,
                    int n_ranks, const int64_t* d_count, int64_t* d_n, int64_t* d_lo, int64_t* d_hi, double* d_x_lo,
                    double* d_x_hi, double* d_value, int32_t* d_failed, pmx_stream stream);
/* The all-gather of the pmx_bsel_* exchanges as stores into peer memory (no counterpart in the
 * reference).  d_peer_base: device array of n_ranks addresses -- the receive buffer of every rank as
 * mapped into THIS process (symmetric memory; Python: torch.distributed._symmetric_memory's
 * buffer_ptrs_dev).  Stores nbytes from d_src into [dst_offset_bytes + rank * nbytes, + nbytes) of
 * every rank's buffer, own included; sizes, offsets and d_src multiples of 16 bytes.  The caller
 * follows it with a barrier of all ranks before anyone reads, and does not write the same region
 * again before every rank has read it (the exchanges of one step use regions of their own, and a
 * rank that starts the next step has passed this step's last barrier).  Ordinary stores only. */
int pmx_peer_put(const void* d_src, const uint64_t* d_peer_base, int n_ranks, int rank, uint64_t dst_offset_bytes,
                 uint64_t nbytes, pmx_stream stream);
/* Single-GPU convenience: all passes back to back (no collective); hist/state in ctx scratch. */
int pmx_quantile_select(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                        const int32_t* d_seg, int64_t n_seg,
                        const int32_t* d_row_order, int64_t n_rows, const double* d_q, int64_t n_q,
                        int64_t* d_n, int64_t* d_lo, int64_t* d_hi,
                        double* d_x_lo, double* d_x_hi, double* d_value, pmx_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* PMX_H */
