/* hga_b200.h — C ABI of libhga_b200.so: the B200 (sm_100a) implementation of the high-gamma-activity (HGA) front end
 * of cronelab/delayed-speech-synthesis.
 *
 * The reference has no C ABI for this path; its boundary is a Python call surface.  Every entry point
 * below names the reference interface it stands behind (paths relative to /root/reference).  The Python package
 * `delayed_speech_synthesis_b200` binds these with ctypes (see INTEGRATION.md for the stub a maintainer would add to
 * the reference).
 *
 * Conventions
 *   - plain C types only; every function returns an `int` status (HGA_OK == 0) and never throws;
 *     `hga_last_error()` returns a thread-local message for the last non-zero status.
 *   - matrices are float64, row-major, time along rows: x[t * ld + c]  (the reference's (T, C) ndarray layout,
 *     local/units.py:81,145).
 *   - pointers named `*_dev` are device pointers on the handle's device; `*_host` are host pointers; `stream` is a
 *     `cudaStream_t` passed as `void*` (NULL = the legacy default stream).  Device entry points only enqueue work;
 *     host entry points return after the result is in the caller's buffer.
 *   - a handle is used by one host thread at a time; distinct handles are independent.
 */
#ifndef HGA_B200_H
#define HGA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HGA_ABI_VERSION 1
#define HGA_MAX_SECTIONS 8 /* biquads per filter (the reference designs order-8 Butterworth band filters = 8 SOS) */

enum hga_status {
  HGA_OK = 0,
  HGA_ERR_INVALID = 1,     /* bad argument */
  HGA_ERR_CUDA = 2,        /* CUDA runtime error (message has the cudaError string) */
  HGA_ERR_NO_DEVICE = 3,   /* no CUDA device: there is NO CPU fallback */
  HGA_ERR_UNSUPPORTED = 4, /* configuration outside what the kernels implement */
  HGA_ERR_NOMEM = 5
};

enum hga_precision {
  /* Unfused IEEE double arithmetic in the reference's operation order, sequential in time per channel, windows summed
   * in row order: bit-identical to scipy.signal.sosfilt (x2) -> compute_log_power_features up to the final log()
   * (CUDA log is <= 1 ulp; glibc's is correctly rounded in almost all cases).  Used for streaming and short trials. */
  HGA_PREC_EXACT = 0,
  /* float64 with fused multiply-adds.  A recording is cut into time-chunks that are filtered in parallel from zero state;
   * the filter state is then carried from chunk to chunk exactly (end state -> zero-input response added to the head of
   * the next chunk, down to `carry_tol` of the state), window power as the sum of hop-block partial sums.
   * |feature - reference| <= 1e-9 whatever the out-of-band content of the input (tests state it; typical 1e-12). */
  HGA_PREC_F64 = 1,
  /* float32 cascade arithmetic (two channels per register, fma.rn.f32x2) on the float64 input, float32 power sums / log /
   * z-score, same chunking and carry.  Offline extraction only (streaming pushes keep HGA_PREC_EXACT arithmetic).
   * |feature - reference| <= 1e-4 at 1 kHz, <= 5e-4 at 2 kHz on white / ECoG-like input (tests state it; typical 2e-5 /
   * 1e-4).  A constant per (segment, channel) - the median of the segment's first three samples - is removed from x in float64
   * before the conversion and folded into the start state, so an amplifier offset 1e4 .. 1e5 times the in-band amplitude
   * costs the features nothing (same bound; 1e-3 inside the first 8 s of a segment, while the reference's own start
   * transient decays from the offset's size).  Needs the standard cascade (8 + 8 unit-numerator sections,
   * zi = sosfilt_zi), L = 5 hops, hop 10 or 20, x 16-byte aligned with an even leading dimension. */
  HGA_PREC_F32 = 2,
  /* float32 band-pass (poles <= 0.963, FMA pipe) + float64 notch (poles at 0.9976 - 0.9988, FP64 pipe), float64 power sums /
   * log / z-score.  |feature - reference| <= 5e-5 at 1 kHz, <= 2e-4 at 2 kHz (typical 1e-5 / 4e-5).  Same requirements
   * as HGA_PREC_F32. */
  HGA_PREC_MIXED = 3
};

/* Configuration of one extractor = the constructor arguments of HighGammaExtractor (local/units.py:102-132) after the
 * host has designed the filters (local/units.py:124-126) and taken sosfilt_zi (local/units.py:128-129).            */
typedef struct hga_config {
  int32_t abi_version;                   /* HGA_ABI_VERSION */
  int32_t device;                        /* CUDA device ordinal */
  int32_t fs;                            /* sampling rate, Hz */
  int32_t n_channels;                    /* nb_electrodes (C) */
  int32_t frame_len;                     /* L   = int(window_length * fs)   (hga_optimized.pyx:77) */
  int32_t hop;                           /* hop = int(window_shift  * fs)   (hga_optimized.pyx:76) */
  double eps;                            /* 0.01 (hga_optimized.pyx:46) */
  int32_t n_sections_bp;                 /* sections of hg_filter (band-pass 70-170 Hz), <= HGA_MAX_SECTIONS */
  int32_t n_sections_notch;              /* sections of fh_filter (band-stop 118-122 Hz); 0 = no second filter */
  double sos_bp[HGA_MAX_SECTIONS * 6];   /* scipy SOS rows [b0 b1 b2 a0 a1 a2], a0 == 1 */
  double sos_notch[HGA_MAX_SECTIONS * 6];
  double zi_bp[HGA_MAX_SECTIONS * 2];    /* sosfilt_zi(sos): initial (z0, z1) per section, tiled over channels,
                                            NOT scaled by x[0] (local/units.py:128-132) */
  double zi_notch[HGA_MAX_SECTIONS * 2];
  int32_t precision;                     /* enum hga_precision: arithmetic of hga_extract_offline* */
  int32_t warmup;                        /* generic float64 kernel only (non-standard cascades, odd leading dimensions): history
                                            re-filtered per time-chunk; 0 = derive from warmup_tol and the largest pole radius */
  double warmup_tol;                     /* carry_tol: relative size at which a section's zero-input response is dropped from the
                                            chunk-to-chunk carry (0 = default: 1e-11 F64, 1e-6 F32, 1e-7 MIXED) */
  int32_t max_push_rows;                 /* largest n accepted by hga_stream_push (0 = 4096) */
  int32_t reserved;
} hga_config;

typedef struct hga_handle hga_handle;

/* ---- life cycle ------------------------------------------------------------------------------------------------ */
const char* hga_last_error(void);
int hga_abi_version(void);
/* Number of CUDA devices visible (0 on a CPU-only box; never fails). */
int hga_device_count(void);

/* HighGammaExtractor.__init__ (local/units.py:102-132) + WarmStartFrameBuffer.__init__ (hga_optimized.pyx:65-88). */
int hga_create(const hga_config* cfg, hga_handle** out);
int hga_destroy(hga_handle* h);
/* A fresh extractor: zi re-tiled, frame buffer zeroed and re-armed (WarmStartFrameBuffer.reset, pyx:90-94; the
 * reference gets the same effect by constructing a new extractor per trial, prepare_corpus.py:48). */
int hga_reset(hga_handle* h, void* stream);
/* Rows at the head of a time-chunk that the carried state still reaches (the generic kernel: rows of re-filtered history). */
int hga_warmup(const hga_handle* h);
/* How the last hga_extract_offline* call on this handle was cut (for tests and reports): rows per time-chunk, number of
 * chunks, number of chunks that started from a carried state, rows the carry reaches.  Any pointer may be NULL. */
int hga_last_plan(const hga_handle* h, int64_t* chunk_rows, int32_t* n_chunks, int32_t* n_carried, int32_t* carry_rows);

/* ---- offline: batched independent segments (prepare_corpus.py:42-52; baseline_offline.py:45-50) ------------------ */
/* Frames a single call on `rows` filtered rows yields: compute_log_power_features' count (hga_optimized.pyx:36),
 * evaluated with the reference's C-float arithmetic. */
int64_t hga_num_frames(const hga_handle* h, int64_t rows);

/* For each segment i (rows [seg_start[i], seg_start[i]+seg_len[i]) of x): fresh extractor -> extract_features ->
 * optional z-score (prepare_corpus.py:226-227).  seg_len[i] must be >= frame_len (frame-buffer case 1,
 * pyx:104-107).  Features of segment i are written to out rows [out_row[i], out_row[i]+hga_num_frames(seg_len[i])).
 * seg_start/seg_len/out_row are HOST arrays.  mu_dev/sigma_dev: (C,) or NULL.  Does not touch the streaming state. */
int hga_extract_offline(hga_handle* h, const double* x_dev, int64_t ldx, int32_t n_segments, const int64_t* seg_start,
                        const int64_t* seg_len, const int64_t* out_row, const double* mu_dev, const double* sigma_dev,
                        double* out_dev, int64_t ldo, void* stream);
/* Same with HOST buffers (x, mu, sigma, out).  The recording goes up in time slabs on a copy stream while the kernels run
 * on the chunks that have arrived and finished frames go down.  Pageable host memory is staged through two pinned buffers
 * owned by the handle (memcpy of slab k + 1 overlaps the DMA of slab k); page-locked memory (cudaHostAlloc /
 * cudaHostRegister, detected with cudaPointerGetAttributes) is read by the DMA engine where it lies.  Returns when
 * `out_host` is complete. */
int hga_extract_offline_host(hga_handle* h, const double* x_host, int64_t ldx, int32_t n_segments,
                             const int64_t* seg_start, const int64_t* seg_len, const int64_t* out_row,
                             const double* mu_host, const double* sigma_host, double* out_host, int64_t ldo);

/* ---- streaming: HighGammaExtractor.extract_features on a long-lived extractor (local/units.py:145-161,203-207) --- */
/* Frames the next push of n rows will yield given the handle's frame-buffer state (cases 1-3, pyx:96-131). */
int64_t hga_stream_frames(const hga_handle* h, int64_t n);
/* Filter n new rows (state carried), run the warm-start frame buffer, emit log-power frames, optional z-score.
 * x/out/mu/sigma are device pointers.  *n_frames receives the number of rows written to out. */
int hga_stream_push(hga_handle* h, const double* x_dev, int64_t n, int64_t ldx, const double* mu_dev,
                    const double* sigma_dev, double* out_dev, int64_t ldo, int64_t* n_frames, void* stream);
/* Same with HOST buffers; mu/sigma (host, (C,) or NULL) are cached on the device and re-uploaded only when the
 * pointer's contents change.  One kernel launch per call; returns when out_host holds the frames. */
int hga_stream_push_host(hga_handle* h, const double* x_host, int64_t n, int64_t ldx, const double* mu_host,
                         const double* sigma_host, double* out_host, int64_t ldo, int64_t* n_frames);

/* ---- wire packets at the streaming seam ------------------------------------------------------- */
/* A packet is what ZMQConnector.process receives (local/units.py:78-93) and development_amplifier.py:14-25 packs:
 * struct '=BBB HH' (descriptor, supplement, dtype, n_channels, n_samples; 7 bytes) + n_channels * n_samples float32,
 * channel-major.  hga_packet_header parses and validates it on the host (the payload length must match the header). */
int hga_packet_header(const void* packet_host, int64_t nbytes, int32_t* n_channels, int32_t* n_samples);
/* ZMQConnector.interpret_bytes (local/units.py:78-82) with the result on the device: float32 [C][n] -> float64 (n, C)
 * row-major at x_dev (leading dimension ldx >= C).  Enqueued on `stream`; the packet may be reused on return. */
int hga_packet_decode(int32_t device, const void* packet_host, int64_t nbytes, double* x_dev, int64_t ldx, void* stream);
/* interpret_bytes + HighGammaExtractor.extract_features (local/units.py:78-82,145-161) for one packet on a long-lived
 * extractor: the kernel reads the float32 channel-major payload directly.  out_host receives the float64 frames like
 * hga_stream_push_host; out_f32_dev (device, (n_frames, C) float32, or NULL) additionally receives them as float32 for
 * a consumer on the same GPU (the nVAD's `.float().to(device)`, local/units.py:433).  Packets of 1..1024 samples. */
int hga_stream_push_packet(hga_handle* h, const void* packet_host, int64_t nbytes, const double* mu_host,
                           const double* sigma_host, double* out_host, int64_t ldo, float* out_f32_dev, int64_t* n_frames);

/* Streaming state, in the reference's own shapes: hg_state/fh_state (n_sections, 2, C) as local/units.py:131-132,
 * remainder (frame_len - hop, C) and the first_frame flag as hga_optimized.pyx:78-88.  HOST buffers; NULL = skip. */
int hga_get_state(hga_handle* h, double* hg_state, double* fh_state, double* remainder, int32_t* first_frame);
int hga_set_state(hga_handle* h, const double* hg_state, const double* fh_state, const double* remainder,
                  const int32_t* first_frame);

/* ---- building blocks (the extension module's own functions) ---------------------------------------------------- */
/* scipy.signal.sosfilt(sos, x, axis=0, zi=state) for one filter of n_sections <= 16 (local/units.py:151-152):
 * y (T, C) and state (n_sections, 2, C) updated in place.  Unfused double arithmetic (bit-identical to scipy). */
int hga_sosfilt(int32_t device, const double* sos /*host (n_sections,6)*/, int32_t n_sections, const double* x_dev,
                int64_t T, int32_t C, int64_t ldx, double* state_dev, double* y_dev, int64_t ldy, void* stream);
/* compute_log_power_features (hga_optimized.pyx:27-47) on a device matrix of `rows` rows preceded by `zero_rows`
 * virtual rows of zeros: out[w, c] = log(sum_{r in [start_w, stop_w)} data[r, c]^2 / (stop_w - start_w) + eps),
 * sequential sums.  Regular windows: start_w = w * hop, stop_w = start_w + frame_len (start_dev/stop_dev NULL);
 * irregular windows: explicit device index arrays (n_frames,).  Optional z-score with mu_dev/sigma_dev. */
int hga_log_power(int32_t device, const double* data_dev, int64_t rows, int64_t zero_rows, int32_t C, int64_t ld,
                  int64_t n_frames, int32_t frame_len, int32_t hop, const int64_t* start_dev, const int64_t* stop_dev,
                  double eps, const double* mu_dev, const double* sigma_dev, double* out_dev, int64_t ldo,
                  void* stream);
/* ZScoreNormalization.__call__ (local/common.py:375-376): out = (f - mu) / sigma, in place allowed. */
int hga_zscore(int32_t device, const double* f_dev, int64_t F, int32_t C, int64_t ld, const double* mu_dev,
               const double* sigma_dev, double* out_dev, int64_t ldo, void* stream);

/* ---- baseline statistics (prepare_corpus.py:190-199; baseline_offline.py:45-60) --------------------------------- */
/* acc_dev is (3, C) float64: rows n, mean, M2 (sum of squared deviations).  Zero-initialise to start.
 * accumulate: Chan-merges the F frames of feats into acc.  merge: acc_a <- merge(acc_a, acc_b) for `count`
 * consecutive accumulators in acc_b (fixed order: deterministic).  finalize: mu = mean, sigma = sqrt(M2 / n)
 * (population std, ddof = 0, as np.std). */
int hga_stats_accumulate(int32_t device, const double* feats_dev, int64_t F, int32_t C, int64_t ld, double* acc_dev,
                         void* stream);
int hga_stats_merge(int32_t device, double* acc_a_dev, const double* acc_b_dev, int32_t count, int32_t C,
                    void* stream);
/* hga_stats_accumulate that allocates nothing.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same result bit for bit; the block partials go to a 1 MiB static device array instead of the library's
 * grow-only scratch, so the call issues no cudaMalloc / cudaFree.  Limit: row tiles x 3 x C doubles <= 131072 (on a
 * 148-SM device: any C up to 43690 channels); larger shapes return HGA_ERR_INVALID with the sizes in hga_last_error().
 * Every call on a device shares the array (it keeps the last call's partials; nothing needs clearing, every word read
 * is written first): calls that overlap on the GPU - different streams without ordering - race on it. */
int hga_stats_accumulate_static(int32_t device, const double* feats_dev, int64_t F, int32_t C, int64_t ld,
                                double* acc_dev, void* stream);
int hga_stats_finalize(int32_t device, const double* acc_dev, int32_t C, double* mu_dev, double* sigma_dev,
                       void* stream);
/* The day-baseline exchange of a sharded corpus in one launch (prepare_corpus.py:183-199 over ranks): gathered_dev holds
 * (n_src, n_days, 3, C) accumulators, rank-major as an all-gather leaves them; per (day, channel) they are merged in source
 * order (identical on every rank) and finalised into mu_dev / sigma_dev, each (n_days, C).  Days without frames: 0 / 1. */
int hga_stats_merge_days(int32_t device, const double* gathered_dev, int32_t n_src, int32_t n_days, int32_t C, double* mu_dev,
                         double* sigma_dev, void* stream);

/* ---- channel-axis transforms either side of the path -------------------------------------- */
/* One row-wise map  out[t, j] = alpha[j] * x[t, sel[j]] + beta[j] * mean_{c in group grp[j]} x[t, c]  covers
 *   - SelectElectrodesFromBothGrids / SelectElectrodesOverSpeechAreas.__call__ (local/common.py:31-32,54-55): alpha 1, beta 0;
 *   - CommonAverageReferencing.__call__ (local/common.py:336-343): alpha 1, beta -1, group = the grid's computation mask;
 *   - BadChannelCorrection.__call__ (local/common.py:279-284): alpha 0, beta 1, group = the bad channel's neighbours;
 *   - and gather -> CAR -> gather chains (prepare_corpus.py:152-160, decode_online.py:69-82) folded by the host into
 *     raw-column indices.  sel[j] = -1 / grp[j] = -1 drop the respective term.  Groups are CSR: columns
 *     gidx[gptr[g] .. gptr[g+1]) (raw input columns), at most 64 groups.  All table arguments are HOST arrays, copied
 *     at creation.  In place (out == x, ldo == ldx) is allowed. */
typedef struct hga_chanmap hga_chanmap;
int hga_chanmap_create(int32_t device, int32_t c_in, int32_t c_out, const int32_t* sel, const int32_t* grp,
                       const double* alpha, const double* beta, int32_t n_groups, const int32_t* gptr,
                       const int32_t* gidx, hga_chanmap** out);
int hga_chanmap_destroy(hga_chanmap* m);
int hga_chanmap_apply(hga_chanmap* m, const double* x_dev, int64_t T, int64_t ldx, double* out_dev, int64_t ldo,
                      void* stream);
/* Raw ingest: the same map applied to amplifier counts of another element type, each column scaled at the load,
 * double(x[t, c]) * gain[c] -- `BCI2000MatFile.signals()` = `signal * gain` (/root/reference/local/common.py:456-459)
 * without materialising the float64 recording on the host: the upload is the raw element size.  x_dev: [T, ldx]
 * elements of `dtype`; gain_dev: [c_in] float64 in device memory, or NULL for no scaling. */
enum { HGA_RAW_F64 = 0, HGA_RAW_F32 = 1, HGA_RAW_I32 = 2, HGA_RAW_I16 = 3 };
int hga_chanmap_apply_raw(hga_chanmap* m, const void* x_dev, int32_t dtype, const double* gain_dev, int64_t T, int64_t ldx,
                          double* out_dev, int64_t ldo, void* stream);
int hga_chanmap_apply_host(hga_chanmap* m, const double* x_host, int64_t T, int64_t ldx, double* out_host, int64_t ldo);
/* hga_extract_offline_host fed with raw amplifier counts: x_host is [rows, ldx] elements of `dtype` (HGA_RAW_*), map->c_in
 * columns wide; every upload slab travels at the raw element size and is expanded on the device, double(x) * gain[c]
 * followed by the map (identity, or the folded selection / CAR chain), into the float64 rows the extraction reads.  This is
 * `FeatureExtractionPipeline.get_features(BCI2000MatFile.signals(), trial_indices)` (/root/reference/prepare_corpus.py:42-52,
 * local/common.py:456-459) without the float64 recording ever existing on the host.  gain_host: [c_in] float64 or NULL.
 * Everything else as hga_extract_offline_host. */
int hga_extract_offline_raw_host(hga_handle* h, hga_chanmap* map, const void* x_host, int32_t dtype, const double* gain_host,
                                 int64_t ldx, int32_t n_segments, const int64_t* seg_start, const int64_t* seg_len,
                                 const int64_t* out_row, const double* mu_host, const double* sigma_host, double* out_host,
                                 int64_t ldo);

/* hga_extract_offline that also accumulates the baseline statistics of the frames it writes: the day-baseline flow of
 * /root/reference/prepare_corpus.py:190-199 and baseline_offline.py:45-60 (features of the baseline recording's trials,
 * then np.mean / np.std over all frames) in one pass.  acc_dev is a (3, C) accumulator as for hga_stats_accumulate
 * (zero-initialise to start; finish with hga_stats_finalize).  With HGA_PREC_F32 handles on layouts the lane kernels take,
 * the kernels leave (n, mean, M2) partials of the frames they emit and one small kernel merges them in a fixed order
 * (*fused = 1: the features are not read back; the accumulation runs on the otherwise idle float64 pipe); otherwise
 * extraction and hga_stats_accumulate run one after the other behind the same call (*fused = 0: in the float64 kernels
 * the accumulation would cost more than the second read).  fused may be NULL. */
int hga_extract_offline_stats(hga_handle* h, const double* x_dev, int64_t ldx, int32_t n_segments, const int64_t* seg_start,
                              const int64_t* seg_len, const int64_t* out_row, const double* mu_dev, const double* sigma_dev,
                              double* out_dev, int64_t ldo, double* acc_dev, void* stream, int32_t* fused);

/* hga_extract_offline with the pre-transform chain inside the extraction kernel:
 * `HighGammaExtractor.extract_features` applies `pre_transform` and hands the result to the filters
 * (/root/reference/local/units.py:147-152; chains as wired in prepare_corpus.py:152-160).  x_raw_dev: float64 device rows
 * [rows, ldx], map->c_in columns wide, BEFORE the channel map; segments are given in those rows; everything else as
 * hga_extract_offline.  When the handle computes in packed float32 (HGA_PREC_F32) at hop 10 and the map fits the kernel
 * (raw rows of up to ~180 columns, at most 2 averaging groups with 512 members together) the mapped recording never exists
 * in memory: mapper warps form the kernel's input tiles from the raw rows (*fused = 1).  Otherwise the map runs as a pass
 * of its own into scratch and the plain extraction follows (*fused = 0).  fused may be NULL. */
int hga_extract_offline_mapped(hga_handle* h, hga_chanmap* map, const double* x_raw_dev, int64_t ldx, int32_t n_segments,
                               const int64_t* seg_start, const int64_t* seg_len, const int64_t* out_row,
                               const double* mu_dev, const double* sigma_dev, double* out_dev, int64_t ldo, void* stream,
                               int32_t* fused);

/* ---- introspection for bench.py / tests ------------------------------------------------------------------------- */
/* Kernel launches issued through this library by the calling process since load (all handles). */
int64_t hga_launch_count(void);
/* Device time (ms, CUDA events on the launching stream) of the kernel sequence of the last hga_extract_offline call
 * on this handle when timing was enabled with hga_set_timing(h, 1); -1 if unavailable. */
int hga_set_timing(hga_handle* h, int32_t enable);
double hga_last_kernel_ms(hga_handle* h);
/* The same time split at the end of the dominant kernel (k_hga_lanes / k_hga_lane): main_ms = that kernel alone,
 * rest_ms = state carry + head fix-up + seam frames behind it.  HGA_ERR_INVALID when the last timed call did not
 * take the lane kernels. */
int hga_last_kernel_split_ms(hga_handle* h, double* main_ms, double* rest_ms);

/* Test instrumentation (compute-sanitizer is not available on every pool): with guards enabled, every device buffer the
 * library allocates from then on carries 64 KB of a fill pattern in front of and behind it.  hga_debug_check_guards
 * synchronises the device and returns the number of guard bands that no longer hold the pattern (0 = no kernel wrote
 * outside the library's own scratch), hga_debug_guarded_buffers how many buffers are being watched. */
int hga_debug_guards(int32_t enable);
int64_t hga_debug_check_guards(void);
int64_t hga_debug_guarded_buffers(void);

#ifdef __cplusplus
}
#endif
#endif /* HGA_B200_H */
