/*
 * libspeechpy_b200.so -- C ABI of the B200-native speechpy front end and
 * cosine / EER scoring path.
 *
 * The reference (MingmChen/Speaker_Verification) has no FFI layer: the hot
 * path sits behind the speechpy *Python module API* (numpy in / numpy out).
 * Each entry point below is what a ctypes binding for one reference function
 * calls; the file:line it replaces is cited per function ("sfe/" =
 * speech_feature_extraction/).  The Python side of the binding lives in
 * speaker_verification_b200/speechpy/ and INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - every function returns an spb_status (0 == SPB_OK); the message for the
 *     last failure on a context is spb_last_error(ctx) (ctx == NULL: the last
 *     failure of a call that had no context, thread-local);
 *   - the caller allocates every output buffer; the library never frees or
 *     keeps caller memory;
 *   - plain pointers and sizes only, no torch / numpy types;
 *   - functions without a "_dev" suffix take HOST pointers and perform the
 *     host<->device copies themselves (pinned staging inside the context);
 *     "_dev" functions take DEVICE pointers valid on the context's device and
 *     are asynchronous on the context's stream (spb_sync to wait);
 *   - a context is bound to one CUDA device and owns one stream; it may be
 *     used by one thread at a time, different contexts are independent;
 *   - there is NO CPU fallback: without a usable CUDA device spb_init fails.
 *
 * Arithmetic: float32 on device (FFT, mel, log, DCT), float64 accumulators
 * for cmvn / cmvnw, float64 end to end for stack_frames / cmvn / delta cube.
 */
#ifndef SPEECHPY_B200_H_
#define SPEECHPY_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPB_VERSION 100 /* 0.1.0 */

#if defined(__GNUC__)
#define SPB_API __attribute__((visibility("default")))
#else
#define SPB_API
#endif

typedef enum spb_status {
  SPB_OK = 0,
  SPB_ERR_INVALID = 1,      /* bad argument (maps to AssertionError / ValueError in Python) */
  SPB_ERR_NEGATIVE_DIM = 2, /* N < frame length: numpy's "negative dimensions are not allowed" */
  SPB_ERR_CUDA = 3,         /* a CUDA runtime call failed */
  SPB_ERR_UNSUPPORTED = 4,  /* legal in the reference, not implemented here (documented) */
  SPB_ERR_NOMEM = 5
} spb_status;

typedef enum spb_dtype { SPB_F32 = 0, SPB_F64 = 1, SPB_I16 = 2, SPB_I32 = 3 } spb_dtype;

typedef struct spb_context spb_context;
typedef struct spb_plan spb_plan;
typedef struct spb_batch spb_batch;

/* ------------------------------------------------------------------ context */
SPB_API int spb_version(void);
SPB_API int spb_init(int device, spb_context **ctx);
SPB_API int spb_destroy(spb_context *ctx);
SPB_API const char *spb_last_error(spb_context *ctx);
SPB_API int spb_sync(spb_context *ctx);
/* cudaStream_t of the context, as an integer (for torch.cuda.ExternalStream). */
SPB_API int spb_stream(spb_context *ctx, uint64_t *stream);
/* sm_count, compute capability major / minor, total global memory in bytes. */
SPB_API int spb_device_info(spb_context *ctx, int32_t *sm_count, int32_t *cc_major,
                    int32_t *cc_minor, uint64_t *total_mem);
/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
SPB_API int spb_launch_count(spb_context *ctx, uint64_t *count);
/* Per-kernel device timing: when enabled, every launch of the hot kernels is
 * bracketed by CUDA events on the context's stream.  spb_timing_read synchronises,
 * sums the elapsed times recorded since the last read and clears them.
 * kernel: 0 = features (framing+FFT+mel+log), 1 = cmvnw stage 1, 2 = cmvnw stage 2,
 * 3 = cosine GEMM, 4 = cmvnw single streaming pass (option cmvnw_kernel 3), 5 = cmvnw with the utterance resident in
 * shared memory (both stages in one pass: the default for utterances of <= 930 rows). */
SPB_API int spb_timing_enable(spb_context *ctx, int32_t on);
SPB_API int spb_timing_read(spb_context *ctx, int32_t kernel, double *total_ms, uint64_t *launches);
/* Process-wide kernel selection and guard knobs.  The defaults are the product path; the other
 * values exist so that tests can cross-check one kernel family against another on the same call
 * (parity of every variant) and so that the guard can be calibrated.
 *   "feature_kernel"  0 auto | 1 generic radix-2 FFT kernel | 2 CUDA-core FFT kernels only
 *   "cmvnw_kernel"    0 auto (utterances of <= 930 rows resident in shared memory, both stages in one pass; longer ones
 *                     large-tile two-pass) | 1 small-tile two-pass | 2 large-tile two-pass |
 *                     3 single streaming pass (window 301, 40 columns, variance normalisation) |
 *                     4 as 0, and utterances of 931 .. 1 860 rows resident in the shared memory of a cluster of two CTAs
 *                     (halves exchanged over DSMEM; one read + one write per row, but measured slower than two passes)
 *   "gemm_multicast"  1 cosine GEMM on 2-CTA clusters with the T tiles multicast (default) | 0 one CTA per tile
 *   "tc_version"      2 (default) | 1: generation of the tensor-core lmfe kernel
 *   "tc_epilogue"     -1 auto | 0 generic | 1 runs of four filters (when the bank allows) | 2 static
 *   "redo"            1 (default) | 0: float64 recomputation of low-energy filters (the guard that
 *                     keeps every log-mel value within 1e-4 of feature.py:216-217,256)
 *   "tau_tc", "tau_fft"  guard thresholds (fractions of the frame's level)
 * Unknown names / out-of-range values return SPB_ERR_INVALID. */
SPB_API int spb_set_option(const char *name, double value);
SPB_API int spb_get_option(const char *name, double *value);
/* Pinned host memory, so callers can hand page-locked buffers to the host API. */
SPB_API int spb_host_alloc(void **ptr, size_t bytes);
SPB_API int spb_host_free(void *ptr);

/* ------------------------------------------------------- a1  stack_frames
 * sfe/speechpy/processing.py:61-139.  Host integer math only:
 * F = ceil((N-L)/S) if zero_padding else floor((N-L)/S); SPB_ERR_NEGATIVE_DIM
 * when that is negative (the reference raises ValueError). */
SPB_API int spb_num_frames(int64_t num_samples, int32_t frame_len, int32_t frame_stride,
                   int32_t zero_padding, int64_t *num_frames);
/* out[f*L + j] = sig[f*S + j] * window[j] in float64 (window == NULL: ones).
 * num_frames must equal spb_num_frames(...). Samples at or beyond num_samples
 * read as 0 (processing.py:108-110). */
SPB_API int spb_stack_frames(spb_context *ctx, const double *sig, int64_t num_samples,
                     int32_t frame_len, int32_t frame_stride, const double *window,
                     int64_t num_frames, double *out);

/* ---------------------------------- a2  fft_spectrum / power_spectrum
 * sfe/speechpy/processing.py:142-159, 162-174.  frames is (F, L) row-major
 * float64; rows are zero-padded (L < nfft) or truncated (L > nfft) to nfft as
 * numpy's rfft(n=) does.  out is (F, nfft/2+1) float64.
 * mode 0: |X|^2/nfft   mode 1: |X|
 * mode 2: log_power_spectrum(normalize=False) = 10 log10(max(P, 1e-20)) (processing.py:177-198)
 * mode 3: the same minus its global maximum (normalize=True).
 * Any nfft >= 1 (numpy's rfft(n=) takes them all): powers of two in [32, 4096] run the FFT kernels --
 * deviation from the reference: in float32 on the device (frames are narrowed after upload, the result
 * widened to float64), |error| <= 1e-5 of the row maximum; every other size runs a direct DFT in float64. */
SPB_API int spb_power_spectrum(spb_context *ctx, const double *frames, int64_t num_frames,
                       int32_t frame_len, int32_t nfft, int32_t mode, double *out);

/* ------------------------------------------------- 8f-2  preemphasis
 * sfe/speechpy/processing.py:45-58: out[i] = sig[i] - cof * sig[(i - shift) mod n]
 * (np.roll is circular: sample 0 uses the last sample).  Computed in the dtype numpy
 * uses: float32 in -> float32 out; float64 / int16 / int32 in -> float64 out. */
SPB_API int spb_preemphasis(spb_context *ctx, const void *sig, int32_t dtype, int64_t n, int64_t shift,
                    double cof, void *out);

/* -------------------------------------------------------------- feature plan
 * Everything mfe / lmfe / mfcc need that does not depend on the signal.  The
 * mel filterbank is built by the caller in float64 exactly as
 * sfe/speechpy/feature.py:55-97 does (integer bin edges must be bit-exact, so
 * they are never recomputed on device) and passed dense, (num_filters, nfft/2+1)
 * row-major.  Plans are immutable and may be shared by calls on their context. */
typedef struct spb_feature_params {
  int32_t frame_len;     /* L = int(round(fs*frame_length))          processing.py:93-97 */
  int32_t frame_stride;  /* S = round(fs*frame_stride)               processing.py:98   */
  int32_t nfft;          /* fft_length                               feature.py:199     */
  int32_t num_filters;   /* M                                        feature.py:208     */
  int32_t num_cepstral;  /* C (mfcc only; <= M)                      feature.py:147     */
  int32_t dc_elimination;/* c0 <- ln(frame energy)                   feature.py:151-152 */
  int32_t cmvn_window;   /* odd win_size for cmvnw (fused batch)     processing.py:296-299 */
  int32_t cmvn_variance; /* variance_normalization                   processing.py:310  */
} spb_feature_params;

SPB_API int spb_plan_create(spb_context *ctx, const spb_feature_params *params,
                    const double *filterbank, spb_plan **plan);
SPB_API int spb_plan_destroy(spb_plan *plan);
/* 8f-2: pre-emphasis (sfe/speechpy/processing.py:45-58, shift = 1) as a prologue of the framing kernels: every
 * feature call on this plan (single signal or batch) then works on y[i] = x[i] - cof * x[i-1], y[0] = x[0] -
 * cof * x[N-1] (np.roll is circular) of each utterance, computed in float32 while a tile's samples are staged
 * in shared memory (CUDA-core FFT kernels) or applied in the spectral domain in the epilogue of the tensor-core
 * kernel (log-mel rows of the default shape) -- the emphasised signal is never written.  cof = 0 (the default)
 * switches it off.  Set it before the plan is used by concurrent calls. */
SPB_API int spb_plan_set_preemphasis(spb_plan *plan, double cof);

/* --------------------------------------- a4-a6  mfe / lmfe / mfcc (one signal)
 * sfe/speechpy/feature.py:156-219 (mfe), 222-258 (lmfe), 102-153 (mfcc).
 * sig: num_samples values of `dtype` (converted like signal.astype(float));
 * frames = floor((N-L)/S) (zero_padding=False, feature.py:193).  Outputs are
 * float64: feat (F, M), energy (F) [may be NULL], ceps (F, C).  Exact zeros
 * become float64 eps before the log (functions.py:55-62).
 * Deviation from the reference: framing, FFT and filterbank run in float32 on the device (the
 * signal is narrowed after upload, results widened to float64); log-mel / MFCC values are within
 * 1e-4 of the float64 reference on every value (filters whose energy is too low for float32 -- for
 * mfcc: the rows that hold one -- are recomputed in float64, csrc/redo_kernels.cu). */
SPB_API int spb_mfe(spb_context *ctx, const spb_plan *plan, const void *sig, int32_t dtype,
            int64_t num_samples, double *feat, double *energy);
SPB_API int spb_lmfe(spb_context *ctx, const spb_plan *plan, const void *sig, int32_t dtype,
             int64_t num_samples, double *feat);
SPB_API int spb_mfcc(spb_context *ctx, const spb_plan *plan, const void *sig, int32_t dtype,
             int64_t num_samples, double *ceps);

/* ------------------------------------ a7  extract_derivative_feature
 * sfe/speechpy/feature.py:261-282 + processing.py:201-236 (as written: forward
 * weighted sum along the feature axis, DeltaWindows = 2).  feat (F, Mc) float64
 * -> out (F, Mc, 3) float64. */
SPB_API int spb_delta_cube(spb_context *ctx, const double *feat, int64_t num_frames,
                   int32_t num_cols, double *out);
/* processing.derivative_extraction itself, any DeltaWindows >= 1: (F, Mc) -> (F, Mc). */
SPB_API int spb_derivative(spb_context *ctx, const double *feat, int64_t num_frames,
                   int32_t num_cols, int32_t delta_windows, double *out);

/* ------------------------------------------------------- a8 / a9  cmvn, cmvnw
 * sfe/speechpy/processing.py:239-271: float64 in, float64 out.
 * sfe/speechpy/processing.py:274-327: float64 in, float32 out; win_size odd.  The window means are
 * float64 sums of the float64 input and x - mean is rounded to float32 once, as in the reference
 * (processing.py:302-307), so a large common offset costs no precision; stage 2 (window std of the
 * float32 rows) runs in float32 with float64 block sums: var = E[x^2] - mean^2 anchored every 30 rows
 * in float64 and slid with float32 deltas -- within 1e-4 for feature-like data (|x - mean| spanning
 * <~ 3 decades inside one window); numpy's two-pass float32 std is more accurate beyond that. */
SPB_API int spb_cmvn(spb_context *ctx, const double *vec, int64_t rows, int32_t cols,
             int32_t variance_normalization, double *out);
SPB_API int spb_cmvnw(spb_context *ctx, const double *vec, int64_t rows, int32_t cols,
              int32_t win_size, int32_t variance_normalization, float *out);

/* ------------------------------------------ ragged batches (the hot path)
 * A batch is U utterances inside one float32 sample arena: utterance u is
 * samples[starts[u] .. starts[u]+lengths[u]).  starts that are multiples of 4
 * (16 bytes) take the bulk-copy path.  The descriptor object precomputes the
 * frame offsets and the tile table on the device for (plan, starts, lengths).
 * Output row r of utterance u is row frame_offsets[u] + r of the (total_frames,
 * width) float32 output. */
SPB_API int spb_batch_create(spb_context *ctx, const spb_plan *plan, const int64_t *starts,
                     const int64_t *lengths, int64_t num_utterances, spb_batch **batch);
SPB_API int spb_batch_destroy(spb_batch *batch);
SPB_API int spb_batch_total_frames(const spb_batch *batch, int64_t *total_frames);
/* frame_offsets has num_utterances+1 entries. */
SPB_API int spb_batch_frame_offsets(const spb_batch *batch, int64_t *frame_offsets);

/* lmfe (feature.py:222-258) followed by cmvnw(win, variance) (processing.py:274-327),
 * device pointers: d_samples float32 arena, d_out (total_frames, M) float32. */
SPB_API int spb_lmfe_cmvnw_batch_dev(spb_context *ctx, const spb_batch *batch,
                             const float *d_samples, int64_t arena_len, float *d_out);
/* lmfe only (no normalisation): d_out (total_frames, M) float32. */
SPB_API int spb_lmfe_batch_dev(spb_context *ctx, const spb_batch *batch, const float *d_samples,
                       int64_t arena_len, float *d_out);
/* mfcc (feature.py:102-153) + extract_derivative_feature (feature.py:261-282):
 * d_out (total_frames, C, 3) float32. */
SPB_API int spb_mfcc_delta_batch_dev(spb_context *ctx, const spb_batch *batch,
                             const float *d_samples, int64_t arena_len, float *d_out);
/* Host-pointer variants, pipelined over chunks of whole utterances: the host->device copy
 * of chunk k+1, the kernels of chunk k and the device->host copy of chunk k-1 run on three
 * streams, so a call costs about max(H2D, compute, D2H).  Pinned buffers (spb_host_alloc)
 * give full PCIe rate; pageable buffers work but copy synchronously. */
SPB_API int spb_lmfe_cmvnw_batch(spb_context *ctx, const spb_batch *batch, const float *samples,
                         int64_t arena_len, float *out);
SPB_API int spb_mfcc_delta_batch(spb_context *ctx, const spb_batch *batch, const float *samples,
                         int64_t arena_len, float *out);
/* Same, with the sample arena in any spb_dtype (wav ingest).  The device
 * computes x = float(sample) * sample_scale before framing:
 *   sample_scale = 1            `signal.astype(float)` (feature.py:182) of an integer array;
 *   SPB_I16, 1.0/32768          the float32 array librosa.load (utils.py:170-173 load_wav,
 *                               called at load_data.py:53) decodes from a PCM16 wav.
 * The PCM arena crosses PCIe at 2 bytes per sample; widening and power-of-two scaling are
 * exact in float32.  starts/lengths/arena_len are in samples. */
SPB_API int spb_lmfe_cmvnw_batch_dtype(spb_context *ctx, const spb_batch *batch, const void *samples,
                               int dtype, double sample_scale, int64_t arena_len, float *out);
SPB_API int spb_mfcc_delta_batch_dtype(spb_context *ctx, const spb_batch *batch, const void *samples,
                               int dtype, double sample_scale, int64_t arena_len, float *out);

/* ------------------------------------------------- 8f-1  training feature cubes
 * utils.py:325-397, the DataLoader transforms that follow lmfe (train.py:194-210):
 *   CMVN        : derivative != 0 -> extract_derivative_feature (F, M, 3);
 *                 normalize  != 0 -> each (F, M) channel <- cmvn(channel, True) over the utterance
 *   FeatureCube : out[u][ch][k][t][c] = feature_u[crop_starts[u*num_crops + k] + t][c][ch]
 * out: (U, channels, num_crops, crop_frames, M) float32, channels = derivative ? 3 : 1
 * (utterance u's slice is the reference's FeatureCube3C / FeatureCube[None] output).
 * crop_starts: HOST int32 [U * num_crops], each in [0, F_u - crop_frames] -- the reference
 * draws them with np.random.randint(F - frames, size=crops) (utils.py:373); the caller owns
 * the RNG.  The log-mel features never leave the device. */
SPB_API int spb_feature_cubes_batch_dev(spb_context *ctx, const spb_batch *batch, const float *d_samples,
                                int64_t arena_len, int32_t derivative, int32_t normalize,
                                const int32_t *crop_starts, int32_t num_crops, int32_t crop_frames,
                                float *d_out);
SPB_API int spb_feature_cubes_batch(spb_context *ctx, const spb_batch *batch, const void *samples, int32_t dtype,
                            double sample_scale, int64_t arena_len, int32_t derivative, int32_t normalize,
                            const int32_t *crop_starts, int32_t num_crops, int32_t crop_frames, float *out);

/* ------------------------------------------------------- a10  cosine scores
 * evaluation.py:76-77 over all (enrolment, test) pairs at once:
 * S[i, j] = <E_i, T_j> / (|E_i| |T_j|); zero-norm rows give 0 (sklearn).
 * E (m, d), T (n, d), S (m, n) row-major float32; any d >= 1 (padded with zeros to a multiple of 16).
 * Tensor-core path: tcgen05 kind::tf32, 3-term split (hi*hi + hi*lo + lo*hi),
 * fp32 accumulation in TMEM. */
SPB_API int spb_cosine_scores(spb_context *ctx, const float *E, int64_t m, const float *T,
                      int64_t n, int32_t d, float *S);
SPB_API int spb_cosine_scores_dev(spb_context *ctx, const float *d_E, int64_t m, const float *d_T,
                          int64_t n, int32_t d, float *d_S);
/* Allocation-free variant of spb_cosine_scores_dev.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * spb_cosine_scores_dev, but the normalised TF32 hi/lo split of E and T lives in a static
 * __device__ array of SPB_NOALLOC_SCRATCH_BYTES (8 MiB) inside the library instead of the
 * context's grow-only scratch, and the call performs no cudaMalloc / cudaFree / allocating library
 * routine (the kernel-timing events of spb_timing_enable are not recorded for it).  Limit:
 * 2 * (pad256(m * dp) + pad256(n * dp)) * 4 bytes <= SPB_NOALLOC_SCRATCH_BYTES with dp = d rounded
 * up to a multiple of 16 and pad256 rounding up to 256 floats (e.g. m + n <= 2 048 rows at d = 512);
 * larger problems return SPB_ERR_INVALID with a message naming both sizes.  Every call on a device
 * shares the one array (each call rewrites all of it that it reads, so nothing has to be cleared):
 * calls that overlap on the GPU -- different contexts / streams of the same device -- race; calls
 * on one context are stream-ordered and safe.  The first call on a device resolves the array's
 * device address (cudaGetSymbolAddress); make it outside a graph capture. */
#define SPB_NOALLOC_SCRATCH_BYTES (8u << 20)
SPB_API int spb_cosine_scores_dev_noalloc(spb_context *ctx, const float *d_E, int64_t m, const float *d_T,
                                  int64_t n, int32_t d, float *d_S);
/* out[t] = S[trial_i[t], trial_j[t]] gathered from a device score matrix. */
SPB_API int spb_trial_scores_dev(spb_context *ctx, const float *d_S, int64_t m, int64_t n,
                         const int32_t *d_trial_i, const int32_t *d_trial_j,
                         int64_t num_trials, float *d_out);
/* argmax over each row of S (evaluation.py:79,126-130). */
SPB_API int spb_argmax_rows_dev(spb_context *ctx, const float *d_S, int64_t m, int64_t n,
                        int32_t *d_out);

/* ------------------------------------------------------------ a11  EER / AUC
 * evaluation.py:47-52: roc_curve(label, score, pos_label=1) -> brentq root of
 * 1 - x - tpr(x); roc_auc_score.  Device sort (descending, distinct-threshold
 * groups) + integer cumulative counts + one linear solve.  labels are 0/1. */
SPB_API int spb_eer(spb_context *ctx, const float *scores, const int32_t *labels, int64_t n,
            double *eer, double *auc);
SPB_API int spb_eer_dev(spb_context *ctx, const float *d_scores, const int32_t *d_labels,
                int64_t n, double *eer, double *auc);

/* ROC curve export: sklearn.metrics.roc_curve(label, score, pos_label=1) as
 * called at evaluation.py:48 (drop_intermediate=True; first point (0, 0, +inf)).
 * fpr / tpr / thresholds: caller-allocated HOST arrays of `capacity` doubles (n + 1 always
 * suffices; thresholds may be NULL); *count receives the number of points (also when the
 * capacity is too small, which returns SPB_ERR_INVALID).  _dev: scores/labels on the device. */
SPB_API int spb_roc_curve(spb_context *ctx, const float *scores, const int32_t *labels, int64_t n,
                  double *fpr, double *tpr, double *thresholds, int64_t capacity, int64_t *count);
SPB_API int spb_roc_curve_dev(spb_context *ctx, const float *d_scores, const int32_t *d_labels, int64_t n,
                      double *fpr, double *tpr, double *thresholds, int64_t capacity, int64_t *count);

#ifdef __cplusplus
}
#endif
#endif /* SPEECHPY_B200_H_ */
