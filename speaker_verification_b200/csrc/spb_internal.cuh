// Internal declarations shared by the kernels and the C-ABI layer.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "speechpy_b200.h"

#ifndef SPB_FFT_NW
#define SPB_FFT_NW 8   // warps per CTA of the NFFT-512 feature kernel; a tile is 4 frames per warp
#endif

namespace spb {
constexpr int kFftWarps = SPB_FFT_NW;
constexpr int kFftTileFrames = 4 * kFftWarps;
constexpr int kTcGroupFrames = 31;   // tensor-core feature kernel: 32 hop rows -> 31 frames

// ---------------------------------------------------------------- errors
void set_global_error(const std::string &msg);

struct DeviceBuffer {
  void *ptr = nullptr;
  size_t bytes = 0;
};

struct FftTables {
  // generic kernel: tw[k] = exp(-2*pi*i*k/nfft), k < nfft/2
  float2 *d_tw = nullptr;
  // fast 512 kernel: twA[l*16+k1] = W256^(l*k1); twB[k1*9+k2] = W512^(k1+16*k2)
  float2 *d_twA = nullptr;
  float2 *d_twB = nullptr;
  // float64 redo kernels: tw64[r] = (cos, -sin)(2 pi r / nfft), r < nfft
  double2 *d_tw64 = nullptr;
};

}  // namespace spb

struct spb_context {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;  // host->device leg of the pipelined host batch calls
  cudaStream_t out_stream = nullptr;   // device->host leg
  std::vector<cudaEvent_t> pipe_events;  // per-chunk events of the pipelined host batch calls
  cudaEvent_t ev_a = nullptr, ev_b = nullptr;
  int sm_count = 0, cc_major = 0, cc_minor = 0;
  size_t total_mem = 0;
  uint64_t launches = 0;
  std::string last_error;
  // grow-only device scratch, keyed by role
  spb::DeviceBuffer scratch[12];
  // pinned staging (two halves for double buffering)
  void *pinned[2] = {nullptr, nullptr};
  size_t pinned_bytes = 0;
  std::map<int, spb::FftTables> fft_tables;  // by nfft
  std::vector<spb_plan *> plans;
  // (kernel function, dynamic smem) -> resident CTAs per SM, set up once per context
  std::map<std::pair<const void *, size_t>, int> kernel_cfg;
  std::map<const void *, size_t> kernel_smem_max;
  // optional per-kernel timing (spb_timing_enable)
  bool timing = false;
  std::vector<cudaEvent_t> ev_pool;
  size_t ev_used = 0;
  std::vector<std::pair<size_t, size_t>> ev_pairs[6];  // indices into ev_pool
};

struct spb_plan {
  spb_context *ctx = nullptr;
  spb_feature_params p{};
  int K = 0;  // nfft/2+1
  // band-compressed filterbank (float32 on device)
  int32_t *d_band_lo = nullptr;   // [M] first bin with non-zero weight
  int32_t *d_band_n = nullptr;    // [M] number of bins in the band
  int32_t *d_band_off = nullptr;  // [M] offset into d_band_w
  float *d_band_w = nullptr;      // [sum band_n]
  int band_total = 0;
  int max_bin = -1;               // largest bin index carrying weight
  float *d_dct = nullptr;         // [C][M] orthonormal DCT-II rows
  // segment form of a triangular bank (every bin feeds at most the rising side of filter j and
  // the falling side of filter j-1): per bin {rise weight, fall weight, end-of-segment flag, 0},
  // bins [seg_kmin, seg_kmin + seg_nbins); 8 warp ranges {first bin, end bin, first segment, 0}.
  // seg_nbins == 0: the bank does not have that structure (generic taps are used).
  float4 *d_seg_rd = nullptr;
  int4 *d_seg_ranges = nullptr;
  int seg_kmin = 0, seg_nbins = 0;
  std::vector<int32_t> h_band_lo, h_band_n, h_band_off;
  // tensor-core hop-DFT kernel (feat_tc_kernels.cu): fp16 hi/lo DFT matrix in its shared-memory
  // layout, per-bin {rise, fall, cos, sin}; tc_ok == false: plan shape not covered
  bool tc_ok = false;
  bool tc_built = false;   // build_tc_plan succeeded
  double pre_cof = 0.0;    // pre-emphasis fused into the framing kernels (spb_plan_set_preemphasis); 0: off
  void *d_tc_whi = nullptr, *d_tc_wlo = nullptr;
  std::vector<float> h_tc_tab;        // [2][128][4]: per epilogue half {rise, fall, cos, sin}
  uint32_t tc_endmask[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
  int tc_half[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};  // {first chunk, end chunk, first segment, first filter}
  bool tc_vec4 = false;               // filter ranges of both halves are multiples of four
  uint64_t tc_serial = 0;             // identifies these tables in the per-device constant bank
  // second-generation tensor-core kernel (feat_tc2.cuh): folded DFT matrices C_hi | C_lo | S_hi | S_lo;
  // tc2_static: the bank is the speechpy default one and takes the compile-time epilogue with its own
  // split of the filters (tables *_s)
  bool tc2_ok = false, tc2_static = false;
  void *d_tc2_w = nullptr;
  std::vector<float> h_tc_tab_s;
  uint32_t tc_endmask_s[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
  int tc_half_s[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
  // 2 KB per plan: int flags [0] unsplittable sample seen by the tensor-core kernel, [1] / [2] count /
  // overflow of its redo list, [3] / [4] the same for the CUDA-core FFT kernels; zeros from byte 256 on
  int *d_tc_flag = nullptr;
};

namespace spb {
// everything the FFT kernel needs to know about one 32-frame tile (one 32-byte load, no
// dependent descriptor chain)
struct __align__(16) FftTile {
  int utt, f0, nf, pad;
  long long gstart;  // first sample of the tile in the arena
  long long row0;    // first output row of the tile
};

// tensor-core feature kernel: 32 consecutive hops of one utterance (31 frames); 4 per tile
struct __align__(16) TcGroup {
  long long gstart;  // first sample of the group's first hop in the arena
  long long row0;    // first output row
  int nf;            // frames (0: padding group)
  int pad;
  long long prior;   // arena index of the sample in front of the group's first hop: gstart - 1, or -- first group of an
                     // utterance -- the utterance's LAST sample (what np.roll puts in front of sample 0; pre-emphasis)
};

// Exactness guard of the log-mel kernels: a filter whose fp32 energy is below tau x the level of
// its frame is listed here and recomputed in float64 by redo_list_kernel (redo_kernels.cu).
struct __align__(16) RedoEntry {
  long long soff;  // first sample of the frame in the arena
  long long oidx;  // output element index: row * M + filter
};
struct RedoList {
  RedoEntry *entries = nullptr;  // nullptr: guard off
  int *count = nullptr;          // entries appended so far (may exceed cap)
  int *overflow = nullptr;       // set when an entry did not fit: the guarded redo_all kernel takes over
  int cap = 0;
  float tau = 0.f;               // threshold factor (kernel-specific meaning of the level)
};
__device__ __forceinline__ void redo_push(const RedoList &r, long long soff, long long oidx) {
  const int i = atomicAdd(r.count, 1);
  if (i < r.cap) {
    RedoEntry e;
    e.soff = soff;
    e.oidx = oidx;
    r.entries[i] = e;
  } else {
    *r.overflow = 1;
  }
}

// a run of whole utterances whose intermediates fit in L2
struct BatchChunk {
  int64_t u0, u1;        // utterances [u0, u1)
  int64_t tile0, tile1;  // fft tiles
  int64_t cm0, cm1;      // cmvnw tiles
  int64_t row0, row1;    // output rows
  int64_t big0 = 0, big1 = 0;  // large-tile cmvnw tiles
  int64_t s_lo = 0, s_hi = 0;  // sample span [min start, max end) of the chunk's utterances
  int64_t tc0 = 0, tc1 = 0;    // TcGroup range (multiples of 4)
  int64_t st0 = 0, st1 = 0;    // single-pass cmvnw streams (utterances of >= 150 rows), longest first
  int64_t rt0 = 0, rt1 = 0;    // large-tile cmvnw tiles of the utterances the stream kernel does not take
  int64_t rs0 = 0, rs1 = 0;    // resident single-pass cmvnw: utterances of <= cmvnw_resident_max_rows() rows, longest first
  int64_t lt0 = 0, lt1 = 0;    // large-tile cmvnw tiles of the utterances too long for the resident kernel (> 930 rows)
  // option cmvnw_kernel = 4 only:
  int64_t r20 = 0, r21 = 0;    // resident cmvnw on 2-CTA clusters: utterances of 931 .. 1 860 rows, longest first (in d_res_tiles)
  int64_t vl0 = 0, vl1 = 0;    // large-tile cmvnw tiles of the utterances too long for the cluster pairs as well
};
}  // namespace spb

struct spb_batch {
  spb_context *ctx = nullptr;
  const spb_plan *plan = nullptr;
  int64_t U = 0;
  int64_t total_frames = 0;
  int64_t total_tiles = 0;       // fft tiles
  int64_t total_cm_tiles = 0;    // cmvnw tiles
  int tile_frames = 0;
  int cm_tile_rows = 0;
  std::vector<int64_t> h_frame_offsets;  // [U+1]
  int64_t max_end = 0;                   // max(starts+lengths): arena must be at least this long
  int64_t *d_starts = nullptr;           // [U]
  int32_t *d_nframes = nullptr;          // [U]
  int64_t *d_frame_offsets = nullptr;    // [U+1]
  int64_t *d_last = nullptr;             // [U] arena index of each utterance's last sample (pre-emphasis wraps to it)
  int2 *d_tiles = nullptr;               // [total_tiles] {utt, first frame}
  spb::FftTile *d_ftiles = nullptr;      // [total_tiles] full descriptors for the fast kernel
  int2 *d_cm_tiles = nullptr;            // [total_cm_tiles] {utt, first row}
  spb::TcGroup *d_tc_groups = nullptr;    // tensor-core feature kernel tiles (nullptr: not usable)
  int4 *d_big_tiles = nullptr;           // large-tile cmvnw: {utt, first row, rows, 0}, longest first per chunk
  int big_rmax = 0;                      // 0: large-tile path not usable for this plan
  int4 *d_streams = nullptr;             // single-pass cmvnw: {utt, first row, end row, 0} (nullptr: not usable)
  int4 *d_res_tiles = nullptr;           // resident cmvnw: two int4 per utterance {utt, F, 0, 0}, {first global row lo, hi, 0, 0}
  bool resident_pairs = false;           // the device takes 2-CTA clusters of the resident kernel (option cmvnw_kernel = 4)
  bool has_short_utts = false;           // some utterance has fewer rows than the stream kernel takes
  std::vector<spb::BatchChunk> chunks;   // pipeline units of the host-pointer calls
  std::vector<spb::BatchChunk> groups;   // launch units of the device-pointer calls (runs of chunks)
  int64_t max_group_rows = 0;
  int64_t max_chunk_rows = 0;
  bool chunk_spans_disjoint = true;  // no two chunks' sample spans overlap: the host path may upload chunk by chunk
  int64_t chunk_span_total = 0;  // sum of the chunks' sample spans (== arena use when utterances are packed in order)
};

namespace spb {

// process-wide kernel selection / guard knobs behind spb_set_option (declared in the public header)
struct Options {
  int feature_kernel = 0;  // 0 auto, 1 generic radix-2 kernel, 2 CUDA-core FFT kernels only (no tensor cores)
  int cmvnw_kernel = 0;    // 0 auto (resident single pass for utterances of <= 930 rows, large-tile two-pass for the rest),
                           // 1 small-tile two-pass, 2 large-tile two-pass, 3 single streaming pass
  int tc_version = 2;      // tensor-core lmfe kernel generation: 2 (folded hop DFT) or 1
  int tc_epilogue = -1;    // -1 auto, 0 / 1 / 2: force that epilogue flavour of generation 2 where legal
  int redo = 1;            // float64 recomputation of low-energy filters (the 1e-4 guard)
  int gemm_multicast = 1;  // cosine GEMM: 1 = 2-CTA clusters with the T tiles multicast (109 us vs 114-116 us, same process), 0 = one CTA per tile
  double tau_tc = 2e-5;    // flag when energy < tau x 80 (max_f^2 + max_f+1^2)  (tensor-core kernels)
  double tau_fft = 1e-4;   // flag when energy < tau x sum x^2 of the frame      (FFT kernels)
};
Options &options();

int fail(spb_context *ctx, int code, const std::string &msg);
// timing helpers: begin returns an index to pass to end (or SIZE_MAX when disabled)
size_t timing_begin(spb_context *ctx);
void timing_end(spb_context *ctx, int kernel, size_t begin_idx);
int cuda_fail(spb_context *ctx, cudaError_t e, const char *what);
// returns device pointer of at least `bytes` (grow-only), or nullptr (error set)
void *scratch(spb_context *ctx, int slot, size_t bytes);
int get_fft_tables(spb_context *ctx, int nfft, FftTables *out);
// opt in to `smem` bytes of dynamic shared memory and query occupancy, cached per context
int configure_kernel(spb_context *ctx, const void *func, int threads, size_t smem, int *per_sm);

#define SPB_CUDA(ctx, call)                                        \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return spb::cuda_fail(ctx, e__, #call); \
  } while (0)

#define SPB_LAUNCH_CHECK(ctx, name)                                  \
  do {                                                               \
    (ctx)->launches++;                                               \
    cudaError_t e__ = cudaGetLastError();                            \
    if (e__ != cudaSuccess) return spb::cuda_fail(ctx, e__, name);   \
  } while (0)

// ------------------------------------------------------------ kernel entry points
enum FeatMode { FEAT_POWER = 0, FEAT_MAG = 1, FEAT_MFE = 2, FEAT_LMFE = 3, FEAT_MFCC = 4,
                FEAT_MFCC_DELTA = 5 };

struct FeatArgs {
  const float *samples;      // arena
  int64_t arena_len;
  const int64_t *starts;     // [U]
  const int32_t *nframes;    // [U]
  const int64_t *frame_offsets;  // [U+1]
  const int64_t *last = nullptr; // [U] index of each utterance's last sample (pre-emphasis only)
  int64_t num_utts = 0;
  double pre_cof = 0.0;          // pre-emphasis coefficient (0: off); processing.py:45-58 with shift = 1
  const int2 *tiles;         // [num_tiles]
  const FftTile *ftiles;     // [num_tiles] or nullptr
  const TcGroup *tc_groups = nullptr;  // tensor-core kernel: groups of this launch (4 per tile) or nullptr
  int64_t num_tc_groups = 0;
  float4 *tc_hopsum = nullptr;         // tensor-core kernel: per-hop sums out (mfcc energy), [num_tc_groups * 32]
  RedoList redo_tc, redo_fft;          // exactness guard lists of the tensor-core / FFT kernels (entries == nullptr: off)
  const spb_plan *plan = nullptr;
  const int *run_if = nullptr;         // fast512: run only when *run_if != 0
  int64_t num_tiles;
  int tile_frames;
  int L, S, nfft, K;
  int M, C, dc_elim;
  const int32_t *band_lo, *band_n, *band_off;
  const float *band_w;
  int band_total;
  int max_bin;               // largest bin with non-zero weight
  const float *dct;          // [C][M]
  const float4 *seg_rd;      // segment form of the bank (or nullptr)
  const int4 *seg_ranges;
  int seg_kmin, seg_nbins;
  int mode;
  float *out;                // power/mag: [frames][K]; mfe/lmfe: [frames][M]; mfcc: [frames][C]; delta: [frames][C][3]
  float *energy;             // [frames] or nullptr (mfe only)
};

int launch_features_generic(spb_context *ctx, const FeatArgs &a, const FftTables &t);
bool fast512_supported(const FeatArgs &a);
int launch_features_fast512(spb_context *ctx, const FeatArgs &a, const FftTables &t);
bool features_tc_supported(const FeatArgs &a);
int launch_features_tc(spb_context *ctx, const FeatArgs &a);
bool features_tc_has_guard(const FeatArgs &a);
// float64 recomputation of the values a feature launch listed in `list` (redo_kernels.cu)
int launch_redo(spb_context *ctx, const FeatArgs &a, const RedoList &list, const int *skip_if, const FftTables &t);
// mfcc / mfcc+delta rows from log-mel rows (global row addressing) + frame energies from the samples
int launch_mfcc_post(spb_context *ctx, const FeatArgs &a, const float *d_lmfe);
void tc_plan_apply_preemphasis(spb_plan *pl);  // refresh the tensor-core epilogue's pre-emphasis constants from pl->pre_cof
int build_tc_plan(spb_context *ctx, spb_plan *pl, const float4 *seg_rd);  // seg_rd: host copy, seg_nbins entries
bool fast1024_supported(const FeatArgs &a);
int launch_features_fast1024(spb_context *ctx, const FeatArgs &a, const FftTables &t);

int launch_cmvnw_f32(spb_context *ctx, const float *in, float *scratch_ms, float *out,
                     const int32_t *nframes, const int64_t *frame_offsets, const int2 *tiles,
                     int64_t num_tiles, int tile_rows, int cols, int ld, int win, int varnorm, int first_stage = 1);
int cmvnw_tile_rows(int cols, int win);
int cmvnw_big_rmax(int cols, int win);
int launch_cmvnw_big(spb_context *ctx, const float *in, float *scratch_ms, float *out,
                     const int32_t *nframes, const int64_t *frame_offsets, const int4 *tiles,
                     int64_t num_tiles, int rmax, int cols, int ld, int win, int varnorm, int first_stage = 1);
// stage 1 of cmvnw from float64 rows (float64 window means, one rounding to float32): the public cmvnw call
int launch_cmvnw_stage1_f64(spb_context *ctx, const double *in, int64_t rows, int cols, int win, float *out);
// single-pass streaming cmvnw (cmvn_stream_kernels.cu): window 301, variance normalisation, 40 columns
bool cmvnw_stream_supported(int cols, int win, int varnorm);
int cmvnw_stream_min_rows();
void append_streams(std::vector<int4> &v, int utt, int64_t F);
int launch_cmvnw_stream(spb_context *ctx, const float *in, float *out, const int32_t *nframes,
                        const int64_t *frame_offsets, const int4 *streams, int64_t num_streams);
// whole-utterance-resident single-pass cmvnw (cmvn_resident_kernels.cu): window 301, variance normalisation, 40 columns
bool cmvnw_resident_supported(int cols, int win, int varnorm);
int cmvnw_resident_max_rows();
int launch_cmvnw_resident(spb_context *ctx, const float *in, float *out, const int4 *tiles, int64_t num_tiles, int cluster = 1);
int cmvnw_resident_clusters(spb_context *ctx);  // 2-CTA clusters of the resident kernel the device runs at once (0: none)
// balanced tiles of one utterance of F rows: appends {utt, r0, rows, 0}
void append_big_tiles(std::vector<int4> &v, int utt, int64_t F, int rmax);
bool cmvnw_supported(int cols, int win);
int launch_cmvn_f64(spb_context *ctx, const double *in, int64_t rows, int cols, int varnorm,
                    double *out);
int launch_stack_frames_f64(spb_context *ctx, const double *sig, int64_t n, int L, int S,
                            const double *window, int64_t F, double *out);
int launch_derivative_f64(spb_context *ctx, const double *in, int64_t rows, int cols, int dw,
                          double *out);
int launch_delta_cube_f64(spb_context *ctx, const double *in, int64_t rows, int cols,
                          double *out);
int launch_spectrum_dft(spb_context *ctx, const double *d_frames, int64_t F, int L, int nfft, int mode, double *d_out,
                        double *d_tmp);
int launch_convert(spb_context *ctx, const void *in, int dtype_in, void *out, int dtype_out,
                   int64_t n);
int launch_preemphasis(spb_context *ctx, const void *in, int dtype, long long n, long long shift, double cof,
                       void *out);
int launch_log_power(spb_context *ctx, const float *p, long long n, int normalize, unsigned int *d_max_bits,
                     double *out);
int launch_feature_cubes(spb_context *ctx, const float *d_lmfe, const int64_t *d_frame_offsets, int64_t U, int M,
                         int channels, int normalize, const int32_t *d_crop_starts, int num_crops,
                         int crop_frames, double2 *d_stats, float *d_out);
int launch_ingest(spb_context *ctx, const void *in, int dtype_in, float *out, int64_t n, float scale);
int launch_build_tiles(spb_context *ctx, const int32_t *nframes, const int64_t *tile_offsets,
                       int64_t U, int tile, int2 *tiles);
int launch_build_fft_tiles(spb_context *ctx, const int32_t *nframes, const int64_t *tile_offsets,
                           const int64_t *starts, const int64_t *frame_offsets, int64_t U, int tile,
                           int stride, FftTile *tiles);

int launch_cosine_scores(spb_context *ctx, const float *d_E, int64_t m, const float *d_T,
                         int64_t n, int d, float *d_S);
int launch_trial_gather(spb_context *ctx, const float *d_S, int64_t m, int64_t n,
                        const int32_t *ti, const int32_t *tj, int64_t nt, float *out);
int launch_argmax_rows(spb_context *ctx, const float *d_S, int64_t m, int64_t n, int32_t *out);
int run_eer(spb_context *ctx, const float *d_scores, const int32_t *d_labels, int64_t n,
            double *eer, double *auc);

}  // namespace spb
