// C-ABI layer of libspeechpy_b200.so (declared in include/speechpy_b200.h).
#include <math.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "spb_internal.cuh"

namespace spb {

static thread_local std::string g_last_error;

Options &options() {
  static Options o;
  return o;
}

void set_global_error(const std::string &msg) { g_last_error = msg; }

int fail(spb_context *ctx, int code, const std::string &msg) {
  if (ctx) ctx->last_error = msg;
  g_last_error = msg;
  return code;
}

int cuda_fail(spb_context *ctx, cudaError_t e, const char *what) {
  std::string msg = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
  cudaGetLastError();  // clear sticky-less errors
  return fail(ctx, SPB_ERR_CUDA, msg);
}

void *scratch(spb_context *ctx, int slot, size_t bytes) {
  DeviceBuffer &b = ctx->scratch[slot];
  if (b.bytes >= bytes && b.ptr) return b.ptr;
  if (b.ptr) {
    cudaStreamSynchronize(ctx->stream);
    cudaFree(b.ptr);
    b.ptr = nullptr;
    b.bytes = 0;
  }
  size_t want = std::max(bytes, (size_t)1 << 20);
  want = (want + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
  cudaError_t e = cudaMalloc(&b.ptr, want);
  if (e != cudaSuccess) {
    cuda_fail(ctx, e, "cudaMalloc(scratch)");
    b.ptr = nullptr;
    return nullptr;
  }
  b.bytes = want;
  return b.ptr;
}

static size_t timing_event(spb_context *ctx) {
  if (ctx->ev_used == ctx->ev_pool.size()) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return (size_t)-1;
    ctx->ev_pool.push_back(e);
  }
  const size_t i = ctx->ev_used++;
  cudaEventRecord(ctx->ev_pool[i], ctx->stream);
  return i;
}

size_t timing_begin(spb_context *ctx) {
  if (!ctx->timing) return (size_t)-1;
  return timing_event(ctx);
}

void timing_end(spb_context *ctx, int kernel, size_t begin_idx) {
  if (!ctx->timing || begin_idx == (size_t)-1) return;
  const size_t e = timing_event(ctx);
  if (e != (size_t)-1) ctx->ev_pairs[kernel].push_back({begin_idx, e});
}

int configure_kernel(spb_context *ctx, const void *func, int threads, size_t smem, int *per_sm) {
  auto key = std::make_pair(func, smem);
  auto it = ctx->kernel_cfg.find(key);
  if (it != ctx->kernel_cfg.end()) {
    *per_sm = it->second;
    return SPB_OK;
  }
  // the attribute is a per-function maximum: only ever raise it
  size_t &cur_max = ctx->kernel_smem_max[func];
  if (smem > cur_max) {
    SPB_CUDA(ctx, cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cur_max = smem;
  }
  int n = 0;
  SPB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, func, threads, smem));
  ctx->kernel_cfg[key] = n;
  *per_sm = n;
  return SPB_OK;
}

static bool is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

int get_fft_tables(spb_context *ctx, int nfft, FftTables *out) {
  auto it = ctx->fft_tables.find(nfft);
  if (it != ctx->fft_tables.end()) {
    *out = it->second;
    return SPB_OK;
  }
  FftTables t;
  const double two_pi = 6.283185307179586476925286766559;
  {
    std::vector<float2> tw(nfft / 2);
    for (int k = 0; k < nfft / 2; ++k) {
      const double a = -two_pi * (double)k / (double)nfft;
      tw[k] = make_float2((float)cos(a), (float)sin(a));
    }
    SPB_CUDA(ctx, cudaMalloc(&t.d_tw, tw.size() * sizeof(float2)));
    SPB_CUDA(ctx, cudaMemcpy(t.d_tw, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice));
  }
  {
    std::vector<double2> tw((size_t)nfft);
    for (int r = 0; r < nfft; ++r) {
      const double a = two_pi * (double)r / (double)nfft;
      tw[r] = make_double2(cos(a), -sin(a));
    }
    SPB_CUDA(ctx, cudaMalloc(&t.d_tw64, tw.size() * sizeof(double2)));
    SPB_CUDA(ctx, cudaMemcpy(t.d_tw64, tw.data(), tw.size() * sizeof(double2), cudaMemcpyHostToDevice));
  }
  if (nfft == 512) {
    std::vector<float2> a(16 * 16), b(16 * 9);
    for (int l = 0; l < 16; ++l)
      for (int k1 = 0; k1 < 16; ++k1) {
        const double ang = -two_pi * (double)((l * k1) % 256) / 256.0;
        a[l * 16 + k1] = make_float2((float)cos(ang), (float)sin(ang));
      }
    for (int k1 = 0; k1 < 16; ++k1)
      for (int k2 = 0; k2 < 9; ++k2) {
        const double ang = -two_pi * (double)(k1 + 16 * k2) / 512.0;
        b[k1 * 9 + k2] = make_float2((float)cos(ang), (float)sin(ang));
      }
    SPB_CUDA(ctx, cudaMalloc(&t.d_twA, a.size() * sizeof(float2)));
    SPB_CUDA(ctx, cudaMemcpy(t.d_twA, a.data(), a.size() * sizeof(float2), cudaMemcpyHostToDevice));
    SPB_CUDA(ctx, cudaMalloc(&t.d_twB, b.size() * sizeof(float2)));
    SPB_CUDA(ctx, cudaMemcpy(t.d_twB, b.data(), b.size() * sizeof(float2), cudaMemcpyHostToDevice));
  }
  ctx->fft_tables[nfft] = t;
  *out = t;
  return SPB_OK;
}

// Python-compatible frame count (processing.py:103-104, 115-116): float division, then
// ceil / floor.  Returns false when the count is negative.
static bool frame_count(int64_t n, int L, int S, int zero_padding, int64_t *F) {
  const double q = (double)(n - L) / (double)S;
  const double f = zero_padding ? ceil(q) : floor(q);
  *F = (int64_t)f;
  return f >= 0.0;
}

void append_big_tiles(std::vector<int4> &v, int utt, int64_t F, int rmax) {
  if (F <= 0) return;
  const int64_t nt = (F + rmax - 1) / rmax;
  int64_t r = (F + nt - 1) / nt;
  r = (r + 29) / 30 * 30;  // multiple of the 30-row block
  if (r > rmax) r = rmax;
  for (int64_t r0 = 0; r0 < F; r0 += r) v.push_back(make_int4(utt, (int)r0, (int)std::min<int64_t>(r, F - r0), 0));
}

}  // namespace spb

using namespace spb;

#define SPB_REQUIRE(ctx, cond, msg) \
  do {                              \
    if (!(cond)) return fail(ctx, SPB_ERR_INVALID, msg); \
  } while (0)

extern "C" {

int spb_version(void) { return SPB_VERSION; }

int spb_init(int device, spb_context **out) {
  if (!out) return fail(nullptr, SPB_ERR_INVALID, "spb_init: ctx pointer is NULL");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    return fail(nullptr, SPB_ERR_CUDA,
                std::string("spb_init: no usable CUDA device (") +
                    (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                    "); libspeechpy_b200 has no CPU fallback");
  }
  if (device < 0 || device >= count) return fail(nullptr, SPB_ERR_INVALID, "spb_init: bad device ordinal");
  spb_context *ctx = new (std::nothrow) spb_context();
  if (!ctx) return fail(nullptr, SPB_ERR_NOMEM, "spb_init: out of host memory");
  ctx->device = device;
  cudaDeviceProp prop;
  if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) {
    delete ctx;
    return cuda_fail(nullptr, e, "cudaSetDevice/cudaGetDeviceProperties");
  }
  ctx->sm_count = prop.multiProcessorCount;
  ctx->cc_major = prop.major;
  ctx->cc_minor = prop.minor;
  ctx->total_mem = prop.totalGlobalMem;
  if (prop.major < 10) {
    delete ctx;
    return fail(nullptr, SPB_ERR_UNSUPPORTED,
                "spb_init: this library is built for sm_100a (B200) only");
  }
  if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaStreamCreateWithFlags(&ctx->out_stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaEventCreateWithFlags(&ctx->ev_a, cudaEventDisableTiming)) != cudaSuccess ||
      (e = cudaEventCreateWithFlags(&ctx->ev_b, cudaEventDisableTiming)) != cudaSuccess) {
    delete ctx;
    return cuda_fail(nullptr, e, "stream/event creation");
  }
  *out = ctx;
  return SPB_OK;
}

int spb_destroy(spb_context *ctx) {
  if (!ctx) return SPB_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  cudaStreamSynchronize(ctx->copy_stream);
  cudaStreamSynchronize(ctx->out_stream);
  for (auto &b : ctx->scratch)
    if (b.ptr) cudaFree(b.ptr);
  for (int i = 0; i < 2; ++i)
    if (ctx->pinned[i]) cudaFreeHost(ctx->pinned[i]);
  for (auto e : ctx->ev_pool) cudaEventDestroy(e);
  for (auto e : ctx->pipe_events) cudaEventDestroy(e);
  for (auto &kv : ctx->fft_tables) {
    if (kv.second.d_tw) cudaFree(kv.second.d_tw);
    if (kv.second.d_twA) cudaFree(kv.second.d_twA);
    if (kv.second.d_twB) cudaFree(kv.second.d_twB);
    if (kv.second.d_tw64) cudaFree(kv.second.d_tw64);
  }
  cudaEventDestroy(ctx->ev_a);
  cudaEventDestroy(ctx->ev_b);
  cudaStreamDestroy(ctx->stream);
  cudaStreamDestroy(ctx->copy_stream);
  cudaStreamDestroy(ctx->out_stream);
  delete ctx;
  return SPB_OK;
}

const char *spb_last_error(spb_context *ctx) {
  return ctx ? ctx->last_error.c_str() : g_last_error.c_str();
}

int spb_sync(spb_context *ctx) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_sync: NULL context");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

int spb_stream(spb_context *ctx, uint64_t *stream) {
  SPB_REQUIRE(ctx, ctx && stream, "spb_stream: NULL argument");
  *stream = (uint64_t)(uintptr_t)ctx->stream;
  return SPB_OK;
}

int spb_device_info(spb_context *ctx, int32_t *sm_count, int32_t *cc_major, int32_t *cc_minor,
                    uint64_t *total_mem) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_device_info: NULL context");
  if (sm_count) *sm_count = ctx->sm_count;
  if (cc_major) *cc_major = ctx->cc_major;
  if (cc_minor) *cc_minor = ctx->cc_minor;
  if (total_mem) *total_mem = ctx->total_mem;
  return SPB_OK;
}

int spb_launch_count(spb_context *ctx, uint64_t *count) {
  SPB_REQUIRE(ctx, ctx && count, "spb_launch_count: NULL argument");
  *count = ctx->launches;
  return SPB_OK;
}

int spb_timing_enable(spb_context *ctx, int32_t on) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_timing_enable: NULL context");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->timing = on != 0;
  ctx->ev_used = 0;
  for (auto &v : ctx->ev_pairs) v.clear();
  return SPB_OK;
}

int spb_timing_read(spb_context *ctx, int32_t kernel, double *total_ms, uint64_t *launches) {
  SPB_REQUIRE(ctx, ctx && total_ms && launches, "spb_timing_read: NULL argument");
  SPB_REQUIRE(ctx, kernel >= 0 && kernel < 6, "spb_timing_read: kernel must be 0..5");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  double tot = 0.0;
  for (auto &pr : ctx->ev_pairs[kernel]) {
    float ms = 0.f;
    SPB_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_pool[pr.first], ctx->ev_pool[pr.second]));
    tot += ms;
  }
  *total_ms = tot;
  *launches = ctx->ev_pairs[kernel].size();
  ctx->ev_pairs[kernel].clear();
  bool all_empty = true;
  for (auto &v : ctx->ev_pairs) all_empty = all_empty && v.empty();
  if (all_empty) ctx->ev_used = 0;
  return SPB_OK;
}

int spb_host_alloc(void **ptr, size_t bytes) {
  if (!ptr) return fail(nullptr, SPB_ERR_INVALID, "spb_host_alloc: NULL pointer");
  cudaError_t e = cudaHostAlloc(ptr, bytes, cudaHostAllocDefault);
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaHostAlloc");
  return SPB_OK;
}

int spb_host_free(void *ptr) {
  if (!ptr) return SPB_OK;
  cudaError_t e = cudaFreeHost(ptr);
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaFreeHost");
  return SPB_OK;
}

// ------------------------------------------------------------------ stack_frames
int spb_num_frames(int64_t num_samples, int32_t frame_len, int32_t frame_stride,
                   int32_t zero_padding, int64_t *num_frames) {
  if (!num_frames) return fail(nullptr, SPB_ERR_INVALID, "spb_num_frames: NULL output");
  if (frame_len < 0 || frame_stride < 1 || num_samples < 0)
    return fail(nullptr, SPB_ERR_INVALID, "spb_num_frames: frame_len >= 0, frame_stride >= 1, N >= 0 required");
  int64_t F;
  if (!frame_count(num_samples, frame_len, frame_stride, zero_padding, &F)) {
    *num_frames = F;
    return fail(nullptr, SPB_ERR_NEGATIVE_DIM, "negative dimensions are not allowed");
  }
  *num_frames = F;
  return SPB_OK;
}

int spb_stack_frames(spb_context *ctx, const double *sig, int64_t num_samples, int32_t frame_len,
                     int32_t frame_stride, const double *window, int64_t num_frames, double *out) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_stack_frames: NULL context");
  SPB_REQUIRE(ctx, frame_len >= 1 && frame_stride >= 1, "spb_stack_frames: bad frame geometry");
  SPB_REQUIRE(ctx, num_frames >= 0 && num_samples >= 0, "spb_stack_frames: negative size");
  if (num_frames == 0) return SPB_OK;
  SPB_REQUIRE(ctx, sig && out, "spb_stack_frames: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t out_elems = (size_t)num_frames * frame_len;
  double *d_sig = (double *)scratch(ctx, 0, (size_t)std::max<int64_t>(num_samples, 1) * 8);
  double *d_out = (double *)scratch(ctx, 1, out_elems * 8);
  double *d_win = nullptr;
  if (!d_sig || !d_out) return SPB_ERR_CUDA;
  if (window) {
    d_win = (double *)scratch(ctx, 4, (size_t)frame_len * 8);
    if (!d_win) return SPB_ERR_CUDA;
    SPB_CUDA(ctx, cudaMemcpyAsync(d_win, window, (size_t)frame_len * 8, cudaMemcpyHostToDevice, ctx->stream));
  }
  SPB_CUDA(ctx, cudaMemcpyAsync(d_sig, sig, (size_t)num_samples * 8, cudaMemcpyHostToDevice, ctx->stream));
  int rc = launch_stack_frames_f64(ctx, d_sig, num_samples, frame_len, frame_stride, d_win, num_frames, d_out);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, out_elems * 8, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

// ------------------------------------------------------------------ plans
int spb_plan_create(spb_context *ctx, const spb_feature_params *params, const double *filterbank,
                    spb_plan **plan) {
  SPB_REQUIRE(ctx, ctx && params && plan, "spb_plan_create: NULL argument");
  *plan = nullptr;
  const spb_feature_params &q = *params;
  SPB_REQUIRE(ctx, q.frame_len >= 1 && q.frame_stride >= 1, "plan: frame_len and frame_stride must be >= 1");
  if (!(is_pow2(q.nfft) && q.nfft >= 32 && q.nfft <= 4096))
    return fail(ctx, SPB_ERR_UNSUPPORTED, "plan: fft_length must be a power of two in [32, 4096]");
  SPB_REQUIRE(ctx, q.num_filters >= 0 && q.num_filters <= 256, "plan: num_filters must be in [0, 256]");
  SPB_REQUIRE(ctx, q.num_cepstral >= 0 && q.num_cepstral <= std::max(q.num_filters, 0),
              "plan: num_cepstral must be in [0, num_filters]");
  SPB_REQUIRE(ctx, q.num_filters == 0 || filterbank != nullptr, "plan: filterbank is NULL");
  if (q.cmvn_window != 0) {
    SPB_REQUIRE(ctx, q.cmvn_window > 0 && (q.cmvn_window & 1), "Windows size must be odd!");
  }
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  spb_plan *pl = new (std::nothrow) spb_plan();
  if (!pl) return fail(ctx, SPB_ERR_NOMEM, "plan: out of host memory");
  pl->ctx = ctx;
  pl->p = q;
  pl->K = q.nfft / 2 + 1;
  const int M = q.num_filters, K = pl->K, C = q.num_cepstral;
  std::vector<float> w;
  pl->h_band_lo.assign(M, 0);
  pl->h_band_n.assign(M, 0);
  pl->h_band_off.assign(M, 0);
  for (int m = 0; m < M; ++m) {
    int lo = -1, hi = -1;
    for (int k = 0; k < K; ++k)
      if (filterbank[(size_t)m * K + k] != 0.0) {
        if (lo < 0) lo = k;
        hi = k;
      }
    pl->h_band_off[m] = (int)w.size();
    if (lo >= 0) {
      pl->h_band_lo[m] = lo;
      // taps zero-padded to a multiple of four: the kernels read four weights per load
      const int n4 = (hi - lo + 1 + 3) & ~3;
      pl->h_band_n[m] = n4;
      for (int k = lo; k < lo + n4; ++k)
        w.push_back(k <= hi ? (float)filterbank[(size_t)m * K + k] : 0.f);
      pl->max_bin = std::max(pl->max_bin, hi);
    }
  }
  pl->band_total = (int)w.size();
  // ---- segment form (see spb_plan): validated against the dense bank, else left empty
  std::vector<float4> seg_rd;
  std::vector<int4> seg_ranges;
  if (M >= 1) {
    std::vector<int> mid(M, -1);
    for (int m = 0; m < M; ++m) {
      double best = 0.0;
      for (int k = 0; k < K; ++k)
        if (filterbank[(size_t)m * K + k] > best) { best = filterbank[(size_t)m * K + k]; mid[m] = k; }
    }
    int kmin = K, kmax = -1;
    std::vector<int> seg(K, -1);
    std::vector<float> rw(K, 0.f), dw(K, 0.f);
    bool ok = true;
    for (int k = 0; k < K && ok; ++k) {
      int cnt = 0, jmax = -1;
      for (int m = 0; m < M; ++m)
        if (filterbank[(size_t)m * K + k] != 0.0) { ++cnt; jmax = m; }
      if (cnt == 0) continue;
      if (cnt > 2) { ok = false; break; }
      kmin = std::min(kmin, k);
      kmax = std::max(kmax, k);
      if (cnt == 2) {
        if (filterbank[(size_t)(jmax - 1) * K + k] == 0.0) { ok = false; break; }
        seg[k] = jmax;
        rw[k] = (float)filterbank[(size_t)jmax * K + k];
        dw[k] = (float)filterbank[(size_t)(jmax - 1) * K + k];
      } else if (k <= mid[jmax]) {
        seg[k] = jmax;
        rw[k] = (float)filterbank[(size_t)jmax * K + k];
      } else {
        seg[k] = jmax + 1;
        dw[k] = (float)filterbank[(size_t)jmax * K + k];
      }
    }
    if (ok && kmax >= kmin) {
      // bins without weight inside the range inherit the previous segment; segments must run
      // 0, 1, ..., M without gaps
      int prev = 0;
      for (int k = kmin; k <= kmax && ok; ++k) {
        if (seg[k] < 0) seg[k] = prev;
        if (seg[k] != prev && seg[k] != prev + 1) ok = false;
        if (k == kmin && seg[k] != 0) ok = false;
        prev = seg[k];
      }
      if (ok && prev != M) ok = false;
      // exact reconstruction of the fp32 bank
      for (int m = 0; m < M && ok; ++m)
        for (int k = 0; k < K && ok; ++k) {
          float wgt = 0.f;
          if (k >= kmin && k <= kmax) {
            if (seg[k] == m) wgt = rw[k];
            else if (seg[k] == m + 1) wgt = dw[k];
          }
          if (wgt != (float)filterbank[(size_t)m * K + k]) ok = false;
        }
    } else {
      ok = false;
    }
    if (ok) {
      const int nb = kmax - kmin + 1;
      seg_rd.resize(nb);
      for (int k = kmin; k <= kmax; ++k) {
        const bool end = (k == kmax) || (seg[k + 1] != seg[k]);
        seg_rd[k - kmin] = make_float4(rw[k], dw[k], end ? 1.f : 0.f, 0.f);
      }
      // eight warp ranges of whole segments, balanced by bin count
      seg_ranges.assign(8, make_int4(kmax + 1, kmax + 1, M + 1, 0));
      int k = kmin;
      for (int wv = 0; wv < 8 && k <= kmax; ++wv) {
        const int remaining_w = 8 - wv;
        const int target = (kmax + 1 - k + remaining_w - 1) / remaining_w;
        int e = k;
        while (e <= kmax && ((e - k) < target || seg_rd[e - 1 - kmin].z == 0.f)) ++e;  // stop after a segment end
        if (wv == 7) e = kmax + 1;
        seg_ranges[wv] = make_int4(k, e, seg[k], 0);
        k = e;
      }
      pl->seg_kmin = kmin;
      pl->seg_nbins = nb;
    }
  }
  std::vector<float> dct((size_t)C * M);
  const double pi = 3.14159265358979323846264338327950288;
  for (int c = 0; c < C; ++c) {
    const double s = (c == 0) ? sqrt(1.0 / M) : sqrt(2.0 / M);
    for (int m = 0; m < M; ++m) dct[(size_t)c * M + m] = (float)(s * cos(pi * c * (2.0 * m + 1.0) / (2.0 * M)));
  }
  auto up = [&](void **dptr, const void *src, size_t bytes) -> cudaError_t {
    cudaError_t e = cudaMalloc(dptr, std::max(bytes, (size_t)16));
    if (e != cudaSuccess) return e;
    if (bytes) e = cudaMemcpy(*dptr, src, bytes, cudaMemcpyHostToDevice);
    return e;
  };
  cudaError_t e;
  if ((e = up((void **)&pl->d_band_lo, pl->h_band_lo.data(), (size_t)M * 4)) != cudaSuccess ||
      (e = up((void **)&pl->d_band_n, pl->h_band_n.data(), (size_t)M * 4)) != cudaSuccess ||
      (e = up((void **)&pl->d_band_off, pl->h_band_off.data(), (size_t)M * 4)) != cudaSuccess ||
      (e = up((void **)&pl->d_band_w, w.data(), w.size() * 4)) != cudaSuccess ||
      (e = up((void **)&pl->d_seg_rd, seg_rd.data(), seg_rd.size() * sizeof(float4))) != cudaSuccess ||
      (e = up((void **)&pl->d_seg_ranges, seg_ranges.data(), seg_ranges.size() * sizeof(int4))) != cudaSuccess ||
      (e = up((void **)&pl->d_dct, dct.data(), dct.size() * 4)) != cudaSuccess) {
    spb_plan_destroy(pl);
    return cuda_fail(ctx, e, "plan upload");
  }
  if ((e = cudaMalloc(&pl->d_tc_flag, 2048)) != cudaSuccess || (e = cudaMemset(pl->d_tc_flag, 0, 2048)) != cudaSuccess) {
    spb_plan_destroy(pl);
    return cuda_fail(ctx, e, "plan flags");
  }
  FftTables t;
  int rc = get_fft_tables(ctx, q.nfft, &t);
  if (!rc && pl->seg_nbins > 0) rc = build_tc_plan(ctx, pl, seg_rd.data());
  if (rc) {
    spb_plan_destroy(pl);
    return rc;
  }
  *plan = pl;
  return SPB_OK;
}

// Pre-emphasis as a prologue of the framing (processing.py:45-58 with shift = 1): the
// feature launches of this plan see y[i] = x[i] - cof x[i-1] (y[0] wraps to the utterance's last sample, as
// np.roll does) without the emphasised signal ever being written.  Fused into the CUDA-core FFT kernels
// (staged tile transformed in shared memory) and, for log-mel rows of the default bank, into the epilogue of the
// tensor-core kernel (spectral form, feat_tc2.cuh t2_pre_power).  cof = 0 switches it off again.
int spb_plan_set_preemphasis(spb_plan *pl, double cof) {
  if (!pl) return SPB_ERR_INVALID;
  spb_context *ctx = pl->ctx;
  SPB_REQUIRE(ctx, isfinite(cof), "plan_set_preemphasis: cof must be finite");
  pl->pre_cof = cof;
  tc_plan_apply_preemphasis(pl);
  return SPB_OK;
}

int spb_plan_destroy(spb_plan *pl) {
  if (!pl) return SPB_OK;
  if (pl->ctx) cudaSetDevice(pl->ctx->device);
  cudaFree(pl->d_band_lo);
  cudaFree(pl->d_band_n);
  cudaFree(pl->d_band_off);
  cudaFree(pl->d_band_w);
  cudaFree(pl->d_seg_rd);
  cudaFree(pl->d_seg_ranges);
  cudaFree(pl->d_dct);
  cudaFree(pl->d_tc_whi);
  cudaFree(pl->d_tc_wlo);
  cudaFree(pl->d_tc2_w);
  cudaFree(pl->d_tc_flag);
  delete pl;
  return SPB_OK;
}

// ------------------------------------------------------------------ batches
static constexpr int kFftTile = spb::kFftTileFrames;
static constexpr int64_t kChunkRowsDefault = 192 * 1024;  // pipeline unit of the host-pointer calls (~30 MB of rows)
static constexpr int64_t kGroupRowsDefault = 32 * 1024 * 1024;  // launch unit of the device-pointer calls (5.4 GB x 2 of scratch at most; measured r02ba: 4 Mi rows 48.2 ms, 8 Mi 48.1, 16 Mi 47.8, 32 Mi 47.5 per cfg3 step)
// tuning knobs (read when a batch is created): SPB_CHUNK_ROWS, SPB_GROUP_ROWS, SPB_BIG_RMAX
static int64_t env_int(const char *name, int64_t dflt) {
  const char *v = getenv(name);
  if (!v || !*v) return dflt;
  char *end = nullptr;
  const long long x = strtoll(v, &end, 10);
  return (end && *end == 0 && x > 0) ? (int64_t)x : dflt;
}

int spb_batch_create(spb_context *ctx, const spb_plan *plan, const int64_t *starts,
                     const int64_t *lengths, int64_t U, spb_batch **out) {
  SPB_REQUIRE(ctx, ctx && plan && out, "spb_batch_create: NULL argument");
  SPB_REQUIRE(ctx, U >= 0 && (U == 0 || (starts && lengths)), "spb_batch_create: bad utterance table");
  SPB_REQUIRE(ctx, U < ((int64_t)1 << 31), "spb_batch_create: too many utterances");
  *out = nullptr;
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const int L = plan->p.frame_len, S = plan->p.frame_stride;
  spb_batch *b = new (std::nothrow) spb_batch();
  if (!b) return fail(ctx, SPB_ERR_NOMEM, "batch: out of host memory");
  b->ctx = ctx;
  b->plan = plan;
  b->U = U;
  b->tile_frames = kFftTile;
  const int win = plan->p.cmvn_window;
  const int64_t chunk_rows = env_int("SPB_CHUNK_ROWS", kChunkRowsDefault);
  b->cm_tile_rows = (win > 0) ? cmvnw_tile_rows(plan->p.num_filters, win) : 0;
  std::vector<int32_t> nframes(U);
  std::vector<int64_t> tile_off(U + 1, 0), cm_off(U + 1, 0);
  b->h_frame_offsets.assign(U + 1, 0);
  spb_batch &ch = *b;
  BatchChunk cur;
  cur.u0 = cur.u1 = cur.tile0 = cur.tile1 = cur.cm0 = cur.cm1 = cur.row0 = cur.row1 = 0;
  for (int64_t u = 0; u < U; ++u) {
    int64_t F;
    if (starts[u] < 0 || lengths[u] < 0) {
      delete b;
      return fail(ctx, SPB_ERR_INVALID, "batch: negative start or length");
    }
    if (!frame_count(lengths[u], L, S, 0, &F)) {
      delete b;
      return fail(ctx, SPB_ERR_NEGATIVE_DIM, "negative dimensions are not allowed");
    }
    if (F >= ((int64_t)1 << 31) - 64) {
      delete b;
      return fail(ctx, SPB_ERR_INVALID, "batch: utterance too long");
    }
    nframes[u] = (int32_t)F;
    b->h_frame_offsets[u + 1] = b->h_frame_offsets[u] + F;
    tile_off[u + 1] = tile_off[u] + (F + kFftTile - 1) / kFftTile;
    cm_off[u + 1] = cm_off[u] + (b->cm_tile_rows ? (F + b->cm_tile_rows - 1) / b->cm_tile_rows : 0);
    b->max_end = std::max(b->max_end, starts[u] + lengths[u]);
    // chunking by whole utterances
    if (b->h_frame_offsets[u + 1] - cur.row0 > chunk_rows && u > cur.u0) {
      cur.u1 = u;
      cur.tile1 = tile_off[u];
      cur.cm1 = cm_off[u];
      cur.row1 = b->h_frame_offsets[u];
      ch.chunks.push_back(cur);
      ch.max_chunk_rows = std::max(ch.max_chunk_rows, cur.row1 - cur.row0);
      {
        BatchChunk nx;
        nx.u0 = nx.u1 = u;
        nx.tile0 = nx.tile1 = tile_off[u];
        nx.cm0 = nx.cm1 = cm_off[u];
        nx.row0 = nx.row1 = cur.row1;
        cur = nx;
      }
    }
  }
  cur.u1 = U;
  cur.tile1 = tile_off[U];
  cur.cm1 = cm_off[U];
  cur.row1 = b->h_frame_offsets[U];
  if (cur.u1 > cur.u0) {
    ch.chunks.push_back(cur);
    ch.max_chunk_rows = std::max(ch.max_chunk_rows, cur.row1 - cur.row0);
  }
  b->total_frames = b->h_frame_offsets[U];
  b->total_tiles = tile_off[U];
  for (BatchChunk &c : ch.chunks) {
    c.s_lo = INT64_MAX;
    c.s_hi = 0;
    for (int64_t u = c.u0; u < c.u1; ++u) {
      c.s_lo = std::min(c.s_lo, starts[u]);
      c.s_hi = std::max(c.s_hi, starts[u] + lengths[u]);
    }
    b->chunk_span_total += c.s_hi - c.s_lo;
  }
  b->chunk_spans_disjoint = true;
  for (size_t k = 1; k < ch.chunks.size(); ++k)
    for (size_t j = 0; j < k && b->chunk_spans_disjoint; ++j)
      if (ch.chunks[k].s_lo < ch.chunks[j].s_hi && ch.chunks[j].s_lo < ch.chunks[k].s_hi) b->chunk_spans_disjoint = false;
  b->total_cm_tiles = cm_off[U];
  // large-tile cmvnw table: balanced tiles per utterance, longest first inside each chunk
  std::vector<int4> big;
  b->big_rmax = (win > 0) ? cmvnw_big_rmax(plan->p.num_filters, win) : 0;
  const int tile_target = (int)std::min<int64_t>(b->big_rmax, std::max<int64_t>(30, env_int("SPB_BIG_RMAX", b->big_rmax) / 30 * 30));
  // single-pass stream table (+ large tiles for the utterances too short for it), per chunk / group
  std::vector<int4> streams, res;
  const bool resident_ok = b->big_rmax > 0 && cmvnw_resident_supported(plan->p.num_filters, win, plan->p.cmvn_variance);
  const bool resident_pairs = resident_ok && cmvnw_resident_clusters(ctx) >= 8;
  b->resident_pairs = resident_pairs;
  const bool stream_ok = b->big_rmax > 0 && cmvnw_stream_supported(plan->p.num_filters, win, plan->p.cmvn_variance);
  auto add_cmvnw_tables = [&](BatchChunk &c) {
    if (b->big_rmax <= 0) return;
    c.big0 = (int64_t)big.size();
    for (int64_t u = c.u0; u < c.u1; ++u) append_big_tiles(big, (int)u, nframes[u], tile_target);
    c.big1 = (int64_t)big.size();
    std::stable_sort(big.begin() + c.big0, big.begin() + c.big1, [](const int4 &x, const int4 &y) { return x.z > y.z; });
    if (resident_ok) {
      c.rs0 = (int64_t)res.size() / 2;
      std::vector<std::pair<int, int>> order;  // (F, utt)
      for (int64_t u = c.u0; u < c.u1; ++u)
        if (nframes[u] >= 1 && nframes[u] <= cmvnw_resident_max_rows()) order.push_back({nframes[u], (int)u});
      std::stable_sort(order.begin(), order.end(), [](const std::pair<int, int> &x, const std::pair<int, int> &y) { return x.first > y.first; });
      for (const auto &o : order) {
        const long long base = b->h_frame_offsets[o.second];
        res.push_back(make_int4(o.second, o.first, 0, 0));
        res.push_back(make_int4((int)(unsigned)(base & 0xffffffffll), (int)(base >> 32), 0, 0));
      }
      c.rs1 = (int64_t)res.size() / 2;
      // the longer ones: large tiles of the two-pass kernels
      const int max1 = cmvnw_resident_max_rows();
      c.lt0 = (int64_t)big.size();
      for (int64_t u = c.u0; u < c.u1; ++u)
        if (nframes[u] > max1) append_big_tiles(big, (int)u, nframes[u], tile_target);
      c.lt1 = (int64_t)big.size();
      std::stable_sort(big.begin() + c.lt0, big.begin() + c.lt1, [](const int4 &x, const int4 &y) { return x.z > y.z; });
      // option cmvnw_kernel = 4: utterances of up to twice the resident limit go to a cluster of two
      // CTAs each (when the device takes such clusters), only those beyond to the two-pass kernels
      c.r20 = c.r21 = (int64_t)res.size() / 2;
      c.vl0 = c.vl1 = (int64_t)big.size();
      if (resident_pairs) {
        order.clear();
        for (int64_t u = c.u0; u < c.u1; ++u)
          if (nframes[u] > max1 && nframes[u] <= 2 * max1) order.push_back({nframes[u], (int)u});
        std::stable_sort(order.begin(), order.end(), [](const std::pair<int, int> &x, const std::pair<int, int> &y) { return x.first > y.first; });
        for (const auto &o : order) {
          const long long base = b->h_frame_offsets[o.second];
          res.push_back(make_int4(o.second, o.first, 0, 0));
          res.push_back(make_int4((int)(unsigned)(base & 0xffffffffll), (int)(base >> 32), 0, 0));
        }
        c.r21 = (int64_t)res.size() / 2;
        for (int64_t u = c.u0; u < c.u1; ++u)
          if (nframes[u] > 2 * max1) append_big_tiles(big, (int)u, nframes[u], tile_target);
        c.vl1 = (int64_t)big.size();
        std::stable_sort(big.begin() + c.vl0, big.begin() + c.vl1, [](const int4 &x, const int4 &y) { return x.z > y.z; });
      }
    }
    if (!stream_ok) return;
    c.st0 = (int64_t)streams.size();
    for (int64_t u = c.u0; u < c.u1; ++u) append_streams(streams, (int)u, nframes[u]);
    c.st1 = (int64_t)streams.size();
    std::stable_sort(streams.begin() + c.st0, streams.begin() + c.st1,
                     [](const int4 &x, const int4 &y) { return x.z - x.y > y.z - y.y; });
    c.rt0 = (int64_t)big.size();
    for (int64_t u = c.u0; u < c.u1; ++u)
      if (nframes[u] < cmvnw_stream_min_rows()) append_big_tiles(big, (int)u, nframes[u], tile_target);
    c.rt1 = (int64_t)big.size();
    if (c.rt1 > c.rt0) b->has_short_utts = true;
  };
  for (BatchChunk &c : ch.chunks) add_cmvnw_tables(c);
  // tensor-core feature kernel: groups of 32 hops (31 frames), 4 per tile, each chunk padded to
  // whole tiles; needs 16-byte aligned utterance starts
  std::vector<TcGroup> tcg;
  bool tc_use = plan->tc_ok;
  for (int64_t u = 0; u < U && tc_use; ++u) tc_use = (starts[u] & 3) == 0;
  if (tc_use) {
    for (BatchChunk &c : ch.chunks) {
      c.tc0 = (int64_t)tcg.size();
      for (int64_t u = c.u0; u < c.u1; ++u)
        for (int64_t f0 = 0; f0 < nframes[u]; f0 += kTcGroupFrames) {
          TcGroup g;
          g.gstart = starts[u] + f0 * S;
          g.row0 = b->h_frame_offsets[u] + f0;
          g.nf = (int)std::min<int64_t>(kTcGroupFrames, nframes[u] - f0);
          g.pad = 0;
          g.prior = (f0 == 0) ? starts[u] + std::max<int64_t>(lengths[u], 1) - 1 : g.gstart - 1;
          tcg.push_back(g);
        }
      while (tcg.size() & 3) {
        TcGroup g;
        g.gstart = 0; g.row0 = 0; g.nf = 0; g.pad = 0; g.prior = 0;
        tcg.push_back(g);
      }
      c.tc1 = (int64_t)tcg.size();
    }
  }
  // launch groups of the device-pointer path: runs of chunks of up to group_rows rows.  A launch
  // over ~200 Ki rows is only ~2 tiles per SM for the cmvnw kernels and pays a full tail per
  // kernel; measured, large groups win although their intermediates no longer fit L2.
  {
    const int64_t group_rows = env_int("SPB_GROUP_ROWS", kGroupRowsDefault);
    BatchChunk g;
    bool open = false;
    for (const BatchChunk &c : ch.chunks) {
      if (open && c.row1 - g.row0 > group_rows) {
        b->groups.push_back(g);
        open = false;
      }
      if (!open) {
        g = c;
        open = true;
      } else {
        g.u1 = c.u1; g.tile1 = c.tile1; g.cm1 = c.cm1; g.row1 = c.row1; g.tc1 = c.tc1;
        g.s_lo = std::min(g.s_lo, c.s_lo);
        g.s_hi = std::max(g.s_hi, c.s_hi);
      }
    }
    if (open) b->groups.push_back(g);
    for (BatchChunk &c : b->groups) {
      b->max_group_rows = std::max(b->max_group_rows, c.row1 - c.row0);
      add_cmvnw_tables(c);  // group tables follow the chunk tables in d_big_tiles / d_streams
    }
  }
  auto cleanup = [&](cudaError_t e, const char *what) {
    spb_batch_destroy(b);
    return cuda_fail(ctx, e, what);
  };
  cudaError_t e;
  const size_t u_bytes = (size_t)std::max<int64_t>(U, 1);
  int64_t *d_tile_off = nullptr, *d_cm_off = nullptr;
  if ((e = cudaMalloc(&b->d_starts, u_bytes * 8)) != cudaSuccess ||
      (e = cudaMalloc(&b->d_nframes, u_bytes * 4)) != cudaSuccess ||
      (e = cudaMalloc(&b->d_frame_offsets, (u_bytes + 1) * 8)) != cudaSuccess ||
      (e = cudaMalloc(&b->d_last, u_bytes * 8)) != cudaSuccess ||
      (e = cudaMalloc(&b->d_tiles, (size_t)std::max<int64_t>(b->total_tiles, 1) * sizeof(int2))) != cudaSuccess ||
      (e = cudaMalloc(&b->d_ftiles, (size_t)std::max<int64_t>(b->total_tiles, 1) * sizeof(FftTile))) != cudaSuccess ||
      (e = cudaMalloc(&b->d_cm_tiles, (size_t)std::max<int64_t>(b->total_cm_tiles, 1) * sizeof(int2))) != cudaSuccess ||
      (e = cudaMalloc(&b->d_big_tiles, std::max<size_t>(big.size(), 1) * sizeof(int4))) != cudaSuccess ||
      (resident_ok && (e = cudaMalloc(&b->d_res_tiles, std::max<size_t>(res.size(), 1) * sizeof(int4))) != cudaSuccess) ||
      (stream_ok && (e = cudaMalloc(&b->d_streams, std::max<size_t>(streams.size(), 1) * sizeof(int4))) != cudaSuccess) ||
      (!tcg.empty() && (e = cudaMalloc(&b->d_tc_groups, tcg.size() * sizeof(TcGroup))) != cudaSuccess) ||
      (e = cudaMalloc(&d_tile_off, (u_bytes + 1) * 8)) != cudaSuccess ||
      (e = cudaMalloc(&d_cm_off, (u_bytes + 1) * 8)) != cudaSuccess)
    return cleanup(e, "cudaMalloc(batch)");
  std::vector<int64_t> h_last((size_t)std::max<int64_t>(U, 1), 0);
  for (int64_t u = 0; u < U; ++u) h_last[u] = starts[u] + std::max<int64_t>(lengths[u], 1) - 1;
  if (U > 0) {
    if ((e = cudaMemcpyAsync(b->d_starts, starts, (size_t)U * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess ||
        (e = cudaMemcpyAsync(b->d_last, h_last.data(), (size_t)U * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess ||
        (e = cudaMemcpyAsync(b->d_nframes, nframes.data(), (size_t)U * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess ||
        (e = cudaMemcpyAsync(b->d_frame_offsets, b->h_frame_offsets.data(), (size_t)(U + 1) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess ||
        (e = cudaMemcpyAsync(d_tile_off, tile_off.data(), (size_t)(U + 1) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess ||
        (e = cudaMemcpyAsync(d_cm_off, cm_off.data(), (size_t)(U + 1) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess)
      return cleanup(e, "cudaMemcpyAsync(batch)");
    if (!tcg.empty() &&
        (e = cudaMemcpyAsync(b->d_tc_groups, tcg.data(), tcg.size() * sizeof(TcGroup), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess)
      return cleanup(e, "cudaMemcpyAsync(tc groups)");
    if (!big.empty() &&
        (e = cudaMemcpyAsync(b->d_big_tiles, big.data(), big.size() * sizeof(int4), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess)
      return cleanup(e, "cudaMemcpyAsync(big tiles)");
    if (b->d_res_tiles && !res.empty() &&
        (e = cudaMemcpyAsync(b->d_res_tiles, res.data(), res.size() * sizeof(int4), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess)
      return cleanup(e, "cudaMemcpyAsync(resident tiles)");
    if (b->d_streams && !streams.empty() &&
        (e = cudaMemcpyAsync(b->d_streams, streams.data(), streams.size() * sizeof(int4), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess)
      return cleanup(e, "cudaMemcpyAsync(streams)");
    int rc = launch_build_tiles(ctx, b->d_nframes, d_tile_off, U, kFftTile, b->d_tiles);
    if (!rc)
      rc = launch_build_fft_tiles(ctx, b->d_nframes, d_tile_off, b->d_starts, b->d_frame_offsets, U, kFftTile, S,
                                  b->d_ftiles);
    if (!rc && b->cm_tile_rows)
      rc = launch_build_tiles(ctx, b->d_nframes, d_cm_off, U, b->cm_tile_rows, b->d_cm_tiles);
    if (rc) {
      spb_batch_destroy(b);
      return rc;
    }
  }
  // the host vectors above are pageable: wait for the copies before they go out of scope
  e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d_tile_off);
  cudaFree(d_cm_off);
  if (e != cudaSuccess) return cleanup(e, "cudaStreamSynchronize(batch)");
  *out = b;
  return SPB_OK;
}

int spb_batch_destroy(spb_batch *b) {
  if (!b) return SPB_OK;
  if (b->ctx) cudaSetDevice(b->ctx->device);
  cudaFree(b->d_starts);
  cudaFree(b->d_nframes);
  cudaFree(b->d_frame_offsets);
  cudaFree(b->d_last);
  cudaFree(b->d_tiles);
  cudaFree(b->d_ftiles);
  cudaFree(b->d_tc_groups);
  cudaFree(b->d_cm_tiles);
  cudaFree(b->d_big_tiles);
  cudaFree(b->d_streams);
  cudaFree(b->d_res_tiles);
  delete b;
  return SPB_OK;
}

int spb_batch_total_frames(const spb_batch *b, int64_t *total_frames) {
  if (!b || !total_frames) return fail(nullptr, SPB_ERR_INVALID, "spb_batch_total_frames: NULL argument");
  *total_frames = b->total_frames;
  return SPB_OK;
}

int spb_batch_frame_offsets(const spb_batch *b, int64_t *frame_offsets) {
  if (!b || !frame_offsets) return fail(nullptr, SPB_ERR_INVALID, "spb_batch_frame_offsets: NULL argument");
  memcpy(frame_offsets, b->h_frame_offsets.data(), (size_t)(b->U + 1) * 8);
  return SPB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ feature launches
namespace spb {

static void fill_feat_args(FeatArgs &a, const spb_batch *b, const float *d_samples, int64_t arena_len,
                           int mode, float *d_out, float *d_energy) {
  const spb_plan *pl = b->plan;
  a.samples = d_samples;
  a.arena_len = arena_len;
  a.starts = b->d_starts;
  a.nframes = b->d_nframes;
  a.frame_offsets = b->d_frame_offsets;
  a.last = b->d_last;
  a.num_utts = b->U;
  a.pre_cof = pl->pre_cof;
  a.tiles = b->d_tiles;
  a.ftiles = b->d_ftiles;
  a.plan = pl;
  a.tc_groups = nullptr;  // set per launch range (chunk / group)
  a.num_tc_groups = 0;
  a.run_if = nullptr;
  a.num_tiles = b->total_tiles;
  a.tile_frames = b->tile_frames;
  a.L = pl->p.frame_len;
  a.S = pl->p.frame_stride;
  a.nfft = pl->p.nfft;
  a.K = pl->K;
  a.M = pl->p.num_filters;
  a.C = pl->p.num_cepstral;
  a.dc_elim = pl->p.dc_elimination;
  a.band_lo = pl->d_band_lo;
  a.band_n = pl->d_band_n;
  a.band_off = pl->d_band_off;
  a.band_w = pl->d_band_w;
  a.band_total = pl->band_total;
  a.max_bin = pl->max_bin;
  a.seg_rd = pl->seg_nbins > 0 ? pl->d_seg_rd : nullptr;
  a.seg_ranges = pl->d_seg_ranges;
  a.seg_kmin = pl->seg_kmin;
  a.seg_nbins = pl->seg_nbins;
  a.dct = pl->d_dct;
  a.mode = mode;
  a.out = d_out;
  a.energy = d_energy;
}

// Exactness guard lists of one feature launch: scratch slots 8 (tensor-core kernel) and 9 (FFT kernels)
// hold the entries, the plan's flag block the counters.  Capacity: a quarter of the launch's rows
// (noise flags ~1e-4 of the values); what does not fit sets the overflow flag (redo_all_kernel).
static int setup_redo(spb_context *ctx, FeatArgs &a, int64_t rows, bool tc) {
  a.redo_tc = RedoList();
  a.redo_fft = RedoList();
  const Options &opt = options();
  if (!opt.redo || !a.plan || a.M < 1 || a.mode < FEAT_MFE) return SPB_OK;
  // mfcc modes: the FFT kernels list whole rows (redo_kernels.cu recomputes them in float64); the tensor-core kernel
  // only ever produces mfe / lmfe rows
  if (a.mode >= FEAT_MFCC && (a.C < 1 || a.dct == nullptr)) return SPB_OK;
  if (a.mode >= FEAT_MFCC) tc = false;
  const int64_t cap = std::min<int64_t>(std::max<int64_t>(rows / 4, 65536), (int64_t)1 << 26);
  int *flags = a.plan->d_tc_flag;
  if (tc) {
    RedoEntry *e = (RedoEntry *)scratch(ctx, 8, (size_t)cap * sizeof(RedoEntry));
    if (!e) return SPB_ERR_CUDA;
    a.redo_tc.entries = e;
    a.redo_tc.count = flags + 1;
    a.redo_tc.overflow = flags + 2;
    a.redo_tc.cap = (int)cap;
    a.redo_tc.tau = (float)opt.tau_tc;
  }
  RedoEntry *e = (RedoEntry *)scratch(ctx, 9, (size_t)cap * sizeof(RedoEntry));
  if (!e) return SPB_ERR_CUDA;
  a.redo_fft.entries = e;
  a.redo_fft.count = flags + 3;
  a.redo_fft.overflow = flags + 4;
  a.redo_fft.cap = (int)cap;
  a.redo_fft.tau = (float)opt.tau_fft;
  return SPB_OK;
}

static int run_features(spb_context *ctx, FeatArgs &a) {
  if (a.num_tiles <= 0) return SPB_OK;
  FftTables t;
  int rc = get_fft_tables(ctx, a.nfft, &t);
  if (rc) return rc;
  const Options &opt = options();
  const bool generic = opt.feature_kernel == 1;
  const bool use_tc = opt.feature_kernel == 0 && features_tc_supported(a) && fast512_supported(a);
  const int64_t rows = a.num_tiles * (int64_t)a.tile_frames;
  if ((rc = setup_redo(ctx, a, rows, use_tc && features_tc_has_guard(a)))) return rc;
  const size_t tb = timing_begin(ctx);
  if (a.plan) SPB_CUDA(ctx, cudaMemsetAsync(a.plan->d_tc_flag, 0, 32, ctx->stream));  // flag + guard counters
  if (use_tc) {
    // tensor-core kernel; samples it cannot split (|x| >= 2, NaN, inf) raise the plan's flag
    // and the guarded CUDA-core kernel redoes the launch
    rc = launch_features_tc(ctx, a);
    if (!rc) {
      a.run_if = a.plan->d_tc_flag;
      rc = launch_features_fast512(ctx, a, t);
      a.run_if = nullptr;
    }
    if (!rc) rc = launch_redo(ctx, a, a.redo_tc, a.plan->d_tc_flag, t);
  } else if (!generic && fast512_supported(a)) rc = launch_features_fast512(ctx, a, t);
  else if (!generic && fast1024_supported(a)) rc = launch_features_fast1024(ctx, a, t);
  else rc = launch_features_generic(ctx, a, t);
  if (!rc) rc = launch_redo(ctx, a, a.redo_fft, nullptr, t);
  timing_end(ctx, 0, tb);
  return rc;
}

static int check_batch(spb_context *ctx, const spb_batch *b, int64_t arena_len, const void *in,
                       const void *out) {
  SPB_REQUIRE(ctx, ctx && b, "batch call: NULL context or batch");
  SPB_REQUIRE(ctx, b->ctx == ctx, "batch call: batch belongs to another context");
  SPB_REQUIRE(ctx, arena_len >= b->max_end, "batch call: sample arena shorter than starts+lengths");
  if (b->total_frames > 0) SPB_REQUIRE(ctx, in && out, "batch call: NULL buffer");
  return SPB_OK;
}

static bool cmvnw_single_pass(const spb_batch *b, const float *d_out) {
  return b->d_streams != nullptr && options().cmvnw_kernel == 3 && (((size_t)d_out) & 15) == 0;
}

// lmfe -> cmvnw of one run of whole utterances (a chunk of the host-pointer pipeline or a launch
// group of the device-pointer path); the lmfe intermediate of the run (and, for the two-pass cmvnw
// kernels, the mean-subtracted one) lives in a scratch buffer sized for the largest run.
static int lmfe_cmvnw_chunk(spb_context *ctx, const spb_batch *b, const BatchChunk &c,
                            const float *d_samples, int64_t arena_len, float *d_lmfe, float *d_ms,
                            float *d_out) {
  const spb_plan *pl = b->plan;
  const int M = pl->p.num_filters, win = pl->p.cmvn_window;
  FeatArgs a;
  // rows are addressed by their global index; bias the scratch pointers so that row
  // c.row0 lands at the start of the chunk buffer
  float *lmfe_biased = d_lmfe - (ptrdiff_t)c.row0 * M;
  fill_feat_args(a, b, d_samples, arena_len, FEAT_LMFE, lmfe_biased, nullptr);
  a.tiles = b->d_tiles + c.tile0;
  a.ftiles = b->d_ftiles + c.tile0;
  a.num_tiles = c.tile1 - c.tile0;
  if (b->d_tc_groups) {
    a.tc_groups = b->d_tc_groups + c.tc0;
    a.num_tc_groups = c.tc1 - c.tc0;
  }
  int rc = run_features(ctx, a);
  if (rc) return rc;
  float *ms_biased = d_ms ? d_ms - (ptrdiff_t)c.row0 * M : nullptr;
  if (cmvnw_single_pass(b, d_out)) {
    // one streaming pass (cmvn_stream_kernels.cu, opt-in: measured slower than the two passes);
    // the rare utterance of < 150 rows takes the large-tile kernels
    rc = launch_cmvnw_stream(ctx, lmfe_biased, d_out, b->d_nframes, b->d_frame_offsets, b->d_streams + c.st0,
                             c.st1 - c.st0);
    if (rc || c.rt1 == c.rt0) return rc;
    return launch_cmvnw_big(ctx, lmfe_biased, ms_biased, d_out, b->d_nframes, b->d_frame_offsets,
                            b->d_big_tiles + c.rt0, c.rt1 - c.rt0, b->big_rmax, M, M, win, pl->p.cmvn_variance);
  }
  const int cmk = options().cmvnw_kernel;
  if (b->d_res_tiles != nullptr && (cmk == 0 || cmk == 4) && (((size_t)d_out) & 15) == 0) {
    // utterances that fit in shared memory: both stages in one resident pass (cmvn_resident_kernels.cu);
    // the longer ones: the two-pass large-tile kernels (default), or -- option 4, measured slower --
    // a cluster of two CTAs up to twice that length and the two-pass kernels beyond
    const bool pairs = cmk == 4 && b->resident_pairs;
    rc = launch_cmvnw_resident(ctx, lmfe_biased, d_out, b->d_res_tiles + 2 * c.rs0, c.rs1 - c.rs0, 1);
    if (!rc && pairs) rc = launch_cmvnw_resident(ctx, lmfe_biased, d_out, b->d_res_tiles + 2 * c.r20, c.r21 - c.r20, 2);
    const int64_t t0 = pairs ? c.vl0 : c.lt0, t1 = pairs ? c.vl1 : c.lt1;
    if (rc || t1 == t0) return rc;
    return launch_cmvnw_big(ctx, lmfe_biased, ms_biased, d_out, b->d_nframes, b->d_frame_offsets,
                            b->d_big_tiles + t0, t1 - t0, b->big_rmax, M, M, win, pl->p.cmvn_variance);
  }
  if (b->big_rmax > 0 && options().cmvnw_kernel != 1 && (((size_t)d_out) & 15) == 0)
    return launch_cmvnw_big(ctx, lmfe_biased, ms_biased, d_out, b->d_nframes, b->d_frame_offsets,
                            b->d_big_tiles + c.big0, c.big1 - c.big0, b->big_rmax, M, M, win,
                            pl->p.cmvn_variance);
  return launch_cmvnw_f32(ctx, lmfe_biased, ms_biased, d_out, b->d_nframes, b->d_frame_offsets,
                          b->d_cm_tiles + c.cm0, c.cm1 - c.cm0, b->cm_tile_rows, M, M, win,
                          pl->p.cmvn_variance);
}

static int lmfe_cmvnw_prepare(spb_context *ctx, const spb_batch *b, int64_t max_rows, const float *d_out, float **d_lmfe,
                              float **d_ms) {
  const spb_plan *pl = b->plan;
  const int M = pl->p.num_filters, win = pl->p.cmvn_window;
  SPB_REQUIRE(ctx, win > 0 && (win & 1), "lmfe_cmvnw: plan has no (odd) cmvn_window");
  SPB_REQUIRE(ctx, M >= 1, "lmfe_cmvnw: plan has no filterbank");
  if (!cmvnw_supported(M, win)) return fail(ctx, SPB_ERR_UNSUPPORTED, "cmvnw: window x columns too large for shared memory");
  const size_t chunk_bytes = (size_t)std::max<int64_t>(max_rows, 1) * M * 4;
  *d_lmfe = (float *)scratch(ctx, 2, chunk_bytes);
  // the mean-subtracted intermediate exists only for the two-pass kernels
  const bool need_ms = pl->p.cmvn_variance && (!cmvnw_single_pass(b, d_out) || b->has_short_utts);
  *d_ms = need_ms ? (float *)scratch(ctx, 3, chunk_bytes) : nullptr;
  if (!*d_lmfe || (need_ms && !*d_ms)) return SPB_ERR_CUDA;
  return SPB_OK;
}

static int run_lmfe_cmvnw(spb_context *ctx, const spb_batch *b, const float *d_samples,
                          int64_t arena_len, float *d_out) {
  float *d_lmfe, *d_ms;
  int rc = lmfe_cmvnw_prepare(ctx, b, b->max_group_rows, d_out, &d_lmfe, &d_ms);
  if (rc) return rc;
  for (const BatchChunk &c : b->groups) {
    rc = lmfe_cmvnw_chunk(ctx, b, c, d_samples, arena_len, d_lmfe, d_ms, d_out);
    if (rc) return rc;
  }
  return SPB_OK;
}

// mfcc / mfcc+delta through the tensor-core log-mel kernel: lmfe rows of one run of utterances
// into scratch, then energy + DCT + deltas (mfcc_post_kernel).  Usable when the plan and the batch
// qualify for the tensor-core kernel.
static bool mfcc_hybrid_ok(const spb_batch *b, int mode) {
  return (mode == FEAT_MFCC || mode == FEAT_MFCC_DELTA) && b->d_tc_groups && b->plan->tc_ok && b->plan->pre_cof == 0.0 &&
         options().feature_kernel == 0 && b->plan->p.num_cepstral >= 1;
}

static int mfcc_hybrid_chunk(spb_context *ctx, const spb_batch *b, const BatchChunk &c, const float *d_samples,
                             int64_t arena_len, int mode, float *d_lmfe, float *d_out) {
  const int M = b->plan->p.num_filters;
  FeatArgs a;
  float *lmfe_biased = d_lmfe - (ptrdiff_t)c.row0 * M;  // rows are addressed by their global index
  fill_feat_args(a, b, d_samples, arena_len, FEAT_LMFE, lmfe_biased, nullptr);
  a.tiles = b->d_tiles + c.tile0;
  a.ftiles = b->d_ftiles + c.tile0;
  a.num_tiles = c.tile1 - c.tile0;
  a.tc_groups = b->d_tc_groups + c.tc0;
  a.num_tc_groups = c.tc1 - c.tc0;
  // per-hop sums for the frame energy come out of the tensor-core kernel's converters (scratch slot 3
  // is the cmvnw intermediate of the other batch call: never live at the same time)
  a.tc_hopsum = (float4 *)scratch(ctx, 3, (size_t)std::max<int64_t>(a.num_tc_groups, 1) * 32 * sizeof(float4));
  if (!a.tc_hopsum) return SPB_ERR_CUDA;
  int rc = run_features(ctx, a);
  if (rc) return rc;
  a.mode = mode;
  a.out = d_out;
  return launch_mfcc_post(ctx, a, lmfe_biased);
}

static int run_simple_mode(spb_context *ctx, const spb_batch *b, const float *d_samples,
                           int64_t arena_len, int mode, float *d_out, float *d_energy) {
  if (d_energy == nullptr && mfcc_hybrid_ok(b, mode)) {
    const size_t bytes = (size_t)std::max<int64_t>(b->max_group_rows, 1) * b->plan->p.num_filters * 4;
    float *d_lmfe = (float *)scratch(ctx, 2, bytes);
    if (!d_lmfe) return SPB_ERR_CUDA;
    for (const BatchChunk &c : b->groups) {
      int rc = mfcc_hybrid_chunk(ctx, b, c, d_samples, arena_len, mode, d_lmfe, d_out);
      if (rc) return rc;
    }
    return SPB_OK;
  }
  FeatArgs a;
  fill_feat_args(a, b, d_samples, arena_len, mode, d_out, d_energy);
  if (b->d_tc_groups && !b->chunks.empty()) {
    a.tc_groups = b->d_tc_groups;
    a.num_tc_groups = b->chunks.back().tc1;
  }
  return run_features(ctx, a);
}

// Host signal of any supported dtype -> float32 on device (scratch slot 0).
static int upload_signal(spb_context *ctx, const void *sig, int dtype, int64_t n, float **d_out) {
  size_t esz = (dtype == SPB_F64) ? 8 : (dtype == SPB_I16 ? 2 : 4);
  SPB_REQUIRE(ctx, dtype == SPB_F32 || dtype == SPB_F64 || dtype == SPB_I16 || dtype == SPB_I32,
              "unsupported signal dtype");
  float *d_f32 = (float *)scratch(ctx, 0, (size_t)std::max<int64_t>(n, 1) * 4);
  if (!d_f32) return SPB_ERR_CUDA;
  if (dtype == SPB_F32) {
    SPB_CUDA(ctx, cudaMemcpyAsync(d_f32, sig, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    void *d_raw = scratch(ctx, 5, (size_t)std::max<int64_t>(n, 1) * esz);
    if (!d_raw) return SPB_ERR_CUDA;
    SPB_CUDA(ctx, cudaMemcpyAsync(d_raw, sig, (size_t)n * esz, cudaMemcpyHostToDevice, ctx->stream));
    int rc = launch_convert(ctx, d_raw, dtype, d_f32, SPB_F32, n);
    if (rc) return rc;
  }
  *d_out = d_f32;
  return SPB_OK;
}

// device float32 -> host float64 (widen on the host; D2H carries 4 bytes per value)
static int download_as_f64(spb_context *ctx, const float *d_src, size_t n, double *dst) {
  if (n == 0) return SPB_OK;
  std::vector<float> tmp(n);
  SPB_CUDA(ctx, cudaMemcpyAsync(tmp.data(), d_src, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (size_t i = 0; i < n; ++i) dst[i] = (double)tmp[i];
  return SPB_OK;
}

static int single_signal(spb_context *ctx, const spb_plan *plan, const void *sig, int dtype,
                         int64_t n, int mode, int width, double *out, double *energy) {
  SPB_REQUIRE(ctx, ctx && plan, "feature call: NULL context or plan");
  SPB_REQUIRE(ctx, plan->ctx == ctx, "feature call: plan belongs to another context");
  SPB_REQUIRE(ctx, n >= 0, "feature call: negative length");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  int64_t F;
  if (!frame_count(n, plan->p.frame_len, plan->p.frame_stride, 0, &F))
    return fail(ctx, SPB_ERR_NEGATIVE_DIM, "negative dimensions are not allowed");
  if (F == 0) return SPB_OK;
  SPB_REQUIRE(ctx, sig && out, "feature call: NULL buffer");
  float *d_sig = nullptr;
  int rc = upload_signal(ctx, sig, dtype, n, &d_sig);
  if (rc) return rc;
  const int64_t start = 0;
  spb_batch *b = nullptr;
  rc = spb_batch_create(ctx, plan, &start, &n, 1, &b);
  if (rc) return rc;
  float *d_out = (float *)scratch(ctx, 1, (size_t)F * width * 4);
  float *d_en = energy ? (float *)scratch(ctx, 4, (size_t)F * 4) : nullptr;
  if (!d_out || (energy && !d_en)) {
    spb_batch_destroy(b);
    return SPB_ERR_CUDA;
  }
  rc = run_simple_mode(ctx, b, d_sig, n, mode, d_out, d_en);
  if (!rc) rc = download_as_f64(ctx, d_out, (size_t)F * width, out);
  if (!rc && energy) rc = download_as_f64(ctx, d_en, (size_t)F, energy);
  cudaStreamSynchronize(ctx->stream);
  spb_batch_destroy(b);
  return rc;
}

}  // namespace spb

extern "C" {

int spb_set_option(const char *name, double value) {
  if (!name) return fail(nullptr, SPB_ERR_INVALID, "spb_set_option: NULL name");
  Options &o = options();
  const std::string n(name);
  const int iv = (int)value;
  if (n == "feature_kernel" && iv >= 0 && iv <= 2) o.feature_kernel = iv;
  else if (n == "cmvnw_kernel" && iv >= 0 && iv <= 4) o.cmvnw_kernel = iv;
  else if (n == "tc_version" && (iv == 1 || iv == 2)) o.tc_version = iv;
  else if (n == "tc_epilogue" && iv >= -1 && iv <= 2) o.tc_epilogue = iv;
  else if (n == "redo" && (iv == 0 || iv == 1)) o.redo = iv;
  else if (n == "gemm_multicast" && iv >= 0 && iv <= 1) o.gemm_multicast = iv;
  else if (n == "tau_tc" && value >= 0.0) o.tau_tc = value;
  else if (n == "tau_fft" && value >= 0.0) o.tau_fft = value;
  else return fail(nullptr, SPB_ERR_INVALID, "spb_set_option: unknown option or value out of range: " + n);
  return SPB_OK;
}

int spb_get_option(const char *name, double *value) {
  if (!name || !value) return fail(nullptr, SPB_ERR_INVALID, "spb_get_option: NULL argument");
  const Options &o = options();
  const std::string n(name);
  if (n == "feature_kernel") *value = o.feature_kernel;
  else if (n == "cmvnw_kernel") *value = o.cmvnw_kernel;
  else if (n == "tc_version") *value = o.tc_version;
  else if (n == "tc_epilogue") *value = o.tc_epilogue;
  else if (n == "redo") *value = o.redo;
  else if (n == "gemm_multicast") *value = o.gemm_multicast;
  else if (n == "tau_tc") *value = o.tau_tc;
  else if (n == "tau_fft") *value = o.tau_fft;
  else return fail(nullptr, SPB_ERR_INVALID, "spb_get_option: unknown option: " + n);
  return SPB_OK;
}

int spb_power_spectrum(spb_context *ctx, const double *frames, int64_t F, int32_t L, int32_t nfft,
                       int32_t mode, double *out) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_power_spectrum: NULL context");
  SPB_REQUIRE(ctx, F >= 0 && L >= 1, "spb_power_spectrum: bad shape");
  SPB_REQUIRE(ctx, mode >= 0 && mode <= 3, "spb_power_spectrum: mode must be 0..3");
  SPB_REQUIRE(ctx, nfft >= 1, "spb_power_spectrum: fft_points must be positive");
  if (F == 0) return SPB_OK;
  SPB_REQUIRE(ctx, frames && out, "spb_power_spectrum: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!(is_pow2(nfft) && nfft >= 32 && nfft <= 4096)) {
    // any other size (numpy's rfft(n=) takes them all): direct DFT in float64 on the device
    const int K = nfft / 2 + 1;
    double *d_fr = (double *)scratch(ctx, 5, (size_t)F * L * 8);
    double *d_o = (double *)scratch(ctx, 4, (size_t)F * K * 8 + 256);
    if (!d_fr || !d_o) return SPB_ERR_CUDA;
    SPB_CUDA(ctx, cudaMemcpyAsync(d_fr, frames, (size_t)F * L * 8, cudaMemcpyHostToDevice, ctx->stream));
    int rc = launch_spectrum_dft(ctx, d_fr, F, L, nfft, mode, d_o, (double *)((char *)d_o + (size_t)F * K * 8));
    if (rc) return rc;
    SPB_CUDA(ctx, cudaMemcpyAsync(out, d_o, (size_t)F * K * 8, cudaMemcpyDeviceToHost, ctx->stream));
    SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SPB_OK;
  }
  // (F, L) row-major frames == one signal of F*L samples framed with length L, stride L
  spb_feature_params q{};
  q.frame_len = L;
  q.frame_stride = L;
  q.nfft = nfft;
  spb_plan *pl = nullptr;
  int rc = spb_plan_create(ctx, &q, nullptr, &pl);
  if (rc) return rc;
  const int64_t n = F * L;
  float *d_sig = nullptr;
  rc = upload_signal(ctx, frames, SPB_F64, n, &d_sig);
  spb_batch *b = nullptr;
  if (!rc) {
    // F frames of length L at stride L need N = F*L + L samples under the mfe frame
    // count floor((N-L)/S); describe the signal as one sample-row longer than it is and
    // let the (never read) tail stay virtual: the last frame ends at sample F*L - 1.
    const int64_t start = 0, len = n + L;
    rc = spb_batch_create(ctx, pl, &start, &len, 1, &b);
  }
  const int K = nfft / 2 + 1;
  float *d_out = rc ? nullptr : (float *)scratch(ctx, 1, (size_t)F * K * 4);
  if (!rc && !d_out) rc = SPB_ERR_CUDA;
  if (!rc) {
    FeatArgs a;
    fill_feat_args(a, b, d_sig, n, mode == 1 ? FEAT_MAG : FEAT_POWER, d_out, nullptr);
    rc = run_features(ctx, a);
  }
  if (!rc && mode >= 2) {  // log power spectrum: float64 epilogue on the device
    double *d_log = (double *)scratch(ctx, 4, (size_t)F * K * 8 + 256);
    if (!d_log) rc = SPB_ERR_CUDA;
    unsigned int *d_max = d_log ? (unsigned int *)((char *)d_log + (size_t)F * K * 8) : nullptr;
    if (!rc) rc = launch_log_power(ctx, d_out, (long long)F * K, mode == 3, d_max, d_log);
    if (!rc) {
      cudaError_t e = cudaMemcpyAsync(out, d_log, (size_t)F * K * 8, cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess) rc = cuda_fail(ctx, e, "log power D2H");
    }
  } else if (!rc) {
    rc = download_as_f64(ctx, d_out, (size_t)F * K, out);
  }
  cudaStreamSynchronize(ctx->stream);
  spb_batch_destroy(b);
  spb_plan_destroy(pl);
  return rc;
}

int spb_mfe(spb_context *ctx, const spb_plan *plan, const void *sig, int32_t dtype, int64_t n,
            double *feat, double *energy) {
  if (!plan) return fail(ctx, SPB_ERR_INVALID, "spb_mfe: NULL plan");
  return single_signal(ctx, plan, sig, dtype, n, FEAT_MFE, plan->p.num_filters, feat, energy);
}

int spb_lmfe(spb_context *ctx, const spb_plan *plan, const void *sig, int32_t dtype, int64_t n,
             double *feat) {
  if (!plan) return fail(ctx, SPB_ERR_INVALID, "spb_lmfe: NULL plan");
  return single_signal(ctx, plan, sig, dtype, n, FEAT_LMFE, plan->p.num_filters, feat, nullptr);
}

int spb_mfcc(spb_context *ctx, const spb_plan *plan, const void *sig, int32_t dtype, int64_t n,
             double *ceps) {
  if (!plan) return fail(ctx, SPB_ERR_INVALID, "spb_mfcc: NULL plan");
  SPB_REQUIRE(ctx, plan->p.num_cepstral >= 1, "spb_mfcc: plan has num_cepstral == 0");
  return single_signal(ctx, plan, sig, dtype, n, FEAT_MFCC, plan->p.num_cepstral, ceps, nullptr);
}

// ------------------------------------------------------------------ float64 helpers
static int f64_roundtrip(spb_context *ctx, const double *in, size_t n_in, double **d_in, size_t n_out,
                         double **d_out) {
  *d_in = (double *)scratch(ctx, 0, std::max(n_in, (size_t)1) * 8);
  *d_out = (double *)scratch(ctx, 1, std::max(n_out, (size_t)1) * 8);
  if (!*d_in || !*d_out) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemcpyAsync(*d_in, in, n_in * 8, cudaMemcpyHostToDevice, ctx->stream));
  return SPB_OK;
}

int spb_delta_cube(spb_context *ctx, const double *feat, int64_t F, int32_t cols, double *out) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_delta_cube: NULL context");
  SPB_REQUIRE(ctx, F >= 0 && cols >= 1, "spb_delta_cube: bad shape");
  if (F == 0) return SPB_OK;
  SPB_REQUIRE(ctx, feat && out, "spb_delta_cube: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  double *d_in, *d_out;
  const size_t n = (size_t)F * cols;
  int rc = f64_roundtrip(ctx, feat, n, &d_in, n * 3, &d_out);
  if (rc) return rc;
  rc = launch_delta_cube_f64(ctx, d_in, F, cols, d_out);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * 3 * 8, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

int spb_derivative(spb_context *ctx, const double *feat, int64_t F, int32_t cols, int32_t dw,
                   double *out) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_derivative: NULL context");
  SPB_REQUIRE(ctx, F >= 0 && cols >= 1 && dw >= 1, "spb_derivative: bad shape or DeltaWindows");
  if (F == 0) return SPB_OK;
  SPB_REQUIRE(ctx, feat && out, "spb_derivative: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  double *d_in, *d_out;
  const size_t n = (size_t)F * cols;
  int rc = f64_roundtrip(ctx, feat, n, &d_in, n, &d_out);
  if (rc) return rc;
  rc = launch_derivative_f64(ctx, d_in, F, cols, dw, d_out);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

int spb_cmvn(spb_context *ctx, const double *vec, int64_t rows, int32_t cols, int32_t varnorm,
             double *out) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_cmvn: NULL context");
  SPB_REQUIRE(ctx, rows >= 0 && cols >= 0, "spb_cmvn: bad shape");
  if (rows == 0 || cols == 0) return SPB_OK;
  SPB_REQUIRE(ctx, vec && out, "spb_cmvn: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  double *d_in, *d_out;
  const size_t n = (size_t)rows * cols;
  int rc = f64_roundtrip(ctx, vec, n, &d_in, n, &d_out);
  if (rc) return rc;
  rc = launch_cmvn_f64(ctx, d_in, rows, cols, varnorm, d_out);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

int spb_cmvnw(spb_context *ctx, const double *vec, int64_t rows, int32_t cols, int32_t win,
              int32_t varnorm, float *out) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_cmvnw: NULL context");
  SPB_REQUIRE(ctx, win > 0 && (win & 1), "Windows size must be odd!");
  SPB_REQUIRE(ctx, rows >= 0 && cols >= 0, "spb_cmvnw: bad shape");
  SPB_REQUIRE(ctx, rows < ((int64_t)1 << 30), "spb_cmvnw: too many rows");
  if (rows == 0 || cols == 0) return SPB_OK;
  SPB_REQUIRE(ctx, vec && out, "spb_cmvnw: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  // column chunk that fits shared memory
  int cc = cols;
  while (cc > 1 && !cmvnw_supported(cc, win)) cc = (cc + 1) / 2;
  if (!cmvnw_supported(cc, win)) return fail(ctx, SPB_ERR_UNSUPPORTED, "cmvnw: win_size too large for shared memory");
  const size_t n = (size_t)rows * cols;
  double *d_f64 = (double *)scratch(ctx, 5, n * 8);
  float *d_out = (float *)scratch(ctx, 1, n * 4);
  float *d_ms = (float *)scratch(ctx, 3, n * 4);
  if (!d_f64 || !d_out || !d_ms) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemcpyAsync(d_f64, vec, n * 8, cudaMemcpyHostToDevice, ctx->stream));
  // stage 1 in float64 from the float64 input (processing.py:302-307): a large common offset is
  // subtracted before anything is narrowed; stage 2 on the float32 rows like the reference
  int rc = launch_cmvnw_stage1_f64(ctx, d_f64, rows, cols, win, varnorm ? d_ms : d_out);
  if (rc) return rc;
  if (!varnorm) {
    SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SPB_OK;
  }
  const int32_t nfr = (int32_t)rows;
  const int64_t offs[2] = {0, rows};
  const int big_rmax = options().cmvnw_kernel == 1 ? 0 : cmvnw_big_rmax(cols, win);
  if (big_rmax > 0) {
    // large-tile kernels over all columns at once
    std::vector<int4> tiles;
    append_big_tiles(tiles, 0, rows, big_rmax);
    unsigned char *meta = (unsigned char *)scratch(ctx, 4, 64 + tiles.size() * sizeof(int4));
    if (!meta) return SPB_ERR_CUDA;
    SPB_CUDA(ctx, cudaMemcpyAsync(meta, &nfr, 4, cudaMemcpyHostToDevice, ctx->stream));
    SPB_CUDA(ctx, cudaMemcpyAsync(meta + 16, offs, 16, cudaMemcpyHostToDevice, ctx->stream));
    SPB_CUDA(ctx, cudaMemcpyAsync(meta + 64, tiles.data(), tiles.size() * sizeof(int4), cudaMemcpyHostToDevice, ctx->stream));
    rc = launch_cmvnw_big(ctx, nullptr, d_ms, d_out, (const int32_t *)meta, (const int64_t *)(meta + 16),
                          (const int4 *)(meta + 64), (int64_t)tiles.size(), big_rmax, cols, cols, win, 1, 2);
    if (rc) return rc;
    SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SPB_OK;
  }
  // one "utterance" of `rows` rows
  const int R = cmvnw_tile_rows(cc, win);
  const int64_t ntiles = (rows + R - 1) / R;
  const int64_t toff[2] = {0, ntiles};
  int32_t *d_nfr = (int32_t *)scratch(ctx, 4, 64 + (size_t)ntiles * sizeof(int2));
  if (!d_nfr) return SPB_ERR_CUDA;
  int64_t *d_offs = (int64_t *)((char *)d_nfr + 16);
  int64_t *d_toff = (int64_t *)((char *)d_nfr + 32);
  int2 *d_tiles = (int2 *)((char *)d_nfr + 64);
  SPB_CUDA(ctx, cudaMemcpyAsync(d_nfr, &nfr, 4, cudaMemcpyHostToDevice, ctx->stream));
  SPB_CUDA(ctx, cudaMemcpyAsync(d_offs, offs, 16, cudaMemcpyHostToDevice, ctx->stream));
  SPB_CUDA(ctx, cudaMemcpyAsync(d_toff, toff, 16, cudaMemcpyHostToDevice, ctx->stream));
  rc = launch_build_tiles(ctx, d_nfr, d_toff, 1, R, d_tiles);
  if (rc) return rc;
  for (int c0 = 0; c0 < cols; c0 += cc) {
    const int w = std::min(cc, cols - c0);
    rc = launch_cmvnw_f32(ctx, nullptr, d_ms + c0, d_out + c0, d_nfr, d_offs, d_tiles, ntiles, R,
                          w, cols, win, 1, 2);
    if (rc) return rc;
  }
  SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

// ------------------------------------------------------------------ batch entry points
int spb_lmfe_cmvnw_batch_dev(spb_context *ctx, const spb_batch *b, const float *d_samples,
                             int64_t arena_len, float *d_out) {
  int rc = check_batch(ctx, b, arena_len, d_samples, d_out);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return run_lmfe_cmvnw(ctx, b, d_samples, arena_len, d_out);
}

int spb_lmfe_batch_dev(spb_context *ctx, const spb_batch *b, const float *d_samples,
                       int64_t arena_len, float *d_out) {
  int rc = check_batch(ctx, b, arena_len, d_samples, d_out);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return run_simple_mode(ctx, b, d_samples, arena_len, FEAT_LMFE, d_out, nullptr);
}

int spb_mfcc_delta_batch_dev(spb_context *ctx, const spb_batch *b, const float *d_samples,
                             int64_t arena_len, float *d_out) {
  int rc = check_batch(ctx, b, arena_len, d_samples, d_out);
  if (rc) return rc;
  SPB_REQUIRE(ctx, b->plan->p.num_cepstral >= 1, "mfcc_delta: plan has num_cepstral == 0");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return run_simple_mode(ctx, b, d_samples, arena_len, FEAT_MFCC_DELTA, d_out, nullptr);
}

// Host-buffer batch calls, pipelined over the batch's chunks (runs of whole utterances):
//   copy_stream : H2D of chunk k+1's sample span (raw dtype)
//   stream      : [convert to float32] + kernels of chunk k
//   out_stream  : D2H of chunk k-1's feature rows
// so a call costs about max(H2D, compute, D2H) instead of their sum (PCIe is full duplex).
// `samples` may be float32, int16 PCM, int32 or float64; x = float(sample) * scale on the
// device: scale 1 is signal.astype(float) (feature.py:182), int16 with scale 1/32768 is the
// float32 array librosa.load (utils.py:170-173 load_wav) decodes from a PCM16 file.
static int host_batch_run(spb_context *ctx, const spb_batch *b, const void *samples, int dtype,
                          double scale_d, int64_t arena_len, float *out, int which);

// On an error in the middle of the pipeline, copies into the caller's `out` (and from `samples`) may still
// be in flight on the copy streams: drain all three before the call returns, so that the caller's buffers
// and the context's scratch / events are quiescent whatever the outcome.
static int host_batch(spb_context *ctx, const spb_batch *b, const void *samples, int dtype,
                      double scale_d, int64_t arena_len, float *out, int which) {
  const int rc = host_batch_run(ctx, b, samples, dtype, scale_d, arena_len, out, which);
  if (rc != SPB_OK && ctx != nullptr && ctx->stream != nullptr) {
    const std::string msg = ctx->last_error;  // keep the first error's message
    cudaStreamSynchronize(ctx->copy_stream);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamSynchronize(ctx->out_stream);
    cudaGetLastError();
    ctx->last_error = msg;
  }
  return rc;
}

static int host_batch_run(spb_context *ctx, const spb_batch *b, const void *samples, int dtype,
                          double scale_d, int64_t arena_len, float *out, int which) {
  const float scale = (float)scale_d;
  SPB_REQUIRE(ctx, scale_d == scale_d && scale != 0.f, "batch call: sample_scale must be a non-zero number");
  int rc = check_batch(ctx, b, arena_len, samples, out);
  if (rc) return rc;
  SPB_REQUIRE(ctx, dtype == SPB_F32 || dtype == SPB_F64 || dtype == SPB_I16 || dtype == SPB_I32,
              "batch call: unsupported sample dtype");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (b->total_frames == 0) return SPB_OK;
  const size_t esz = (dtype == SPB_F64) ? 8 : (dtype == SPB_I16 ? 2 : 4);
  const int width = (which == 0) ? b->plan->p.num_filters : b->plan->p.num_cepstral * 3;
  float *d_samples = (float *)scratch(ctx, 0, (size_t)arena_len * 4);
  float *d_out = (float *)scratch(ctx, 1, (size_t)b->total_frames * width * 4);
  char *d_raw = (dtype == SPB_F32) ? (char *)d_samples : (char *)scratch(ctx, 5, (size_t)arena_len * esz);
  if (!d_samples || !d_out || !d_raw) return SPB_ERR_CUDA;
  float *d_lmfe = nullptr, *d_ms = nullptr;
  if (which == 0 && (rc = lmfe_cmvnw_prepare(ctx, b, b->max_chunk_rows, d_out, &d_lmfe, &d_ms))) return rc;
  const bool hybrid = which != 0 && mfcc_hybrid_ok(b, FEAT_MFCC_DELTA);
  if (hybrid) {
    d_lmfe = (float *)scratch(ctx, 2, (size_t)std::max<int64_t>(b->max_chunk_rows, 1) * b->plan->p.num_filters * 4);
    if (!d_lmfe) return SPB_ERR_CUDA;
  }
  const size_t nch = b->chunks.size();
  const bool ingest = dtype != SPB_F32 || scale != 1.0f;  // float32 with scale 1 is used as uploaded
  // piecewise upload only when the chunks' sample spans are disjoint (utterances packed in order): a span that
  // overlaps its neighbour's would be rewritten by the next chunk's H2D while this chunk's kernels read it
  const bool piecewise = b->chunk_spans_disjoint;
  while (ctx->pipe_events.size() < 2 * nch + 1) {
    cudaEvent_t e;
    SPB_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ctx->pipe_events.push_back(e);
  }
  // the copy streams must not overtake earlier work on the compute stream that uses the scratch
  cudaEvent_t ev_start = ctx->pipe_events[2 * nch];
  SPB_CUDA(ctx, cudaEventRecord(ev_start, ctx->stream));
  SPB_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ev_start, 0));
  SPB_CUDA(ctx, cudaStreamWaitEvent(ctx->out_stream, ev_start, 0));
  if (!piecewise) {
    SPB_CUDA(ctx, cudaMemcpyAsync(d_raw, samples, (size_t)arena_len * esz, cudaMemcpyHostToDevice, ctx->copy_stream));
    SPB_CUDA(ctx, cudaEventRecord(ctx->pipe_events[0], ctx->copy_stream));
    SPB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipe_events[0], 0));
    if (ingest && (rc = launch_ingest(ctx, d_raw, dtype, d_samples, arena_len, scale))) return rc;
  }
  for (size_t k = 0; k < nch; ++k) {
    const BatchChunk &c = b->chunks[k];
    cudaEvent_t ev_in = ctx->pipe_events[2 * k], ev_done = ctx->pipe_events[2 * k + 1];
    if (piecewise) {
      const int64_t n = c.s_hi - c.s_lo;
      SPB_CUDA(ctx, cudaMemcpyAsync(d_raw + (size_t)c.s_lo * esz, (const char *)samples + (size_t)c.s_lo * esz,
                                    (size_t)n * esz, cudaMemcpyHostToDevice, ctx->copy_stream));
      SPB_CUDA(ctx, cudaEventRecord(ev_in, ctx->copy_stream));
      SPB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ev_in, 0));
      if (ingest && (rc = launch_ingest(ctx, d_raw + (size_t)c.s_lo * esz, dtype, d_samples + c.s_lo, n, scale)))
        return rc;
    }
    if (which == 0) {
      rc = lmfe_cmvnw_chunk(ctx, b, c, d_samples, arena_len, d_lmfe, d_ms, d_out);
    } else if (hybrid) {
      rc = mfcc_hybrid_chunk(ctx, b, c, d_samples, arena_len, FEAT_MFCC_DELTA, d_lmfe, d_out);
    } else {
      FeatArgs a;
      fill_feat_args(a, b, d_samples, arena_len, FEAT_MFCC_DELTA, d_out, nullptr);
      a.tiles = b->d_tiles + c.tile0;
      a.ftiles = b->d_ftiles + c.tile0;
      a.num_tiles = c.tile1 - c.tile0;
      rc = run_features(ctx, a);
    }
    if (rc) return rc;
    SPB_CUDA(ctx, cudaEventRecord(ev_done, ctx->stream));
    SPB_CUDA(ctx, cudaStreamWaitEvent(ctx->out_stream, ev_done, 0));
    SPB_CUDA(ctx, cudaMemcpyAsync(out + (size_t)c.row0 * width, d_out + (size_t)c.row0 * width,
                                  (size_t)(c.row1 - c.row0) * width * 4, cudaMemcpyDeviceToHost, ctx->out_stream));
  }
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->out_stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

int spb_lmfe_cmvnw_batch(spb_context *ctx, const spb_batch *b, const float *samples,
                         int64_t arena_len, float *out) {
  return host_batch(ctx, b, samples, SPB_F32, 1.0, arena_len, out, 0);
}

int spb_lmfe_cmvnw_batch_dtype(spb_context *ctx, const spb_batch *b, const void *samples, int dtype,
                               double sample_scale, int64_t arena_len, float *out) {
  return host_batch(ctx, b, samples, dtype, sample_scale, arena_len, out, 0);
}

int spb_mfcc_delta_batch(spb_context *ctx, const spb_batch *b, const float *samples,
                         int64_t arena_len, float *out) {
  if (b && b->plan->p.num_cepstral < 1) return fail(ctx, SPB_ERR_INVALID, "mfcc_delta: plan has num_cepstral == 0");
  return host_batch(ctx, b, samples, SPB_F32, 1.0, arena_len, out, 1);
}

int spb_mfcc_delta_batch_dtype(spb_context *ctx, const spb_batch *b, const void *samples, int dtype,
                               double sample_scale, int64_t arena_len, float *out) {
  if (b && b->plan->p.num_cepstral < 1) return fail(ctx, SPB_ERR_INVALID, "mfcc_delta: plan has num_cepstral == 0");
  return host_batch(ctx, b, samples, dtype, sample_scale, arena_len, out, 1);
}

int spb_preemphasis(spb_context *ctx, const void *sig, int32_t dtype, int64_t n, int64_t shift, double cof,
                    void *out) {
  SPB_REQUIRE(ctx, ctx != nullptr, "spb_preemphasis: NULL context");
  SPB_REQUIRE(ctx, n >= 0, "spb_preemphasis: negative length");
  SPB_REQUIRE(ctx, dtype == SPB_F32 || dtype == SPB_F64 || dtype == SPB_I16 || dtype == SPB_I32,
              "spb_preemphasis: unsupported dtype");
  if (n == 0) return SPB_OK;
  SPB_REQUIRE(ctx, sig && out, "spb_preemphasis: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t esz_in = (dtype == SPB_F64) ? 8 : (dtype == SPB_I16 ? 2 : 4);
  const size_t esz_out = (dtype == SPB_F32) ? 4 : 8;
  void *d_in = scratch(ctx, 0, (size_t)n * esz_in);
  void *d_out = scratch(ctx, 1, (size_t)n * esz_out);
  if (!d_in || !d_out) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemcpyAsync(d_in, sig, (size_t)n * esz_in, cudaMemcpyHostToDevice, ctx->stream));
  int rc = launch_preemphasis(ctx, d_in, dtype, n, shift, cof, d_out);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, (size_t)n * esz_out, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

// ------------------------------------------------------------------ 8f-1 feature cubes
static int run_feature_cubes(spb_context *ctx, const spb_batch *b, const float *d_samples, int64_t arena_len,
                             int derivative, int normalize, const int32_t *crop_starts, int num_crops,
                             int crop_frames, float *d_out) {
  const int M = b->plan->p.num_filters;
  SPB_REQUIRE(ctx, M >= 1, "feature cubes: plan has no filterbank");
  SPB_REQUIRE(ctx, num_crops >= 1 && crop_frames >= 1 && crop_starts, "feature cubes: bad crop arguments");
  for (int64_t u = 0; u < b->U; ++u) {
    const int64_t F = b->h_frame_offsets[u + 1] - b->h_frame_offsets[u];
    for (int k = 0; k < num_crops; ++k) {
      const int64_t st = crop_starts[u * num_crops + k];
      // np.random.randint(F - frames) yields [0, F - frames) and needs F > frames (utils.py:373)
      SPB_REQUIRE(ctx, st >= 0 && st + crop_frames <= F, "feature cubes: crop outside its utterance");
    }
  }
  const int channels = derivative ? 3 : 1;
  float *d_lmfe = (float *)scratch(ctx, 2, (size_t)std::max<int64_t>(b->total_frames, 1) * M * 4);
  int32_t *d_crops = (int32_t *)scratch(ctx, 4, (size_t)b->U * num_crops * 4);
  double2 *d_stats = normalize ? (double2 *)scratch(ctx, 3, (size_t)b->U * channels * M * sizeof(double2)) : nullptr;
  if (!d_lmfe || !d_crops || (normalize && !d_stats)) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemcpyAsync(d_crops, crop_starts, (size_t)b->U * num_crops * 4, cudaMemcpyHostToDevice, ctx->stream));
  int rc = run_simple_mode(ctx, b, d_samples, arena_len, FEAT_LMFE, d_lmfe, nullptr);
  if (rc) return rc;
  rc = launch_feature_cubes(ctx, d_lmfe, b->d_frame_offsets, b->U, M, channels, normalize, d_crops, num_crops,
                            crop_frames, d_stats, d_out);
  if (rc) return rc;
  // crop_starts is caller memory (pageable): finish its upload before returning
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

int spb_feature_cubes_batch_dev(spb_context *ctx, const spb_batch *b, const float *d_samples, int64_t arena_len,
                                int32_t derivative, int32_t normalize, const int32_t *crop_starts,
                                int32_t num_crops, int32_t crop_frames, float *d_out) {
  int rc = check_batch(ctx, b, arena_len, d_samples, d_out);
  if (rc) return rc;
  if (b->U == 0) return SPB_OK;
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return run_feature_cubes(ctx, b, d_samples, arena_len, derivative, normalize, crop_starts, num_crops,
                           crop_frames, d_out);
}

int spb_feature_cubes_batch(spb_context *ctx, const spb_batch *b, const void *samples, int32_t dtype,
                            double sample_scale, int64_t arena_len, int32_t derivative, int32_t normalize,
                            const int32_t *crop_starts, int32_t num_crops, int32_t crop_frames, float *out) {
  int rc = check_batch(ctx, b, arena_len, samples, out);
  if (rc) return rc;
  SPB_REQUIRE(ctx, dtype == SPB_F32 || dtype == SPB_F64 || dtype == SPB_I16 || dtype == SPB_I32,
              "feature cubes: unsupported sample dtype");
  if (b->U == 0) return SPB_OK;
  SPB_REQUIRE(ctx, num_crops >= 1 && crop_frames >= 1, "feature cubes: bad crop arguments");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t esz = (dtype == SPB_F64) ? 8 : (dtype == SPB_I16 ? 2 : 4);
  const float scale = (float)sample_scale;
  const int M = b->plan->p.num_filters, channels = derivative ? 3 : 1;
  const size_t out_elems = (size_t)b->U * channels * num_crops * crop_frames * M;
  float *d_samples = (float *)scratch(ctx, 0, (size_t)std::max<int64_t>(arena_len, 1) * 4);
  float *d_out = (float *)scratch(ctx, 1, out_elems * 4);
  if (!d_samples || !d_out) return SPB_ERR_CUDA;
  if (dtype == SPB_F32 && scale == 1.0f) {
    SPB_CUDA(ctx, cudaMemcpyAsync(d_samples, samples, (size_t)arena_len * 4, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    void *d_raw = scratch(ctx, 5, (size_t)std::max<int64_t>(arena_len, 1) * esz);
    if (!d_raw) return SPB_ERR_CUDA;
    SPB_CUDA(ctx, cudaMemcpyAsync(d_raw, samples, (size_t)arena_len * esz, cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = launch_ingest(ctx, d_raw, dtype, d_samples, arena_len, scale))) return rc;
  }
  rc = run_feature_cubes(ctx, b, d_samples, arena_len, derivative, normalize, crop_starts, num_crops, crop_frames,
                         d_out);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaMemcpyAsync(out, d_out, out_elems * 4, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

}  // extern "C"
