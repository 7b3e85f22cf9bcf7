// Small element-wise kernels: float64 frame gather (stack_frames), the reference's
// "derivative" operator and delta cube, dtype conversion, tile-table construction.
#include <algorithm>

#include "spb_internal.cuh"

namespace spb {

// speech_feature_extraction/speechpy/processing.py:123-138: frames = signal[idx] * window,
// idx[f, j] = j + f*S.  Samples at or beyond N are the zero padding of processing.py:108-110.
__global__ void stack_frames_kernel(const double *__restrict__ sig, long long n, int L, int S,
                                    const double *__restrict__ window, long long total,
                                    double *__restrict__ out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long f = i / L;
    const int j = (int)(i - f * L);
    const long long idx = f * S + j;
    double v = (idx < n) ? sig[idx] : 0.0;
    if (window) v *= window[j];
    out[i] = v;
  }
}

int launch_stack_frames_f64(spb_context *ctx, const double *sig, int64_t n, int L, int S,
                            const double *window, int64_t F, double *out) {
  const long long total = (long long)F * L;
  if (total <= 0) return SPB_OK;
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)ctx->sm_count * 16) blocks = (long long)ctx->sm_count * 16;
  stack_frames_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(sig, n, L, S, window, total, out);
  SPB_LAUNCH_CHECK(ctx, "stack_frames_kernel");
  return SPB_OK;
}

// processing.py:216-236 as written: DIF[f, c] = sum_{r=1..D} r * x[f, min(c + r, cols-1)],
// divided by 2 * sum r^2 (the subtraction on processing.py:232 is a no-op statement).
__device__ __forceinline__ double deriv_at(const double *row, int c, int cols, int dw, double inv) {
  double acc = 0.0;
  for (int r = 1; r <= dw; ++r) acc += (double)r * row[min(c + r, cols - 1)];
  return acc * inv;
}

__global__ void derivative_kernel(const double *__restrict__ in, long long rows, int cols, int dw,
                                  double scale, double *__restrict__ out) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long f = i / cols;
    const int c = (int)(i - f * cols);
    const double *row = in + f * cols;
    double acc = 0.0;
    for (int r = 1; r <= dw; ++r) acc += (double)r * row[min(c + r, cols - 1)];
    out[i] = acc / scale;
  }
}

int launch_derivative_f64(spb_context *ctx, const double *in, int64_t rows, int cols, int dw,
                          double *out) {
  const long long total = (long long)rows * cols;
  if (total <= 0) return SPB_OK;
  double scale = 0.0;
  for (int r = 1; r <= dw; ++r) scale += 2.0 * r * r;
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)ctx->sm_count * 16) blocks = (long long)ctx->sm_count * 16;
  derivative_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(in, rows, cols, dw, scale, out);
  SPB_LAUNCH_CHECK(ctx, "derivative_kernel");
  return SPB_OK;
}

// feature.py:272-278: cube[f, c, :] = [x, D(x), D(D(x))], DeltaWindows = 2 (scale 10).
__global__ void delta_cube_kernel(const double *__restrict__ in, long long rows, int cols,
                                  double *__restrict__ out) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long f = i / cols;
    const int c = (int)(i - f * cols);
    const double *row = in + f * cols;
    auto X = [&](int j) { return row[min(j, cols - 1)]; };
    auto D1 = [&](int j) {
      j = min(j, cols - 1);
      return (X(j + 1) + 2.0 * X(j + 2)) / 10.0;
    };
    out[i * 3 + 0] = row[c];
    out[i * 3 + 1] = D1(c);
    out[i * 3 + 2] = (D1(c + 1) + 2.0 * D1(c + 2)) / 10.0;
  }
}

int launch_delta_cube_f64(spb_context *ctx, const double *in, int64_t rows, int cols,
                          double *out) {
  const long long total = (long long)rows * cols;
  if (total <= 0) return SPB_OK;
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)ctx->sm_count * 16) blocks = (long long)ctx->sm_count * 16;
  delta_cube_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(in, rows, cols, out);
  SPB_LAUNCH_CHECK(ctx, "delta_cube_kernel");
  return SPB_OK;
}

// ------------------------------------------------------------------ conversions
template <typename TI, typename TO>
__global__ void convert_kernel(const TI *__restrict__ in, TO *__restrict__ out, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = (TO)in[i];
}

int launch_convert(spb_context *ctx, const void *in, int dtype_in, void *out, int dtype_out,
                   int64_t n) {
  if (n <= 0) return SPB_OK;
  long long blocks = (n + 255) / 256;
  if (blocks > (long long)ctx->sm_count * 16) blocks = (long long)ctx->sm_count * 16;
  const unsigned g = (unsigned)blocks;
#define SPB_CVT(TI, TO) \
  convert_kernel<TI, TO><<<g, 256, 0, ctx->stream>>>((const TI *)in, (TO *)out, n)
  if (dtype_in == SPB_F64 && dtype_out == SPB_F32) SPB_CVT(double, float);
  else if (dtype_in == SPB_F32 && dtype_out == SPB_F64) SPB_CVT(float, double);
  else if (dtype_in == SPB_I16 && dtype_out == SPB_F32) SPB_CVT(int16_t, float);
  else if (dtype_in == SPB_I32 && dtype_out == SPB_F32) SPB_CVT(int32_t, float);
  else if (dtype_in == SPB_F32 && dtype_out == SPB_F32) SPB_CVT(float, float);
  else return fail(ctx, SPB_ERR_INVALID, "unsupported dtype conversion");
#undef SPB_CVT
  SPB_LAUNCH_CHECK(ctx, "convert_kernel");
  return SPB_OK;
}

// ------------------------------------------------------------------ sample ingest
// out[i] = float(in[i]) * scale: raw PCM / integer / float64 arenas widened on the device.
// scale = 1 is `signal.astype(float)` (feature.py:182); scale = 1/32768 on int16 is what
// librosa.load (utils.py:170-173) produces from a PCM16 file.  Power-of-two scales are exact.
template <typename TI>
__global__ void ingest_kernel(const TI *__restrict__ in, float *__restrict__ out, long long n,
                              float scale) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = (float)in[i] * scale;
}

// int16: 8 samples (16 B) in, two float4 out per thread and step; both pointers 16-byte aligned
__global__ void ingest_i16x8_kernel(const int4 *__restrict__ in, float4 *__restrict__ out,
                                    long long n8, float scale) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    const int4 v = __ldg(in + i);
    float4 a, b;
    a.x = (float)(short)(v.x & 0xffff) * scale;
    a.y = (float)(v.x >> 16) * scale;
    a.z = (float)(short)(v.y & 0xffff) * scale;
    a.w = (float)(v.y >> 16) * scale;
    b.x = (float)(short)(v.z & 0xffff) * scale;
    b.y = (float)(v.z >> 16) * scale;
    b.z = (float)(short)(v.w & 0xffff) * scale;
    b.w = (float)(v.w >> 16) * scale;
    out[2 * i] = a;
    out[2 * i + 1] = b;
  }
}

int launch_ingest(spb_context *ctx, const void *in, int dtype_in, float *out, int64_t n, float scale) {
  if (n <= 0) return SPB_OK;
  const long long cap = (long long)ctx->sm_count * 16;
  if (dtype_in == SPB_I16 && (((size_t)in | (size_t)out) & 15) == 0 && n >= 8) {
    const long long n8 = n / 8;
    const unsigned g = (unsigned)std::min<long long>((n8 + 255) / 256, cap);
    ingest_i16x8_kernel<<<g, 256, 0, ctx->stream>>>((const int4 *)in, (float4 *)out, n8, scale);
    SPB_LAUNCH_CHECK(ctx, "ingest_i16x8_kernel");
    const long long done = n8 * 8;
    if (done == n) return SPB_OK;
    ingest_kernel<int16_t><<<1, 32, 0, ctx->stream>>>((const int16_t *)in + done, out + done, n - done, scale);
    SPB_LAUNCH_CHECK(ctx, "ingest_kernel");
    return SPB_OK;
  }
  const unsigned g = (unsigned)std::min<long long>((n + 255) / 256, cap);
  if (dtype_in == SPB_I16) ingest_kernel<int16_t><<<g, 256, 0, ctx->stream>>>((const int16_t *)in, out, n, scale);
  else if (dtype_in == SPB_I32) ingest_kernel<int32_t><<<g, 256, 0, ctx->stream>>>((const int32_t *)in, out, n, scale);
  else if (dtype_in == SPB_F64) ingest_kernel<double><<<g, 256, 0, ctx->stream>>>((const double *)in, out, n, scale);
  else if (dtype_in == SPB_F32) ingest_kernel<float><<<g, 256, 0, ctx->stream>>>((const float *)in, out, n, scale);
  else return fail(ctx, SPB_ERR_INVALID, "unsupported sample dtype");
  SPB_LAUNCH_CHECK(ctx, "ingest_kernel");
  return SPB_OK;
}

// ------------------------------------------------------------------ preemphasis (8f-2)
// processing.py:45-58: signal - cof * np.roll(signal, shift): rolled[i] = signal[(i - shift) mod n]
// (circular: sample 0 uses the last sample).  Two separately rounded operations like numpy
// (no FMA contraction), in the dtype numpy computes in: float32 for a float32 signal, float64
// for float64 and integer signals.
template <typename TI, typename TO>
__global__ void preemphasis_kernel(const TI *__restrict__ x, long long n, long long shift, TO cof,
                                   TO *__restrict__ out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    long long j = i - shift;
    if (j < 0) j += n;
    const TO a = (TO)x[i], b = (TO)x[j];
    if (sizeof(TO) == 4) out[i] = (TO)__fsub_rn((float)a, __fmul_rn((float)cof, (float)b));
    else out[i] = (TO)__dsub_rn((double)a, __dmul_rn((double)cof, (double)b));
  }
}

int launch_preemphasis(spb_context *ctx, const void *in, int dtype, long long n, long long shift, double cof,
                       void *out) {
  if (n <= 0) return SPB_OK;
  shift %= n;
  if (shift < 0) shift += n;
  const unsigned g = (unsigned)std::min<long long>((n + 255) / 256, (long long)ctx->sm_count * 16);
  if (dtype == SPB_F32) preemphasis_kernel<float, float><<<g, 256, 0, ctx->stream>>>((const float *)in, n, shift, (float)cof, (float *)out);
  else if (dtype == SPB_F64) preemphasis_kernel<double, double><<<g, 256, 0, ctx->stream>>>((const double *)in, n, shift, cof, (double *)out);
  else if (dtype == SPB_I16) preemphasis_kernel<int16_t, double><<<g, 256, 0, ctx->stream>>>((const int16_t *)in, n, shift, cof, (double *)out);
  else if (dtype == SPB_I32) preemphasis_kernel<int32_t, double><<<g, 256, 0, ctx->stream>>>((const int32_t *)in, n, shift, cof, (double *)out);
  else return fail(ctx, SPB_ERR_INVALID, "preemphasis: unsupported dtype");
  SPB_LAUNCH_CHECK(ctx, "preemphasis_kernel");
  return SPB_OK;
}

// ------------------------------------------------------------------ log power spectrum (8f-2)
// processing.py:177-198: 10 * log10(max(P, 1e-20)), minus the global maximum when normalised.
// log10 is monotonic, so the maximum of the logs is the log of the maximum power.
__global__ void max_f32_kernel(const float *__restrict__ p, long long n, unsigned int *__restrict__ out_bits) {
  float m = 0.f;  // powers are >= 0
  bool bad = false;  // np.max propagates NaN (processing.py:196); fmaxf would drop it
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float v = p[i];
    bad |= (v != v);
    m = fmaxf(m, v);
  }
  for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (__any_sync(0xffffffffu, bad)) m = __uint_as_float(0x7fc00000u);  // orders above every finite value as uint
  if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));  // non-negative floats order like uints
}

__global__ void log_power_kernel(const float *__restrict__ p, long long n, const unsigned int *__restrict__ max_bits,
                                 double *__restrict__ out) {
  double sub = 0.0;
  if (max_bits) {
    const double pm = (double)__uint_as_float(*max_bits);
    sub = __dmul_rn(10.0, log10(pm <= 1e-20 ? 1e-20 : pm));
  }
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const double v = (double)p[i];
    out[i] = __dsub_rn(__dmul_rn(10.0, log10(v <= 1e-20 ? 1e-20 : v)), sub);  // no FMA: the maximum maps to exactly 0
  }
}

int launch_log_power(spb_context *ctx, const float *p, long long n, int normalize, unsigned int *d_max_bits,
                     double *out) {
  if (n <= 0) return SPB_OK;
  const unsigned g = (unsigned)std::min<long long>((n + 255) / 256, (long long)ctx->sm_count * 16);
  if (normalize) {
    cudaError_t e = cudaMemsetAsync(d_max_bits, 0, 4, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, e, "cudaMemsetAsync(log power max)");
    max_f32_kernel<<<g, 256, 0, ctx->stream>>>(p, n, d_max_bits);
    SPB_LAUNCH_CHECK(ctx, "max_f32_kernel");
  }
  log_power_kernel<<<g, 256, 0, ctx->stream>>>(p, n, normalize ? d_max_bits : nullptr, out);
  SPB_LAUNCH_CHECK(ctx, "log_power_kernel");
  return SPB_OK;
}

// ------------------------------------------------------------------ tile tables
// tiles[tile_offsets[u] + t] = {u, t * tile} for t < ceil(nframes[u] / tile)
__global__ void build_tiles_kernel(const int32_t *__restrict__ nframes,
                                   const int64_t *__restrict__ tile_offsets, long long U, int tile,
                                   int2 *__restrict__ tiles) {
  // one warp per utterance
  const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= U) return;
  const int nt = (nframes[w] + tile - 1) / tile;
  const long long base = tile_offsets[w];
  for (int t = lane; t < nt; t += 32) tiles[base + t] = make_int2((int)w, t * tile);
}

int launch_build_tiles(spb_context *ctx, const int32_t *nframes, const int64_t *tile_offsets,
                       int64_t U, int tile, int2 *tiles) {
  if (U <= 0) return SPB_OK;
  const long long threads = U * 32;
  const unsigned blocks = (unsigned)((threads + 255) / 256);
  build_tiles_kernel<<<blocks, 256, 0, ctx->stream>>>(nframes, tile_offsets, U, tile, tiles);
  SPB_LAUNCH_CHECK(ctx, "build_tiles_kernel");
  return SPB_OK;
}

__global__ void build_fft_tiles_kernel(const int32_t *__restrict__ nframes,
                                       const int64_t *__restrict__ tile_offsets,
                                       const int64_t *__restrict__ starts,
                                       const int64_t *__restrict__ frame_offsets, long long U, int tile,
                                       int stride, FftTile *__restrict__ tiles) {
  const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= U) return;
  const int F = nframes[w];
  const int nt = (F + tile - 1) / tile;
  const long long base = tile_offsets[w];
  for (int t = lane; t < nt; t += 32) {
    FftTile d;
    d.utt = (int)w;
    d.f0 = t * tile;
    d.nf = min(tile, F - d.f0);
    d.pad = 0;
    d.gstart = starts[w] + (long long)d.f0 * stride;
    d.row0 = frame_offsets[w] + d.f0;
    tiles[base + t] = d;
  }
}

int launch_build_fft_tiles(spb_context *ctx, const int32_t *nframes, const int64_t *tile_offsets,
                           const int64_t *starts, const int64_t *frame_offsets, int64_t U, int tile,
                           int stride, FftTile *tiles) {
  if (U <= 0) return SPB_OK;
  const long long threads = U * 32;
  const unsigned blocks = (unsigned)((threads + 255) / 256);
  build_fft_tiles_kernel<<<blocks, 256, 0, ctx->stream>>>(nframes, tile_offsets, starts, frame_offsets, U, tile,
                                                          stride, tiles);
  SPB_LAUNCH_CHECK(ctx, "build_fft_tiles_kernel");
  return SPB_OK;
}

// ------------------------------------------------------------------------------------
// power / magnitude / log-power spectrum for ANY fft_points (numpy's rfft(n=) takes any n): direct DFT
// in float64 (processing.py:142-198).  Used for the sizes the FFT kernels do not cover (not a power of
// two, or outside [32, 4096]); one CTA per frame, twiddles e^{-2 pi i r / n} in shared memory,
// a thread per bin.  O(n^2) per frame: the rare case, not a hot path.
//   mode 0: |X|^2 / n   1: |X|   2, 3: 10 log10(max(|X|^2 / n, 1e-20)) (3: the caller subtracts the maximum)
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
spectrum_dft_kernel(const double *__restrict__ frames, long long F, int L, int nfft, int mode, double *__restrict__ out) {
  extern __shared__ __align__(16) unsigned char dft_smem[];
  double2 *tw = reinterpret_cast<double2 *>(dft_smem);  // [nfft]
  double *x = reinterpret_cast<double *>(tw + nfft);     // [Leff]
  const int Leff = min(L, nfft);                          // rows longer than n are truncated, shorter ones zero-padded
  const int K = nfft / 2 + 1;
  for (int r = threadIdx.x; r < nfft; r += blockDim.x) {
    double sn, cs;
    sincospi(2.0 * (double)r / (double)nfft, &sn, &cs);
    tw[r] = make_double2(cs, -sn);
  }
  for (long long f = blockIdx.x; f < F; f += gridDim.x) {
    __syncthreads();
    for (int n = threadIdx.x; n < Leff; n += blockDim.x) x[n] = frames[f * L + n];
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
      double re = 0.0, im = 0.0;
      int r = 0;
      for (int n = 0; n < Leff; ++n) {
        const double2 w = tw[r];
        re = fma(x[n], w.x, re);
        im = fma(x[n], w.y, im);
        r += k;
        if (r >= nfft) r -= nfft;
      }
      const double p2 = re * re + im * im;
      double v;
      if (mode == 1) v = sqrt(p2);
      else {
        v = p2 / (double)nfft;
        if (mode >= 2) v = 10.0 * log10(fmax(v, 1e-20));
      }
      out[f * K + k] = v;
    }
  }
}

__global__ void __launch_bounds__(256) subtract_max_f64_kernel(double *__restrict__ v, long long n, const double *__restrict__ mx) {
  const double m = *mx;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) v[i] -= m;
}
__global__ void __launch_bounds__(256) max_f64_kernel(const double *__restrict__ v, long long n, double *__restrict__ mx) {
  // single CTA: the spectra this path sees are small
  __shared__ double sm[256];
  double m = -INFINITY;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) m = fmax(m, v[i]);
  sm[threadIdx.x] = m;
  __syncthreads();
  for (int o = 128; o >= 1; o >>= 1) {
    if ((int)threadIdx.x < o) sm[threadIdx.x] = fmax(sm[threadIdx.x], sm[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) *mx = sm[0];
}

// d_frames (F, L) float64 on the device -> d_out (F, nfft/2+1) float64 on the device; d_tmp: 8 bytes
int launch_spectrum_dft(spb_context *ctx, const double *d_frames, int64_t F, int L, int nfft, int mode, double *d_out,
                        double *d_tmp) {
  if (F <= 0) return SPB_OK;
  const size_t smem = (size_t)nfft * 16 + (size_t)std::min(L, nfft) * 8;
  if (smem > 200 * 1024) return fail(ctx, SPB_ERR_UNSUPPORTED, "fft_points too large for the direct-DFT spectrum kernel");
  int per_sm = 0;
  int rc = configure_kernel(ctx, (const void *)spectrum_dft_kernel, 256, smem, &per_sm);
  if (rc) return rc;
  const long long grid = std::min<long long>(F, (long long)ctx->sm_count * std::max(per_sm, 1));
  spectrum_dft_kernel<<<(unsigned)grid, 256, smem, ctx->stream>>>(d_frames, F, L, nfft, mode, d_out);
  SPB_LAUNCH_CHECK(ctx, "spectrum_dft_kernel");
  if (mode == 3) {
    const long long n = (long long)F * (nfft / 2 + 1);
    max_f64_kernel<<<1, 256, 0, ctx->stream>>>(d_out, n, d_tmp);
    SPB_LAUNCH_CHECK(ctx, "max_f64_kernel");
    subtract_max_f64_kernel<<<(unsigned)std::min<long long>((n + 255) / 256, 1024), 256, 0, ctx->stream>>>(d_out, n, d_tmp);
    SPB_LAUNCH_CHECK(ctx, "subtract_max_f64_kernel");
  }
  return SPB_OK;
}

}  // namespace spb
