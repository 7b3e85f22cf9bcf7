// Exactness guard of the log-mel kernels (north_star: |log-mel error| <= 1e-4 on EVERY value).
//
// The fp32 feature kernels (tensor-core hop DFT, CUDA-core FFTs) carry an absolute error of ~1e-7
// of the frame's spectral level in X[k]; a filter whose energy happens to be ~1e-5 of the level or
// less (one or two bins wide, about one value in 1e5 on noise) would show it in the log domain.  Those
// kernels compare every filter energy with tau x the level of its frame and append the few that
// fall below to a list; the kernels here recompute exactly those values in float64, straight from
// the samples, as the reference does (feature.py:199-217,256: |rfft|^2 / NFFT, dot with the bank,
// zero_handling, log):
//
//   redo_list_kernel  one warp per listed (frame, filter): direct DFT of the filter's bins over the
//                     frame's samples with a float64 twiddle table, float64 mel sum and log
//   redo_all_kernel   the same for every (frame, filter) of a launch, run only when the list
//                     overflowed (pathological input: e.g. a pure tone far from most filters)
#include <math.h>

#include "spb_internal.cuh"

namespace spb {

struct RedoParams {
  const float *samples;
  RedoList list;
  const int *skip_if;        // list kernel: do nothing when *skip_if != 0 (the fallback kernel rewrote the rows)
  const FftTile *ftiles;     // all kernel: the launch's 32-frame tiles
  long long num_tiles;
  int L, S, nfft, M;
  const int32_t *band_lo, *band_n, *band_off;
  const float *band_w;
  const double2 *tw;         // [nfft] (cos, -sin)(2 pi r / nfft)
  int mfe;                   // 1: energies (no log)
  int mode;                  // FEAT_MFE / FEAT_LMFE: list entries are values; FEAT_MFCC / FEAT_MFCC_DELTA: rows
  int C, dc_elim;            // mfcc: cepstra kept, c0 <- ln(frame energy)
  const float *dct;          // [C][M] DCT-II (ortho) rows
  float *out;
  double pre_cof;            // pre-emphasis of the plan (0: off), applied in float64 like processing.py:45-58
  const int64_t *last;       // [U] index of each utterance's last sample
  const int64_t *frame_offsets;  // [U+1]
  long long num_utts;
};

// the sample in front of a frame that starts at arena offset soff and is output row `row`: the previous sample,
// or -- first frame of its utterance (np.roll is circular) -- the utterance's last one
__device__ __forceinline__ const float *redo_prior(const RedoParams &p, const long long soff, const long long row) {
  long long lo = 0, hi = p.num_utts;  // frame_offsets[lo] <= row < frame_offsets[hi]
  while (hi - lo > 1) {
    const long long mid = (lo + hi) >> 1;
    if (p.frame_offsets[mid] <= row) lo = mid; else hi = mid;
  }
  return (p.frame_offsets[lo] == row) ? p.samples + p.last[lo] : p.samples + soff - 1;
}

// one (frame, filter) by a whole warp; every lane returns the filter energy (zero-handled), float64
__device__ __forceinline__ double redo_energy(const RedoParams &p, const float *x, const float *xprior, const int m, const int lane) {
  const int lo = p.band_lo[m], n = p.band_n[m];
  const float *w = p.band_w + p.band_off[m];
  const int Leff = min(p.L, p.nfft);  // numpy rfft(n=): rows longer than NFFT are truncated
  const int mask = p.nfft - 1;
  // frames of up to 320 samples (the speechpy default): a lane's ten samples are read and widened once, not once per bin
  constexpr int kPer = 10;
  const bool small = Leff <= 32 * kPer;
  double xv[kPer];
  if (small) {
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
      const int i = lane + 32 * j;
      double v = 0.0;
      if (i < Leff) {
        v = (double)x[i];
        if (p.pre_cof != 0.0) v -= p.pre_cof * (double)(i > 0 ? x[i - 1] : *xprior);
      }
      xv[j] = v;
    }
  }
  double acc = 0.0;
  for (int t = 0; t < n; ++t) {
    const double wt = (double)w[t];
    if (wt == 0.0) continue;  // zero padding of the tap table (warp-uniform)
    const int k = lo + t;
    double re = 0.0, im = 0.0;
    if (small) {
#pragma unroll
      for (int j = 0; j < kPer; ++j) {
        const int i = lane + 32 * j;
        const double2 c = p.tw[(k * i) & mask];  // k, i < 4096; rows past Leff carry xv = 0
        re = fma(xv[j], c.x, re);
        im = fma(xv[j], c.y, im);
      }
    } else {
      for (int i = lane; i < Leff; i += 32) {
        const double2 c = p.tw[(int)(((long long)k * i) & mask)];
        double v = (double)x[i];
        if (p.pre_cof != 0.0) v -= p.pre_cof * (double)(i > 0 ? x[i - 1] : *xprior);
        re = fma(v, c.x, re);
        im = fma(v, c.y, im);
      }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      re += __shfl_xor_sync(0xffffffffu, re, o);
      im += __shfl_xor_sync(0xffffffffu, im, o);
    }
    acc = fma(wt, (re * re + im * im) / (double)p.nfft, acc);
  }
  if (acc == 0.0) acc = 2.220446049250313e-16;  // functions.zero_handling (functions.py:55-62)
  return acc;
}

__device__ __forceinline__ float redo_value(const RedoParams &p, const float *x, const float *xprior, const int m, const int lane) {
  const double acc = redo_energy(p, x, xprior, m, lane);
  return (float)(p.mfe ? acc : log(acc));
}

// one mfcc (+ delta) row by a whole warp, float64 all the way (feature.py:102-153, 261-282): log-mel of every
// filter, DCT-II rows, c0 <- ln(frame energy) with energy = sum over the one-sided power spectrum
// (feature.py:203-205), deltas along the feature axis as the reference writes them (processing.py:223-236).
// lm: M doubles, cep: C doubles of per-warp shared memory.
__device__ __forceinline__ void redo_mfcc_row(const RedoParams &p, const float *x, const float *xprior, const long long row,
                                              const int lane, double *lm, double *cep) {
  const int M = p.M, C = p.C;
  for (int m = 0; m < M; ++m) {
    const double e = redo_energy(p, x, xprior, m, lane);
    if (lane == 0) lm[m] = log(e);
  }
  double en = 0.0;
  if (p.dc_elim) {
    const int Leff = min(p.L, p.nfft);
    double s2 = 0.0, s0 = 0.0, sa = 0.0;
    for (int i = lane; i < Leff; i += 32) {
      double v = (double)x[i];
      if (p.pre_cof != 0.0) v -= p.pre_cof * (double)(i > 0 ? x[i - 1] : *xprior);
      s2 = fma(v, v, s2);
      s0 += v;
      sa += (i & 1) ? -v : v;
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
      s0 += __shfl_xor_sync(0xffffffffu, s0, o);
      sa += __shfl_xor_sync(0xffffffffu, sa, o);
    }
    // sum_{k=0}^{N/2} |X_k|^2 / N = (N sum x^2 + X_0^2 + X_{N/2}^2) / (2 N)
    en = ((double)p.nfft * s2 + s0 * s0 + sa * sa) / (2.0 * (double)p.nfft);
    if (en == 0.0) en = 2.220446049250313e-16;
  }
  __syncwarp();
  for (int c = lane; c < C; c += 32) {
    double acc = 0.0;
    for (int m = 0; m < M; ++m) acc = fma((double)p.dct[c * M + m], lm[m], acc);
    if (c == 0 && p.dc_elim) acc = log(en);
    cep[c] = acc;
  }
  __syncwarp();
  auto X = [&](int j) { return cep[min(j, C - 1)]; };
  auto D1 = [&](int j) {
    j = min(j, C - 1);
    return (X(j + 1) + 2.0 * X(j + 2)) / 10.0;
  };
  for (int c = lane; c < C; c += 32) {
    if (p.mode == FEAT_MFCC) {
      p.out[row * C + c] = (float)cep[c];
    } else {
      float *o = p.out + (row * C + c) * 3;
      o[0] = (float)cep[c];
      o[1] = (float)D1(c);
      o[2] = (float)((D1(c + 1) + 2.0 * D1(c + 2)) / 10.0);
    }
  }
  __syncwarp();
}

__global__ void __launch_bounds__(256) redo_list_kernel(const RedoParams p) {
  extern __shared__ double redo_smem[];  // mfcc modes: per warp M + C doubles
  if (p.skip_if != nullptr && *p.skip_if != 0) return;
  const int total = min(*p.list.count, p.list.cap);
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const bool rows = p.mode >= FEAT_MFCC;
  double *lm = redo_smem + (threadIdx.x >> 5) * (p.M + p.C);
  for (int e = blockIdx.x * wpb + (threadIdx.x >> 5); e < total; e += gridDim.x * wpb) {
    const RedoEntry en = p.list.entries[e];
    const long long row = rows ? en.oidx : en.oidx / p.M;
    const float *xp = (p.pre_cof != 0.0) ? redo_prior(p, en.soff, row) : nullptr;
    if (rows) {
      redo_mfcc_row(p, p.samples + en.soff, xp, row, lane, lm, lm + p.M);
    } else {
      const int m = (int)(en.oidx % p.M);
      const float v = redo_value(p, p.samples + en.soff, xp, m, lane);
      if (lane == 0) p.out[en.oidx] = v;
    }
  }
}

__global__ void __launch_bounds__(256) redo_all_kernel(const RedoParams p) {
  extern __shared__ double redo_smem[];
  if (*p.list.overflow == 0) return;
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const bool rows = p.mode >= FEAT_MFCC;
  double *lm = redo_smem + (threadIdx.x >> 5) * (p.M + p.C);
  const int per = rows ? 1 : p.M;
  const long long units = p.num_tiles * kFftTileFrames * per;
  for (long long u = (long long)blockIdx.x * wpb + (threadIdx.x >> 5); u < units; u += (long long)gridDim.x * wpb) {
    const long long tf = u / per;
    const int m = (int)(u - tf * per);
    const long long tile = tf / kFftTileFrames;
    const int f = (int)(tf - tile * kFftTileFrames);
    const FftTile td = p.ftiles[tile];
    if (f >= td.nf) continue;
    const long long soff = td.gstart + (long long)f * p.S;
    const float *xp = (p.pre_cof == 0.0) ? nullptr : (td.f0 + f == 0) ? p.samples + p.last[td.utt] : p.samples + soff - 1;
    if (rows) {
      redo_mfcc_row(p, p.samples + soff, xp, td.row0 + f, lane, lm, lm + p.M);
    } else {
      const float v = redo_value(p, p.samples + soff, xp, m, lane);
      if (lane == 0) p.out[(td.row0 + f) * p.M + m] = v;
    }
  }
}

// Runs the guard's second half for one feature launch: the listed values, then (guarded by the
// overflow flag) everything.  skip_if: device flag that voids the list (tensor-core fallback).
int launch_redo(spb_context *ctx, const FeatArgs &a, const RedoList &list, const int *skip_if, const FftTables &t) {
  if (list.entries == nullptr || a.M < 1) return SPB_OK;
  RedoParams p;
  p.samples = a.samples;
  p.list = list;
  p.skip_if = skip_if;
  p.ftiles = a.ftiles;
  p.num_tiles = a.num_tiles;
  p.L = a.L;
  p.S = a.S;
  p.nfft = a.nfft;
  p.M = a.M;
  p.band_lo = a.band_lo;
  p.band_n = a.band_n;
  p.band_off = a.band_off;
  p.band_w = a.band_w;
  p.tw = t.d_tw64;
  p.mfe = (a.mode == FEAT_MFE) ? 1 : 0;
  p.mode = a.mode;
  p.C = a.C;
  p.dc_elim = a.dc_elim;
  p.dct = a.dct;
  p.out = a.out;
  p.pre_cof = a.pre_cof;
  p.last = a.last;
  p.frame_offsets = a.frame_offsets;
  p.num_utts = a.num_utts;
  const int grid = ctx->sm_count * 4;
  const size_t smem = (a.mode >= FEAT_MFCC) ? (size_t)8 * (a.M + a.C) * sizeof(double) : 0;
  redo_list_kernel<<<grid, 256, smem, ctx->stream>>>(p);
  SPB_LAUNCH_CHECK(ctx, "redo_list");
  if (a.ftiles != nullptr && a.num_tiles > 0) {
    redo_all_kernel<<<grid, 256, smem, ctx->stream>>>(p);
    SPB_LAUNCH_CHECK(ctx, "redo_all");
  }
  return SPB_OK;
}

}  // namespace spb
