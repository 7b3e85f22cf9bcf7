// K1/K2: fused framing -> real FFT -> |X|^2 -> mel -> log -> (DCT, c0, deltas).
//
// Replaces, per frame, the numpy pipeline of
//   speech_feature_extraction/speechpy/processing.py:123-138 (frame gather)
//   processing.py:158-159,174                              (rfft, |.|^2/NFFT)
//   feature.py:202-217,256                                 (energy, mel, zero_handling, log)
//   feature.py:146-152                                     (DCT-II ortho, c0 <- ln energy)
//   processing.py:216-236 via feature.py:272-278           (delta cube, as written)
//
// Two kernels share the output stage:
//   * features_fast512: NFFT = 512, 16 lanes per frame, 256-point complex FFT as
//     16 x 16 register FFTs with one swizzled shared-memory exchange, real-FFT
//     untangle through 16-lane shuffles; 32-frame tiles, persistent CTAs.
//   * features_generic: any power-of-two NFFT in [32, 4096], one warp per frame,
//     radix-2 in shared memory.  Used for NFFT != 512 and as an on-device cross-check.
#include <math.h>

#include "fft_math.cuh"
#include "fft_math2.cuh"
#include "ptx_helpers.cuh"
#include "spb_internal.cuh"

namespace spb {

static constexpr float kEpsF64AsF32 = 2.220446049250313e-16f;  // functions.py:62

struct FeatParams {
  const float *samples;
  const int64_t *starts;
  const int32_t *nframes;
  const int64_t *frame_offsets;
  const int2 *tiles;
  const FftTile *ftiles;
  long long num_tiles;
  int L, S, K;
  int M, C, dc_elim;
  const int32_t *band_lo, *band_n, *band_off;
  const float *band_w;
  int band_total;
  const float *dct;
  const float4 *seg_rd;
  const int4 *seg_ranges;
  int seg_kmin, seg_nbins;
  int mode;
  float *out;
  float *energy;
  const float2 *twA;
  const float2 *twB;
  const float2 *tw;
  int nfft, log2nc;
  int need_energy;
  int need_level;     // per-frame sum x^2 for the exactness guard (redo.entries != nullptr, mfe / lmfe output)
  RedoList redo;      // filters below tau x sum x^2 are listed for the float64 redo (redo_kernels.cu)
  const int *run_if;  // fast512: when set, the launch is a no-op unless *run_if != 0
  float pre_cof;      // pre-emphasis fused into the framing (0: off); processing.py:45-58, shift = 1
  const int64_t *last;  // [U] index of each utterance's last sample: what sample 0 wraps to (np.roll)
};

// Pre-emphasis of a staged tile, in place: y[i] = x[i] - cof x[i-1], `prior` = the sample in front of the
// tile's first one (the utterance's LAST sample for its first tile: np.roll is circular).  Every thread owns
// a run of k consecutive samples (k odd: the strided walks touch 32 different banks) and reads it, with the
// sample in front of it, into registers before anybody stores (longer runs: walked downwards, so that x[i-1] is
// still the original when y[i] is formed).  Called by the whole CTA; ends with a barrier.
__device__ __forceinline__ void preemphasis_tile(float *s, const int count, const float prior, const float cof,
                                                 const int tid, const int nthreads) {
  constexpr int kRun = 24;  // runs up to this length go through registers (the default tile: 5 280 samples / 256 threads = 21)
  const int k = ((count + nthreads - 1) / nthreads) | 1;
  const int lo = tid * k, hi = min(lo + k, count);
  float front = 0.f;
  if (lo < count) front = lo > 0 ? s[lo - 1] : prior;
  if (k <= kRun) {
    float r[kRun];
#pragma unroll
    for (int j = 0; j < kRun; ++j) r[j] = (lo + j < hi) ? s[lo + j] : 0.f;
    __syncthreads();  // every thread holds its run and the sample in front of it
#pragma unroll
    for (int j = 0; j < kRun; ++j)
      if (lo + j < hi) s[lo + j] = fmaf(-cof, j > 0 ? r[j - 1] : front, r[j]);
  } else {
    __syncthreads();
    for (int i = hi - 1; i >= lo; --i) {
      const float pv = (i > lo) ? s[i - 1] : front;
      s[i] = fmaf(-cof, pv, s[i]);
    }
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------
// Output stage shared by both kernels.  s_p: [nf][pstride] UNSCALED |X|^2 rows (bins
// 0..pbins-1), s_en: per-frame energy (already scaled, zero-handled) or unused.
// Called by the whole CTA after a __syncthreads().  s_a / s_b: scratch of TILE*(41|..).
// ------------------------------------------------------------------------------------
// Band weights are zero-padded to multiples of four (offsets too), so the inner loop is a
// 128-bit broadcast load of four weights + four FFMAs; P rows carry >= 3 zeroed pad bins
// after the last real bin so the padded taps read finite values.
__device__ __forceinline__ int tid0_div(int a, int b) { return a / b; }

struct FmIndex {  // tid -> (frame, filter) walk over a [frames][M] tile, hoisted out of the tile loop
  int f0, m0, df, dm;
};

template <int TILE, bool SEG, int MODE_CT = -1>
__device__ __forceinline__ void output_stage(const FeatParams &p, const float *s_p, int pstride,
                                             const float *s_en, float *s_a, float *s_b,
                                             const float *s_bw, const int *s_bmeta,
                                             const float *s_dct, int nf, long long row0,
                                             const float4 *s_rd, const int4 *s_wr,
                                             float *s_U, float *s_D, const FmIndex fm,
                                             const float *s_lvl, const long long gstart) {
  const int tid = threadIdx.x;
  const int nthreads = blockDim.x;
  const float scale = 1.0f / (float)p.nfft;  // |X|^2 / NFFT  (processing.py:174)
  if (p.mode == FEAT_POWER || p.mode == FEAT_MAG) {
    // rows are contiguous in global memory: (row0 + f) * K + k
    const int K = p.K;
    float *dst = p.out + row0 * K;
    int f = tid / K, k = tid - f * K;
    const int df = nthreads / K, dk = nthreads - df * K;
    while (f < nf) {
      const float v = s_p[f * pstride + k];
      dst[f * K + k] = (p.mode == FEAT_POWER) ? v * scale : sqrtf(v);
      k += dk; f += df;
      if (k >= K) { k -= K; ++f; }
    }
    return;
  }
  const int M = p.M;
  const int astride = M + 1;  // odd-ish padding: lane = frame accesses stay conflict-light
  if (SEG) {
    // Segment form (triangular bank): lane = frame, each warp walks its run of whole
    // segments once; bin k feeds the rising sum of filter seg and the falling sum of filter
    // seg-1.  U[f][j] = rising sum of filter j, D[f][j] = falling sum of filter j-1.
    // TILE consecutive threads (a warp, or a half-warp for 16-frame tiles) = the frames of the
    // tile; each such group walks one of the 8 bin ranges
    const int lane = tid % TILE, range = tid / TILE;
    const int upitch = (M + 2) | 1;
    const int4 wr = s_wr[range < 8 ? range : 0];
    if (range < 8 && lane < nf) {
      const float *pp = s_p + lane * pstride + wr.x;
      const float4 *rd = s_rd + (wr.x - p.seg_kmin);
      float *pu = s_U + lane * upitch + wr.z;
      float *pd = s_D + lane * upitch + wr.z;
      float up = 0.f, dn = 0.f;
      const int n = wr.y - wr.x;
#pragma unroll 4
      for (int i = 0; i < n; ++i) {
        const float4 c = rd[i];
        const float pv = pp[i];
        up = fmaf(c.x, pv, up);
        dn = fmaf(c.y, pv, dn);
        if (c.z != 0.f) {
          *pu++ = up;
          *pd++ = dn;
          up = dn = 0.f;
        }
      }
    }
    __syncthreads();
    // thread = (filter m fixed, frames f0, f0 + rpp, ...): rpp = threads / M whole rows per sweep
    const int mode = (MODE_CT >= 0) ? MODE_CT : p.mode;
    const bool direct = (mode == FEAT_MFE || mode == FEAT_LMFE);
    const int rpp = fm.df;
    if (fm.f0 < rpp) {
      const int m = fm.m0;
      const float *pu = s_U + fm.f0 * upitch + m;
      const float *pd = s_D + fm.f0 * upitch + m + 1;
      float *pa = s_a + fm.f0 * astride + m;
      float *dst = p.out + (row0 + fm.f0) * M + m;  // tile rows are contiguous in global memory
      const int ustep = rpp * upitch, astep = rpp * astride, dstep = rpp * M;
      const bool guard = p.need_level != 0;
      for (int f = fm.f0; f < nf; f += rpp) {
        const float raw = *pu + *pd;
        // exactness guard: fp32 cannot resolve a filter this far below the frame's level (mfe / lmfe: the value
        // is listed; mfcc modes: its whole row, which mixes the filters)
        if (guard && raw < p.redo.tau * s_lvl[f])
          redo_push(p.redo, gstart + (long long)f * p.S, direct ? (row0 + f) * (long long)M + m : row0 + f);
        float acc = raw * scale;
        if (acc == 0.f) acc = kEpsF64AsF32;  // functions.zero_handling (feature.py:217)
        const float v = (mode == FEAT_MFE) ? acc : __logf(acc);
        if (direct) *dst = v; else *pa = v;
        pu += ustep; pd += ustep; pa += astep; dst += dstep;
      }
    }
    if (direct) {
      if (mode == FEAT_MFE && p.energy != nullptr)
        for (int ff = tid; ff < nf; ff += nthreads) p.energy[row0 + ff] = s_en[ff];
      return;
    }
  } else {
  // mel energies: thread (frame = tid % TILE, filter slot = tid / TILE)
  {
    const int f = tid % TILE;
    const int slot = tid / TILE;
    const int nslots = nthreads / TILE;
    if (f < nf) {
      const float *prow = s_p + f * pstride;
      for (int m = slot; m < M; m += nslots) {
        const int lo = s_bmeta[m], n = s_bmeta[M + m], off = s_bmeta[2 * M + m];
        const float4 *w4 = reinterpret_cast<const float4 *>(s_bw + off);
        const float *pr = prow + lo;
        float acc = 0.f;
        for (int k = 0; k < n; k += 4) {
          const float4 w = w4[k >> 2];
          acc = fmaf(w.x, pr[k], acc);
          acc = fmaf(w.y, pr[k + 1], acc);
          acc = fmaf(w.z, pr[k + 2], acc);
          acc = fmaf(w.w, pr[k + 3], acc);
        }
        if (p.need_level != 0 && acc < p.redo.tau * s_lvl[f])
          redo_push(p.redo, gstart + (long long)f * p.S,
                    (p.mode == FEAT_MFE || p.mode == FEAT_LMFE) ? (row0 + f) * (long long)M + m : row0 + f);
        acc *= scale;
        if (acc == 0.f) acc = kEpsF64AsF32;  // functions.zero_handling (feature.py:217)
        s_a[f * astride + m] = (p.mode == FEAT_MFE) ? acc : __logf(acc);
      }
    }
  }
  }
  __syncthreads();
  if (p.mode == FEAT_MFE || p.mode == FEAT_LMFE) {
    float *dst = p.out + row0 * M;
    int f = tid / M, m = tid - f * M;
    const int df = nthreads / M, dm = nthreads - df * M;
    while (f < nf) {
      dst[f * M + m] = s_a[f * astride + m];
      m += dm; f += df;
      if (m >= M) { m -= M; ++f; }
    }
    if (p.mode == FEAT_MFE && p.energy != nullptr) {
      for (int ff = tid; ff < nf; ff += nthreads) p.energy[row0 + ff] = s_en[ff];
    }
    return;
  }
  // MFCC: DCT-II (ortho) of the log-mel row, first C coefficients (feature.py:147)
  const int C = p.C;
  const int cstride = C + ((C & 1) ? 0 : 1);
  {
    const int f = tid % TILE;
    const int slot = tid / TILE;
    const int nslots = nthreads / TILE;
    if (f < nf) {
      const float *lrow = s_a + f * astride;
      for (int c = slot; c < C; c += nslots) {
        const float *drow = s_dct + c * M;
        float acc = 0.f;
        for (int m = 0; m < M; ++m) acc = fmaf(drow[m], lrow[m], acc);
        if (c == 0 && p.dc_elim) acc = __logf(s_en[f]);  // feature.py:151-152
        s_b[f * cstride + c] = acc;
      }
    }
  }
  __syncthreads();
  int f = tid / C, c = tid - f * C;
  const int df = nthreads / C, dc = nthreads - df * C;
  if (p.mode == FEAT_MFCC) {
    float *dst = p.out + row0 * C;
    while (f < nf) {
      dst[f * C + c] = s_b[f * cstride + c];
      c += dc; f += df;
      if (c >= C) { c -= C; ++f; }
    }
    return;
  }
  // FEAT_MFCC_DELTA: [x, D(x), D(D(x))] interleaved on a new last axis
  // D(x)[c] = (x[min(c+1,C-1)] + 2 x[min(c+2,C-1)]) / 10   (processing.py:223-236 as written)
  {
    float *dst = p.out + row0 * C * 3;
    while (f < nf) {
      const float *x = s_b + f * cstride;
      auto X = [&](int j) { return x[min(j, C - 1)]; };
      auto D1 = [&](int j) {
        j = min(j, C - 1);
        return (X(j + 1) + 2.f * X(j + 2)) / 10.f;
      };
      float *o = dst + (f * C + c) * 3;
      o[0] = x[c];
      o[1] = D1(c);
      o[2] = (D1(c + 1) + 2.f * D1(c + 2)) / 10.f;
      c += dc; f += df;
      if (c >= C) { c -= C; ++f; }
    }
  }
}

__device__ __forceinline__ void load_plan_to_smem(const FeatParams &p, float *s_bw, int *s_bmeta,
                                                  float *s_dct) {
  if (p.mode < FEAT_MFE) return;
  for (int i = threadIdx.x; i < p.band_total; i += blockDim.x) s_bw[i] = p.band_w[i];
  for (int i = threadIdx.x; i < p.M; i += blockDim.x) {
    s_bmeta[i] = p.band_lo[i];
    s_bmeta[p.M + i] = p.band_n[i];
    s_bmeta[2 * p.M + i] = p.band_off[i];
  }
  if (p.mode >= FEAT_MFCC)
    for (int i = threadIdx.x; i < p.C * p.M; i += blockDim.x) s_dct[i] = p.dct[i];
}

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
// cp_async_commit / cp_async_wait<N>: ptx_helpers.cuh

// ====================================================================================
// Fast kernel, NFFT = 512.
// ====================================================================================
constexpr int kNW = kFftWarps;            // warps per CTA
constexpr int kTile = kFftTileFrames;     // frames per tile (two per half-warp)
constexpr int kFastThreads = 32 * kNW;

template <int NJ, bool FULL>
__global__ void __launch_bounds__(kFastThreads, 16 / kNW)
features_fast512(const FeatParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (p.run_if != nullptr && *p.run_if == 0) return;  // fallback launch of the tensor-core path
  constexpr int PBINS = FULL ? 257 : 129;
  // 4 zeroed pad bins (padded band taps); stride == 5 (mod 32): conflict-free both for the
  // half-warp row writes (rows 16 apart land 16 banks apart) and the lane = frame column reads
  constexpr int PSTRIDE = PBINS + 4;
  const int L = p.L, S = p.S;
  const int span_max = (kTile - 1) * S + L;
  // carve shared memory
  const int xoff = ((span_max + 3) & ~3);
  float *s_xbuf = reinterpret_cast<float *>(smem_raw);              // 2 x [xoff] sample stages
  float2 *s_exch = reinterpret_cast<float2 *>(s_xbuf + 2 * xoff);   // 16 half-warps x 256
  float *s_p = reinterpret_cast<float *>(s_exch + 2 * kNW * 256);        // [kTile][PSTRIDE]
  float *s_bw = s_p + kTile * PSTRIDE;                                 // [band_total] (16 B aligned)
  float *s_en = s_bw + p.band_total;                                // [kTile]
  float *s_lvl = s_en + kTile;                                      // [kTile] sum x^2 per frame (exactness guard)
  float *s_a = s_lvl + kTile;                                       // [kTile][M+1]
  float *s_b = s_a + kTile * (p.M + 1);                             // [kTile][C+1]
  float *s_dct = s_b + kTile * (p.C + 1);                           // [C*M]
  int *s_bmeta = reinterpret_cast<int *>(s_dct + p.C * p.M);        // [3M]
  // segment form of the bank (16-byte aligned: everything before it is a multiple of 4 floats
  // once padded)
  const int used_floats = (int)(reinterpret_cast<float *>(s_bmeta + 3 * p.M) - reinterpret_cast<float *>(smem_raw));
  float4 *s_rd = reinterpret_cast<float4 *>(reinterpret_cast<float *>(smem_raw) + ((used_floats + 3) & ~3));
  int4 *s_wr = reinterpret_cast<int4 *>(s_rd + p.seg_nbins);
  float2 *s_twA = reinterpret_cast<float2 *>(s_wr + 8);            // [256]
  float2 *s_twB = s_twA + 256;                                      // [128]
  // U / D alias the FFT exchange buffer (32 KB), which is idle during the output stage
  float *s_U = reinterpret_cast<float *>(s_exch);                   // [kTile][(M+2)|1]
  float *s_D = s_U + kTile * ((p.M + 2) | 1);
  FmIndex fm;
  {
    const int Mq = p.M > 0 ? p.M : 1;
    fm.f0 = tid0_div(threadIdx.x, Mq);
    fm.m0 = threadIdx.x - fm.f0 * Mq;
    fm.df = kFastThreads / Mq;
    fm.dm = kFastThreads - fm.df * Mq;
  }
  const bool use_seg = p.seg_rd != nullptr && p.mode >= FEAT_MFE &&
                       2 * kTile * ((p.M + 2) | 1) * 4 <= 2 * kNW * 256 * (int)sizeof(float2);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int half = lane >> 4;
  const int l16 = lane & 15;
  const int hw = warp * 2 + half;  // half-warp id 0..15
  float2 *my_exch = s_exch + hw * 256;

  load_plan_to_smem(p, s_bw, s_bmeta, s_dct);
  if (use_seg) {
    for (int i = tid; i < p.seg_nbins; i += kFastThreads) s_rd[i] = p.seg_rd[i];
    if (tid < 8) s_wr[tid] = p.seg_ranges[tid];
  }
  for (int i = tid; i < kTile * 4; i += kFastThreads)
    s_p[(i >> 2) * PSTRIDE + PBINS + (i & 3)] = 0.f;

  // twiddle tables in shared memory, lane-major so a half-warp reads consecutive float2:
  // s_twA[k1*16 + l] = W256^(l*k1), s_twB[k2*16 + l] = W512^(l + 16*k2)
  for (int i = tid; i < 256; i += kFastThreads) s_twA[i] = p.twA[(i & 15) * 16 + (i >> 4)];
  for (int i = tid; i < 128; i += kFastThreads) s_twB[i] = p.twB[(i & 15) * 9 + (i >> 4)];
  const int partner = (16 - l16) & 15;
  const int Leff = min(L, 512);

  // Stage a tile's samples into s_xbuf[b]: one TMA bulk copy issued by thread 0 when the tile
  // starts on a 16-byte boundary (batches packed by batch.pack do), plain loads otherwise.
  // Each sample is fetched from HBM once and used by L/S frames.
  __shared__ __align__(8) uint64_t s_full[2];
  if (tid == 0) {
    mbar_init(&s_full[0], 1);
    mbar_init(&s_full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t ph0 = 0, ph1 = 0;
  // 32-byte tile descriptors travel through a two-slot shared-memory ring, filled two tiles ahead
  // by thread 0 with cp.async (no register dependency, so nobody stalls on the L2 round trip)
  __shared__ __align__(16) int4 s_fd[2][2];
  auto fetch_desc = [&](long long t, int slot) {  // thread 0 only
    const int4 *src = reinterpret_cast<const int4 *>(p.ftiles + t);
    cp_async16(&s_fd[slot][0], src);
    cp_async16(&s_fd[slot][1], src + 1);
    cp_async_commit();
  };
  auto stage = [&](int slot, int b, long long &row0, int &nf, bool &bulk, long long &gs, float &pri) {
    const int4 d0 = s_fd[slot][0];
    const longlong2 d1 = *reinterpret_cast<const longlong2 *>(&s_fd[slot][1]);
    nf = d0.z;
    const long long gstart = d1.x;
    gs = gstart;
    // pre-emphasis: the sample in front of the tile (first tile of an utterance: its last sample), loaded a
    // tile ahead like the tile itself
    if (p.pre_cof != 0.f) pri = __ldg(p.samples + ((d0.y == 0) ? p.last[d0.x] : gstart - 1));
    row0 = d1.y;
    const int count = (nf - 1) * S + L;
    const float *src = p.samples + gstart;
    float *dst = s_xbuf + b * xoff;
    bulk = ((gstart & 3) == 0) && count >= 4;
    if (bulk) {
      const int n4 = count >> 2;
      if (tid == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&s_full[b], (uint32_t)n4 * 16u);
        bulk_load(dst, src, (uint32_t)n4 * 16u, &s_full[b]);
      }
      for (int i = (n4 << 2) + tid; i < count; i += kFastThreads) dst[i] = __ldg(src + i);
    } else {
      for (int i = tid; i < count; i += kFastThreads) dst[i] = __ldg(src + i);
    }
  };

  long long tile = blockIdx.x;
  int buf = 0;
  long long row0 = 0, gs = 0;
  float pri = 0.f;
  int nf = 0;
  bool bulk = false;
  if (tid == 0) {
    if (tile < p.num_tiles) fetch_desc(tile, 0);
    if (tile + gridDim.x < p.num_tiles) fetch_desc(tile + gridDim.x, 1);
    cp_async_wait<0>();
  }
  __syncthreads();
  if (tile < p.num_tiles) stage(0, 0, row0, nf, bulk, gs, pri);
  for (; tile < p.num_tiles; buf ^= 1) {  // descriptor of the tile in sample buffer b sits in ring slot b
    const long long next = tile + gridDim.x;
    long long n_row0 = 0, n_gs = 0;
    float n_pri = 0.f;
    int n_nf = 0;
    bool n_bulk = false;
    // prefetch the next tile while this one is transformed (its buffer was last read in
    // phase B of the previous iteration, which every thread has left)
    if (next < p.num_tiles) stage(buf ^ 1, buf ^ 1, n_row0, n_nf, n_bulk, n_gs, n_pri);
    // slot `buf` (this tile's descriptor, consumed an iteration ago) takes the tile after next;
    // thread 0 waits for it just before the barrier that ends phase B
    const bool refill = tid == 0 && next + gridDim.x < p.num_tiles;
    if (refill) fetch_desc(next + gridDim.x, buf);
    if (bulk) {
      if (buf == 0) { mbar_wait(&s_full[0], ph0); ph0 ^= 1; }
      else { mbar_wait(&s_full[1], ph1); ph1 ^= 1; }
    }
    __syncthreads();  // samples of this tile visible; previous tile's phase C is finished
    const float *s_x = s_xbuf + buf * xoff;
    if (p.pre_cof != 0.f)  // pre-emphasis as a prologue of the framing: the staged tile, in place
      preemphasis_tile(s_xbuf + buf * xoff, (nf - 1) * S + L, pri, p.pre_cof, tid, kFastThreads);

    // ---- phase B: each half-warp transforms two frames at once (packed fp32, fft_math2.cuh)
    {
      const int flA = 2 * (warp + kNW * half), flB = flA + 1;
      const bool actA = flA < nf, actB = flB < nf;
      const float *xa = s_x + (actA ? flA : 0) * S;
      const float *xb = s_x + (actB ? flB : 0) * S;
      C2 v[16];
      // z[m] = x[2m] + i x[2m+1], m = 16 j + l16
      if (NJ < 16) {
        // compile-time pruned: every lane has exactly NJ non-zero inputs, S even, L == 32*NJ
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          if (j < NJ) {
            // four 32-bit loads straight into the halves of the two register pairs (A.re, B.re),
            // (A.im, B.im): 64-bit loads per frame would need four MOVs per sample pair to
            // transpose (and ptxas re-materialises those at every use)
            const int o = 32 * j + 2 * l16;
            v[j] = c2_make(p_make(xa[o], xb[o]), p_make(xa[o + 1], xb[o + 1]));
          } else {
            v[j] = c2_zero();
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int m = 16 * j + l16;
          const bool r_ok = 2 * m < Leff, i_ok = 2 * m + 1 < Leff;
          v[j] = c2_make(p_make(r_ok ? xa[2 * m] : 0.f, r_ok ? xb[2 * m] : 0.f),
                         p_make(i_ok ? xa[2 * m + 1] : 0.f, i_ok ? xb[2 * m + 1] : 0.f));
        }
      }
      float2 sumsq = p_bc(0.f);
      if (p.need_energy || p.need_level) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < NJ) sumsq = p_fma(v[j].re, v[j].re, p_fma(v[j].im, v[j].im, sumsq));
      }
      // step 1: FFT16 over j -> index k1, then twiddle W256^(l16*k1) (table in shared memory)
      fft16_2<NJ>(v);
#pragma unroll
      for (int k1 = 1; k1 < 16; ++k1) v[k1] = c2_mul(v[k1], s_twA[k1 * 16 + l16]);
      // exchange within the half-warp, real parts then imaginary parts through the same 2 KB
      float2 *ex = my_exch;
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) ex[exch_store_index(k1, l16)] = v[k1].re;
      __syncwarp();
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float4 q = *reinterpret_cast<const float4 *>(ex + exch_load_chunk_index(l16, c));
        v[2 * c].re = make_float2(q.x, q.y);
        v[2 * c + 1].re = make_float2(q.z, q.w);
      }
      __syncwarp();
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) ex[exch_store_index(k1, l16)] = v[k1].im;
      __syncwarp();
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float4 q = *reinterpret_cast<const float4 *>(ex + exch_load_chunk_index(l16, c));
        v[2 * c].im = make_float2(q.x, q.y);
        v[2 * c + 1].im = make_float2(q.z, q.w);
      }
      // step 2: lane = k1 holds n2 = 0..15
      fft16_2<16>(v);  // v[k2] = Z[l16 + 16 k2]
      // untangle: partner value Z[256 - k] lives in lane (16 - l16) & 15
      float *prA = s_p + flA * PSTRIDE, *prB = prA + PSTRIDE;
#pragma unroll
      for (int k2 = 0; k2 < 8; ++k2) {
        // sender slot t = 7 - k2: generic lanes send v[8 + t], lane 0 sends v[(9 + t) & 15]
        const int t = 7 - k2;
        const C2 snd = (l16 == 0) ? v[(9 + t) & 15] : v[8 + t];
        C2 zp;
        zp.re.x = __shfl_sync(0xffffffffu, snd.re.x, partner, 16);
        zp.re.y = __shfl_sync(0xffffffffu, snd.re.y, partner, 16);
        zp.im.x = __shfl_sync(0xffffffffu, snd.im.x, partner, 16);
        zp.im.y = __shfl_sync(0xffffffffu, snd.im.y, partner, 16);
        const C2 zk = v[k2];
        const C2 A = c2_make(p_add(zk.re, zp.re), p_sub(zk.im, zp.im));
        const C2 B = c2_make(p_add(zk.im, zp.im), p_sub(zp.re, zk.re));
        const C2 WB = c2_mul(B, s_twB[k2 * 16 + l16]);
        const float2 xr = p_add(A.re, WB.re), xi = p_add(A.im, WB.im);
        const float2 pk = p_muls(p_fma(xr, xr, p_mul(xi, xi)), 0.25f);
        if (actA) prA[l16 + 16 * k2] = pk.x;
        if (actB) prB[l16 + 16 * k2] = pk.y;
        if (FULL) {
          const float2 yr = p_sub(A.re, WB.re), yi = p_sub(A.im, WB.im);
          const float2 pm = p_muls(p_fma(yr, yr, p_mul(yi, yi)), 0.25f);
          if (actA) prA[256 - (l16 + 16 * k2)] = pm.x;  // k = 0 writes bin 256
          if (actB) prB[256 - (l16 + 16 * k2)] = pm.y;
        }
      }
      if (l16 == 0) {
        // bin 128 = Z[128] (its own partner): X = Zr - i Zi  -> |Z|^2
        const C2 z = v[8];
        const float2 pz = p_fma(z.re, z.re, p_mul(z.im, z.im));
        if (actA) prA[128] = pz.x;
        if (actB) prB[128] = pz.y;
      }
      if (p.need_energy || p.need_level) {
        // frame energy = sum_k P[k], all 257 one-sided bins (feature.py:202), via Parseval:
        // (NFFT*sum x^2 + X0^2 + X256^2) / (2*NFFT), X0 = Zr+Zi, X256 = Zr-Zi of Z[0]
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) {
          sumsq.x += __shfl_xor_sync(0xffffffffu, sumsq.x, o, 16);
          sumsq.y += __shfl_xor_sync(0xffffffffu, sumsq.y, o, 16);
        }
        if (l16 == 0 && p.need_level) {
          if (actA) s_lvl[flA] = sumsq.x;
          if (actB) s_lvl[flB] = sumsq.y;
        }
        if (l16 == 0 && p.need_energy) {
          const float2 x0 = p_add(v[0].re, v[0].im), x256 = p_sub(v[0].re, v[0].im);
          float ea = (512.f * sumsq.x + x0.x * x0.x + x256.x * x256.x) * (1.0f / (2.f * 512.f));
          float eb = (512.f * sumsq.y + x0.y * x0.y + x256.y * x256.y) * (1.0f / (2.f * 512.f));
          if (ea == 0.f) ea = kEpsF64AsF32;  // feature.py:205
          if (eb == 0.f) eb = kEpsF64AsF32;
          if (actA) s_en[flA] = ea;
          if (actB) s_en[flB] = eb;
        }
      }
    }
    if (refill) cp_async_wait<0>();
    __syncthreads();
    // ---- phase C
    if (use_seg && p.mode == FEAT_LMFE)
      output_stage<kTile, true, FEAT_LMFE>(p, s_p, PSTRIDE, s_en, s_a, s_b, s_bw, s_bmeta, s_dct, nf, row0, s_rd,
                                           s_wr, s_U, s_D, fm, s_lvl, gs);
    else if (use_seg)
      output_stage<kTile, true>(p, s_p, PSTRIDE, s_en, s_a, s_b, s_bw, s_bmeta, s_dct, nf, row0, s_rd, s_wr,
                                s_U, s_D, fm, s_lvl, gs);
    else
      output_stage<kTile, false>(p, s_p, PSTRIDE, s_en, s_a, s_b, s_bw, s_bmeta, s_dct, nf, row0, s_rd, s_wr,
                                 s_U, s_D, fm, s_lvl, gs);
    tile = next;
    row0 = n_row0; nf = n_nf; bulk = n_bulk; gs = n_gs; pri = n_pri;
  }
}

// ====================================================================================
// Fast kernel, NFFT = 1024: the reference's own call (load_data.py:64-70 with constants.py:
// FRAME_LEN 0.025 -> L = 400, FRAME_STEP 0.01, NUM_FFT 1024, NUM_COEF 40).
//
// Same pipeline as features_fast512 (TMA-staged sample tiles, descriptor ring, P rows in shared
// memory, segment-form mel, direct stores); the transform is one frame per WARP:
//   z[n] = x[2n] + i x[2n+1], n = 32 j + l (l = lane): 16-point register FFT over j (inputs
//   j >= NZ are zero: L <= 448 -> NZ = 7), twiddle W512^(l k1), exchange through shared memory
//   so that lane (k1, b) holds the 16 values l = 2a + b, 16-point FFT over a, and one radix-2
//   step across the two half-warps (shuffle) -> lane (k1, s) holds Z[k1 + 16 q + 256 s].
//   Real-FFT untangle for bins 0..256 (speechpy's (K+1) bin mapping stops at bin 256 here): the
//   lower half-warp owns them, partners Z[512 - k] come from the upper half by shuffles with
//   static register indices.  Index algebra: tests/test_host.py::test_fast1024_index_algebra.
// 16 warps per CTA (two passes per 32-frame tile), one CTA per SM.
// ====================================================================================
constexpr int kF1kWarps = 16;
constexpr int kF1kThreads = 32 * kF1kWarps;
constexpr int kF1kRS = 34;  // exchange row stride (float2): the 16 rows a read touches spread over all banks

template <int NZ>
__global__ void __launch_bounds__(kF1kThreads, 1)
features_fast1024(const FeatParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int PBINS = 257;
  constexpr int PSTRIDE = PBINS + 4;  // 261 == 5 (mod 32): lane = frame column reads are conflict-free
  const int L = p.L, S = p.S;
  const int span_max = (kTile - 1) * S + L;
  const int xoff = ((span_max + 3) & ~3);
  float *s_xbuf = reinterpret_cast<float *>(smem_raw);                      // 2 x [xoff]
  float2 *s_exch = reinterpret_cast<float2 *>(s_xbuf + 2 * xoff);           // [kF1kWarps][16][kF1kRS]
  float *s_p = reinterpret_cast<float *>(s_exch + kF1kWarps * 16 * kF1kRS); // [kTile][PSTRIDE]
  float *s_bw = s_p + kTile * PSTRIDE;
  float *s_en = s_bw + p.band_total;
  float *s_lvl = s_en + kTile;
  float *s_a = s_lvl + kTile;
  float *s_b = s_a + kTile * (p.M + 1);
  float *s_dct = s_b + kTile * (p.C + 1);
  int *s_bmeta = reinterpret_cast<int *>(s_dct + p.C * p.M);
  const int used_floats = (int)(reinterpret_cast<float *>(s_bmeta + 3 * p.M) - reinterpret_cast<float *>(smem_raw));
  float4 *s_rd = reinterpret_cast<float4 *>(reinterpret_cast<float *>(smem_raw) + ((used_floats + 3) & ~3));
  int4 *s_wr = reinterpret_cast<int4 *>(s_rd + p.seg_nbins);
  float2 *s_tw1 = reinterpret_cast<float2 *>(s_wr + 8);  // [16][32]  W512^(l k1) at k1 * 32 + l
  float2 *s_tw2 = s_tw1 + 512;                           // [16][16]  W1024^(k1 + 16 q) at q * 16 + k1
  float2 *s_tw3 = s_tw2 + 256;                           // [16]      W32^q
  float *s_U = reinterpret_cast<float *>(s_exch);        // [kTile][(M+2)|1], aliases the exchange buffer
  float *s_D = s_U + kTile * ((p.M + 2) | 1);
  FmIndex fm;
  {
    const int Mq = p.M > 0 ? p.M : 1;
    fm.f0 = tid0_div(threadIdx.x, Mq);
    fm.m0 = threadIdx.x - fm.f0 * Mq;
    fm.df = kF1kThreads / Mq;
    fm.dm = kF1kThreads - fm.df * Mq;
  }
  const bool use_seg = p.seg_rd != nullptr && p.mode >= FEAT_MFE &&
                       2 * kTile * ((p.M + 2) | 1) * 4 <= kF1kWarps * 16 * kF1kRS * (int)sizeof(float2);
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int k1r = lane & 15, hb = lane >> 4;
  float2 *my_exch = s_exch + warp * 16 * kF1kRS;

  load_plan_to_smem(p, s_bw, s_bmeta, s_dct);
  if (use_seg) {
    for (int i = tid; i < p.seg_nbins; i += kF1kThreads) s_rd[i] = p.seg_rd[i];
    if (tid < 8) s_wr[tid] = p.seg_ranges[tid];
  }
  for (int i = tid; i < kTile * 4; i += kF1kThreads) s_p[(i >> 2) * PSTRIDE + PBINS + (i & 3)] = 0.f;
  // twiddles (float64 sincospi once per CTA; the kernel is persistent)
  for (int i = tid; i < 512 + 256 + 16; i += kF1kThreads) {
    double ang;  // in units of pi
    if (i < 512) ang = -(double)((i & 31) * (i >> 5)) / 256.0;
    else if (i < 768) ang = -(double)(((i - 512) & 15) + 16 * ((i - 512) >> 4)) / 512.0;
    else ang = -(double)(i - 768) / 16.0;
    double sn, cs;
    sincospi(ang, &sn, &cs);
    s_tw1[i] = make_float2((float)cs, (float)sn);  // the three tables are contiguous
  }
  const int Leff = min(L, 1024);

  __shared__ __align__(8) uint64_t s_full[2];
  if (tid == 0) {
    mbar_init(&s_full[0], 1);
    mbar_init(&s_full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t ph0 = 0, ph1 = 0;
  __shared__ __align__(16) int4 s_fd[2][2];
  auto fetch_desc = [&](long long t, int slot) {  // thread 0 only
    const int4 *src = reinterpret_cast<const int4 *>(p.ftiles + t);
    cp_async16(&s_fd[slot][0], src);
    cp_async16(&s_fd[slot][1], src + 1);
    cp_async_commit();
  };
  auto stage = [&](int slot, int b, long long &row0, int &nf, bool &bulk, long long &gs, float &pri) {
    const int4 d0 = s_fd[slot][0];
    const longlong2 d1 = *reinterpret_cast<const longlong2 *>(&s_fd[slot][1]);
    nf = d0.z;
    const long long gstart = d1.x;
    gs = gstart;
    // pre-emphasis: the sample in front of the tile (first tile of an utterance: its last sample), loaded a
    // tile ahead like the tile itself
    if (p.pre_cof != 0.f) pri = __ldg(p.samples + ((d0.y == 0) ? p.last[d0.x] : gstart - 1));
    row0 = d1.y;
    const int count = (nf - 1) * S + L;
    const float *src = p.samples + gstart;
    float *dst = s_xbuf + b * xoff;
    bulk = ((gstart & 3) == 0) && count >= 4;
    if (bulk) {
      const int n4 = count >> 2;
      if (tid == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&s_full[b], (uint32_t)n4 * 16u);
        bulk_load(dst, src, (uint32_t)n4 * 16u, &s_full[b]);
      }
      for (int i = (n4 << 2) + tid; i < count; i += kF1kThreads) dst[i] = __ldg(src + i);
    } else {
      for (int i = tid; i < count; i += kF1kThreads) dst[i] = __ldg(src + i);
    }
  };

  long long tile = blockIdx.x;
  int buf = 0;
  long long row0 = 0, gs = 0;
  float pri = 0.f;
  int nf = 0;
  bool bulk = false;
  if (tid == 0) {
    if (tile < p.num_tiles) fetch_desc(tile, 0);
    if (tile + gridDim.x < p.num_tiles) fetch_desc(tile + gridDim.x, 1);
    cp_async_wait<0>();
  }
  __syncthreads();
  if (tile < p.num_tiles) stage(0, 0, row0, nf, bulk, gs, pri);
  for (; tile < p.num_tiles; buf ^= 1) {
    const long long next = tile + gridDim.x;
    long long n_row0 = 0, n_gs = 0;
    float n_pri = 0.f;
    int n_nf = 0;
    bool n_bulk = false;
    if (next < p.num_tiles) stage(buf ^ 1, buf ^ 1, n_row0, n_nf, n_bulk, n_gs, n_pri);
    const bool refill = tid == 0 && next + gridDim.x < p.num_tiles;
    if (refill) fetch_desc(next + gridDim.x, buf);
    if (bulk) {
      if (buf == 0) { mbar_wait(&s_full[0], ph0); ph0 ^= 1; }
      else { mbar_wait(&s_full[1], ph1); ph1 ^= 1; }
    }
    __syncthreads();  // samples of this tile visible; previous tile's phase C is finished
    const float *s_x = s_xbuf + buf * xoff;
    if (p.pre_cof != 0.f)  // pre-emphasis as a prologue of the framing: the staged tile, in place
      preemphasis_tile(s_xbuf + buf * xoff, (nf - 1) * S + L, pri, p.pre_cof, tid, kF1kThreads);

    // ---- phase B: one frame per warp and pass
    for (int pass = 0; pass < kTile / kF1kWarps; ++pass) {
      const int fl = pass * kF1kWarps + warp;
      if (fl >= nf) continue;  // warp-uniform
      const float *xf = s_x + fl * S;
      float2 v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        if (j < NZ) {
          const int o = 64 * j + 2 * lane;
          v[j] = c_make(o < Leff ? xf[o] : 0.f, o + 1 < Leff ? xf[o + 1] : 0.f);
        } else {
          v[j] = c_make(0.f, 0.f);
        }
      }
      float sumsq = 0.f;
      if (p.need_energy || p.need_level) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < NZ) sumsq = fmaf(v[j].x, v[j].x, fmaf(v[j].y, v[j].y, sumsq));
      }
      fft16<NZ>(v);  // over j -> k1
#pragma unroll
      for (int k1 = 1; k1 < 16; ++k1) v[k1] = c_mul(v[k1], s_tw1[k1 * 32 + lane]);
      __syncwarp();  // the previous pass's reads of my_exch are done
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) my_exch[k1 * kF1kRS + lane] = v[k1];
      __syncwarp();
#pragma unroll
      for (int a = 0; a < 16; ++a) v[a] = my_exch[k1r * kF1kRS + 2 * a + hb];
      fft16<16>(v);  // over a -> q
      if (hb) {
#pragma unroll
        for (int q = 1; q < 16; ++q) v[q] = c_mul(v[q], s_tw3[q]);
      }
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        float2 o;
        o.x = __shfl_xor_sync(0xffffffffu, v[q].x, 16);
        o.y = __shfl_xor_sync(0xffffffffu, v[q].y, 16);
        v[q] = hb ? c_sub(o, v[q]) : c_add(v[q], o);  // lane (k1, s): Z[k1 + 16 q + 256 s]
      }
      // untangle: X[k] = (A + W1024^k B) / 2, A = Z[k] + conj Z[512-k], B = -i (Z[k] - conj Z[512-k])
      const int src = ((16 - k1r) & 15) + 16;
      float2 prev = v[0];  // k1 == 0, q == 0: the partner of Z[0] is Z[512] = Z[0]
      float *prow = s_p + fl * PSTRIDE;
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        float2 r;
        r.x = __shfl_sync(0xffffffffu, v[15 - q].x, src);
        r.y = __shfl_sync(0xffffffffu, v[15 - q].y, src);
        const float2 zp = (k1r == 0) ? prev : r;  // lane k1 == 0 wants element 16 - q: last round's value
        prev = r;
        const float2 zk = v[q];
        const float2 A = c_make(zk.x + zp.x, zk.y - zp.y);
        const float2 B = c_make(zk.y + zp.y, zp.x - zk.x);
        const float2 WB = c_mul(B, s_tw2[q * 16 + k1r]);
        const float xr = A.x + WB.x, xi = A.y + WB.y;
        if (hb == 0) prow[k1r + 16 * q] = 0.25f * fmaf(xr, xr, xi * xi);
      }
      if (lane == 16) prow[256] = fmaf(v[0].x, v[0].x, v[0].y * v[0].y);  // Z[256] is its own partner
      if (p.need_energy || p.need_level) {
        // frame energy = sum of the one-sided power spectrum, via Parseval (feature.py:202)
        for (int o = 16; o >= 1; o >>= 1) sumsq += __shfl_xor_sync(0xffffffffu, sumsq, o);
        if (lane == 0 && p.need_level) s_lvl[fl] = sumsq;
        if (lane == 0 && p.need_energy) {
          const float x0 = v[0].x + v[0].y, xn = v[0].x - v[0].y;
          float e = (1024.f * sumsq + x0 * x0 + xn * xn) * (1.0f / 2048.f);
          if (e == 0.f) e = kEpsF64AsF32;  // feature.py:205
          s_en[fl] = e;
        }
      }
    }
    if (refill) cp_async_wait<0>();
    __syncthreads();
    // ---- phase C
    if (use_seg && p.mode == FEAT_LMFE)
      output_stage<kTile, true, FEAT_LMFE>(p, s_p, PSTRIDE, s_en, s_a, s_b, s_bw, s_bmeta, s_dct, nf, row0, s_rd,
                                           s_wr, s_U, s_D, fm, s_lvl, gs);
    else if (use_seg)
      output_stage<kTile, true>(p, s_p, PSTRIDE, s_en, s_a, s_b, s_bw, s_bmeta, s_dct, nf, row0, s_rd, s_wr,
                                s_U, s_D, fm, s_lvl, gs);
    else
      output_stage<kTile, false>(p, s_p, PSTRIDE, s_en, s_a, s_b, s_bw, s_bmeta, s_dct, nf, row0, s_rd, s_wr,
                                 s_U, s_D, fm, s_lvl, gs);
    tile = next;
    row0 = n_row0; nf = n_nf; bulk = n_bulk; gs = n_gs; pri = n_pri;
  }
}

// ====================================================================================
// Generic kernel: one warp per frame, radix-2 DIT on NFFT/2 complex points in smem.
// ====================================================================================
constexpr int kGenWarps = 4;
constexpr int kGenTile = kGenWarps;  // frames per CTA iteration

__global__ void __launch_bounds__(kGenWarps * 32)
features_generic(const FeatParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int nc = p.nfft >> 1;
  const int K = p.K;
  const int pstride = (K + 4) | 1;  // >= 4 zeroed pad bins after the last real bin
  float2 *s_z = reinterpret_cast<float2 *>(smem_raw);      // [kGenWarps][nc]
  float *s_p = reinterpret_cast<float *>(s_z + kGenWarps * nc);  // [kGenTile][pstride]
  float *s_bw = s_p + kGenTile * pstride;                  // 16 B aligned (kGenTile == 4)
  float *s_en = s_bw + p.band_total;
  float *s_lvl = s_en + kGenTile;
  float *s_a = s_lvl + kGenTile;
  float *s_b = s_a + kGenTile * (p.M + 1);
  float *s_dct = s_b + kGenTile * (p.C + 1);
  int *s_bmeta = reinterpret_cast<int *>(s_dct + p.C * p.M);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  load_plan_to_smem(p, s_bw, s_bmeta, s_dct);
  for (int i = threadIdx.x; i < kGenTile * (pstride - K); i += blockDim.x)
    s_p[(i / (pstride - K)) * pstride + K + i % (pstride - K)] = 0.f;
  const int Leff = min(p.L, p.nfft);
  float2 *z = s_z + warp * nc;

  // tiles here are the same 32-frame tiles as the fast kernel; walk them kGenTile frames at a time
  for (long long tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
    const int2 td = p.tiles[tile];
    const int utt = td.x;
    const int nf_tile = min(kTile, p.nframes[utt] - td.y);
    for (int fsub = 0; fsub < nf_tile; fsub += kGenTile) {
      const int f0 = td.y + fsub;
      const int nf = min(kGenTile, nf_tile - fsub);
      __syncthreads();
      if (warp < nf) {
        const float *x = p.samples + p.starts[utt] + (long long)(f0 + warp) * p.S;
        // pre-emphasis fused into the load: y[i] = x[i] - cof x[i-1]; sample 0 of the utterance wraps to its last
        const float cof = p.pre_cof;
        const float *xw = (cof != 0.f && f0 + warp == 0) ? p.samples + p.last[utt] : x - 1;
        auto ld = [&](const int i) {
          float v = __ldg(x + i);
          if (cof != 0.f) v = fmaf(-cof, __ldg(i > 0 ? x + i - 1 : xw), v);
          return v;
        };
        // bit-reversed load of z[m] = x[2m] + i x[2m+1]
        float sumsq = 0.f;
        for (int m = lane; m < nc; m += 32) {
          float re = (2 * m < Leff) ? ld(2 * m) : 0.f;
          float im = (2 * m + 1 < Leff) ? ld(2 * m + 1) : 0.f;
          sumsq = fmaf(re, re, fmaf(im, im, sumsq));
          const int r = __brev((unsigned)m) >> (32 - p.log2nc);
          z[r] = c_make(re, im);
        }
        __syncwarp();
        for (int s = 0; s < p.log2nc; ++s) {
          const int hlf = 1 << s;
          const int tw_step = (p.nfft >> 1) >> s;  // W_{2*hlf}^pos = tw[pos * nfft/(2*hlf)]
          for (int b = lane; b < (nc >> 1); b += 32) {
            const int pos = b & (hlf - 1);
            const int i = ((b >> s) << (s + 1)) + pos;
            const int j = i + hlf;
            const float2 w = p.tw[pos * tw_step];
            const float2 u = z[i];
            const float2 t = c_mul(z[j], w);
            z[i] = c_add(u, t);
            z[j] = c_sub(u, t);
          }
          __syncwarp();
        }
        // untangle: X[k] = E + W^k O ; bins 0..nc
        float *prow = s_p + warp * pstride;
        for (int k = lane; k <= (nc >> 1); k += 32) {
          const float2 zk = z[k];
          const float2 zp = z[(nc - k) & (nc - 1)];
          const float2 A = c_make(zk.x + zp.x, zk.y - zp.y);
          const float2 B = c_make(zk.y + zp.y, zp.x - zk.x);
          const float2 w = p.tw[k];
          const float2 WB = c_mul(B, w);
          const float xr = A.x + WB.x, xi = A.y + WB.y;
          const float yr = A.x - WB.x, yi = A.y - WB.y;
          prow[k] = 0.25f * fmaf(xr, xr, xi * xi);
          prow[nc - k] = 0.25f * fmaf(yr, yr, yi * yi);
        }
        if (p.need_energy || p.need_level) {
          for (int o = 16; o >= 1; o >>= 1) sumsq += __shfl_xor_sync(0xffffffffu, sumsq, o);
          if (lane == 0 && p.need_level) s_lvl[warp] = sumsq;
          if (lane == 0 && p.need_energy) {
            const float x0 = z[0].x + z[0].y, xn = z[0].x - z[0].y;
            const float nf_ = (float)p.nfft;
            float e = (nf_ * sumsq + x0 * x0 + xn * xn) / (2.f * nf_);
            if (e == 0.f) e = kEpsF64AsF32;
            s_en[warp] = e;
          }
        }
      }
      __syncthreads();
      const long long row0 = p.frame_offsets[utt] + f0;
      output_stage<kGenTile, false>(p, s_p, pstride, s_en, s_a, s_b, s_bw, s_bmeta, s_dct, nf, row0, nullptr,
                                    nullptr, nullptr, nullptr, FmIndex{0, 0, 0, 0}, s_lvl,
                                    p.starts[utt] + (long long)f0 * p.S);
    }
  }
}

// ------------------------------------------------------------------------------------
static FeatParams make_params(const FeatArgs &a, const FftTables &t) {
  FeatParams p;
  p.samples = a.samples;
  p.starts = a.starts;
  p.nframes = a.nframes;
  p.frame_offsets = a.frame_offsets;
  p.tiles = a.tiles;
  p.ftiles = a.ftiles;
  p.run_if = a.run_if;
  p.pre_cof = (float)a.pre_cof;
  p.last = a.last;
  p.num_tiles = a.num_tiles;
  p.L = a.L; p.S = a.S; p.K = a.K;
  p.M = a.M; p.C = a.C; p.dc_elim = a.dc_elim;
  p.band_lo = a.band_lo; p.band_n = a.band_n; p.band_off = a.band_off;
  p.band_w = a.band_w; p.band_total = a.band_total;
  p.dct = a.dct;
  p.seg_rd = a.seg_rd;
  p.seg_ranges = a.seg_ranges;
  p.seg_kmin = a.seg_kmin;
  p.seg_nbins = a.seg_nbins;
  p.mode = a.mode;
  p.out = a.out;
  p.energy = a.energy;
  p.twA = t.d_twA; p.twB = t.d_twB; p.tw = t.d_tw;
  p.nfft = a.nfft;
  int lg = 0;
  while ((1 << lg) < (a.nfft >> 1)) ++lg;
  p.log2nc = lg;
  p.need_energy = (a.mode == FEAT_MFE && a.energy != nullptr) ||
                  (a.mode >= FEAT_MFCC && a.dc_elim);
  p.redo = a.redo_fft;
  p.need_level = (a.redo_fft.entries != nullptr && a.mode >= FEAT_MFE) ? 1 : 0;
  return p;
}

static size_t plan_smem_bytes(const FeatArgs &a, int tile) {
  size_t floats = 2 * (size_t)tile                               // s_en, s_lvl
                  + (size_t)tile * (a.M + 1) + (size_t)tile * (a.C + 1)
                  + (size_t)a.band_total + (size_t)a.C * a.M + 3 * (size_t)a.M;
  return floats * 4;
}

bool fast512_supported(const FeatArgs &a) {
  if (a.nfft != 512) return false;
  if (a.tile_frames != kTile) return false;
  if (a.L < 1 || a.S < 1) return false;
  size_t span = (size_t)(kTile - 1) * a.S + a.L;
  if (span * 4 > 44 * 1024) return false;  // long strides / frames: the generic kernel takes it
  return true;
}

template <int NJ, bool FULL>
static int launch_fast(spb_context *ctx, const FeatParams &p, const FeatArgs &a) {
  const int span_max = (kTile - 1) * a.S + a.L;
  size_t smem = 2 * (size_t)((span_max + 3) & ~3) * 4 + 2 * kNW * 256 * sizeof(float2) +
                (size_t)kTile * ((FULL ? 257 : 129) + 4) * 4 + plan_smem_bytes(a, kTile) +
                16 + (size_t)a.seg_nbins * 16 + 8 * 16 + (256 + 128) * sizeof(float2);
  auto kern = features_fast512<NJ, FULL>;
  int per_sm = 0;
  int rc = configure_kernel(ctx, (const void *)kern, kFastThreads, smem, &per_sm);
  if (rc) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_INVALID, "fast512: kernel does not fit on an SM");
  long long grid = (long long)ctx->sm_count * per_sm;
  if (grid > a.num_tiles) grid = a.num_tiles;
  if (grid < 1) return SPB_OK;
  kern<<<(unsigned)grid, kFastThreads, smem, ctx->stream>>>(p);
  SPB_LAUNCH_CHECK(ctx, "features_fast512");
  return SPB_OK;
}

int launch_features_fast512(spb_context *ctx, const FeatArgs &a, const FftTables &t) {
  FeatParams p = make_params(a, t);
  // FULL spectrum is needed for power/magnitude output; feature modes only read the
  // bins the filterbank touches (speechpy's (K+1) bin mapping never exceeds bin 128,
  // a custom bank that does selects the FULL variant).
  const bool full = (a.mode == FEAT_POWER || a.mode == FEAT_MAG) || a.max_bin > 128;
  const FeatArgs &aa = a;
  const bool pruned320 = (a.L == 320 && (a.S % 2 == 0));
  if (full) {
    return pruned320 ? launch_fast<10, true>(ctx, p, aa) : launch_fast<16, true>(ctx, p, aa);
  }
  return pruned320 ? launch_fast<10, false>(ctx, p, aa) : launch_fast<16, false>(ctx, p, aa);
}

bool fast1024_supported(const FeatArgs &a) {
  if (a.nfft != 1024 || a.tile_frames != kTile) return false;
  if (a.mode < FEAT_MFE || a.max_bin > 256) return false;  // spectrum outputs need all 513 bins: generic kernel
  if (a.L < 1 || a.S < 1) return false;
  size_t span = (size_t)(kTile - 1) * a.S + a.L;
  if (span * 4 > 44 * 1024) return false;
  return true;
}

int launch_features_fast1024(spb_context *ctx, const FeatArgs &a, const FftTables &t) {
  FeatParams p = make_params(a, t);
  const int span_max = (kTile - 1) * a.S + a.L;
  size_t smem = 2 * (size_t)((span_max + 3) & ~3) * 4 + (size_t)kF1kWarps * 16 * kF1kRS * sizeof(float2) +
                (size_t)kTile * (257 + 4) * 4 + plan_smem_bytes(a, kTile) + 16 + (size_t)a.seg_nbins * 16 + 8 * 16 +
                (512 + 256 + 16) * sizeof(float2);
  auto kern = (a.L <= 448) ? features_fast1024<7> : features_fast1024<16>;
  int per_sm = 0;
  int rc = configure_kernel(ctx, (const void *)kern, kF1kThreads, smem, &per_sm);
  if (rc) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_INVALID, "fast1024: kernel does not fit on an SM");
  long long grid = (long long)ctx->sm_count * per_sm;
  if (grid > a.num_tiles) grid = a.num_tiles;
  if (grid < 1) return SPB_OK;
  kern<<<(unsigned)grid, kF1kThreads, smem, ctx->stream>>>(p);
  SPB_LAUNCH_CHECK(ctx, "features_fast1024");
  return SPB_OK;
}

int launch_features_generic(spb_context *ctx, const FeatArgs &a, const FftTables &t) {
  FeatParams p = make_params(a, t);
  const FeatArgs &aa = a;
  const int nc = a.nfft >> 1;
  size_t smem = (size_t)kGenWarps * nc * sizeof(float2) + (size_t)kGenTile * ((a.K + 4) | 1) * 4 +
                plan_smem_bytes(aa, kGenTile);
  int per_sm = 0;
  int rc = configure_kernel(ctx, (const void *)features_generic, kGenWarps * 32, smem, &per_sm);
  if (rc) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_INVALID, "generic fft kernel does not fit on an SM");
  long long grid = (long long)ctx->sm_count * per_sm;
  if (grid > a.num_tiles) grid = a.num_tiles;
  if (grid < 1) return SPB_OK;
  features_generic<<<(unsigned)grid, kGenWarps * 32, smem, ctx->stream>>>(p);
  SPB_LAUNCH_CHECK(ctx, "features_generic");
  return SPB_OK;
}

}  // namespace spb
