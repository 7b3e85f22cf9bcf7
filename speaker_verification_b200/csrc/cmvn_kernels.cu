// K3: cmvnw (sliding-window mean / variance normalisation) and cmvn (global).
//
// cmvnw replaces speech_feature_extraction/speechpy/processing.py:301-322: for every
// row i, mean over the symmetric-padded window [i-h, i+h] (h = (win-1)/2), float32
// store of x - mean (processing.py:302,307), then population std of the float32
// mean-subtracted values over the same window shape and out = ms / (std + 2^-30)
// (processing.py:314-322).  The reference spends 2F numpy reductions of win x cols
// on this; here each stage is one pass with float64 window accumulators:
//
//   tile = R output rows of one utterance x all columns, staged in shared memory
//   together with the +-h halo (symmetric reflection t -> t mod 2F, mirrored above F,
//   which equals np.pad(..., 'symmetric') for any F, including F < h);
//   two-level window sums: float64 sums of G-row blocks, a thread sums win/G block
//   sums (+ win%G rows) for its first row and then slides: W += x[q+win] - x[q].
//
// Stage 1 writes ms (float32) to a scratch array; stage 2 reads it back (L2-resident
// for chunked batches) and needs the same halo of ms, which is why the two stages are
// separate launches rather than one kernel with a 2h halo (that would recompute stage
// 1 for (R+2h)/R times the rows).
#include <algorithm>

#include "ptx_helpers.cuh"
#include "spb_internal.cuh"

namespace spb {

constexpr int kCmG = 15;          // rows per block sum
constexpr int kCmThreads = 320;

// 32-bit variant with the common single-reflection cases first
__device__ __forceinline__ int reflect_row32(int t, int F) {
  if (t >= 0 && t < F) return t;
  if (t < 0 && t >= -F) return -1 - t;
  if (t >= F && t < 2 * F) return 2 * F - 1 - t;
  const int period = 2 * F;
  int m = t % period;
  if (m < 0) m += period;
  return (m < F) ? m : (period - 1 - m);
}

__device__ __forceinline__ int reflect_row(long long t, int F) {
  // np.pad(..., 'symmetric'): ... x1 x0 | x0 x1 ... x(F-1) | x(F-1) x(F-2) ...
  const int period = 2 * F;
  int m = (int)(t % period);
  if (m < 0) m += period;
  return (m < F) ? m : (period - 1 - m);
}

struct CmParams {
  const float *in;
  float *out;
  const int32_t *nframes;
  const int64_t *frame_offsets;
  const int2 *tiles;
  long long num_tiles;
  int R, cols, win;
  int ld;  // row stride (floats) of in / out; cols <= ld selects a column slice
  float eps;
};

// Vector helpers: VEC columns per thread (1 or 2).
template <int VEC> struct CmVec;
template <> struct CmVec<1> {
  static __device__ __forceinline__ void load(const float *p, float (&v)[1]) { v[0] = *p; }
  static __device__ __forceinline__ void store(float *p, const float (&v)[1]) { *p = v[0]; }
};
template <> struct CmVec<2> {
  static __device__ __forceinline__ void load(const float *p, float (&v)[2]) {
    const float2 t = *reinterpret_cast<const float2 *>(p);
    v[0] = t.x; v[1] = t.y;
  }
  static __device__ __forceinline__ void store(float *p, const float (&v)[2]) {
    *reinterpret_cast<float2 *>(p) = make_float2(v[0], v[1]);
  }
};

// Thread layout: column group g = tid % ncg (VEC adjacent columns), row block b = tid / ncg;
// every index below is derived from (g, b) with adds only -- no per-element division.
template <int STAGE, int VEC>
__global__ void __launch_bounds__(kCmThreads)
cmvnw_kernel(const CmParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int cols = p.cols, win = p.win, R = p.R, ld = p.ld;
  const int h = (win - 1) >> 1;
  const int rows_ext_max = R + win - 1;
  const int nb = win / kCmG, rem = win - nb * kCmG;
  const int nblk_max = R / kCmG + nb;  // R is a multiple of kCmG
  float *s = reinterpret_cast<float *>(smem_raw);                       // [rows_ext_max][cols]
  size_t s_bytes = ((size_t)rows_ext_max * cols * 4 + 15) & ~(size_t)15;
  double *bs1 = reinterpret_cast<double *>(smem_raw + s_bytes);         // [nblk_max][cols]
  double *bs2 = bs1 + (size_t)nblk_max * cols;                          // stage 2 only
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  constexpr int kWarps = kCmThreads / 32;
  const int ncg = cols / VEC;                 // column groups
  const int nbpt = kCmThreads / ncg;          // row blocks handled per sweep
  const int g = tid % ncg, b0 = tid / ncg;    // once per thread
  const bool worker = b0 < nbpt;
  const int c0 = g * VEC;
  const double inv_win = 1.0 / (double)win;
  const bool vec4 = ((cols & 3) == 0) && ((ld & 3) == 0) &&
                    ((((size_t)p.in) & 15) == 0);
  const int n4 = cols >> 2;

  for (long long tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
    const int2 td = p.tiles[tile];
    const int utt = td.x, r0 = td.y;
    const int F = p.nframes[utt];
    const long long base = p.frame_offsets[utt];
    const int nrows = min(R, F - r0);
    const int rows_needed = nrows + win - 1;
    __syncthreads();
    // ---- stage rows [r0-h, r0+nrows+h), symmetric reflection at the utterance ends
    const int first = r0 - h;
    const bool interior = first >= 0 && first + rows_needed <= F;
    if (interior && ld == cols && vec4) {
      const float4 *s4 = reinterpret_cast<const float4 *>(p.in + (base + first) * cols);
      float4 *d4 = reinterpret_cast<float4 *>(s);
      const int total4 = (rows_needed * cols) >> 2;
      for (int i = tid; i < total4; i += kCmThreads) d4[i] = __ldg(s4 + i);
    } else if (vec4 && n4 <= 32) {
      // rows_per_pass rows per warp iteration, one float4 per lane
      const int rpp = 32 / n4;
      const int lr = lane / n4, lc = lane - lr * n4;
      const int period = 2 * F;
      if (lr < rpp) {
        for (int t = warp * rpp + lr; t < rows_needed; t += kWarps * rpp) {
          int m = (first + t) % period;
          if (m < 0) m += period;
          const int srow = (m < F) ? m : (period - 1 - m);
          const float4 v = __ldg(reinterpret_cast<const float4 *>(p.in + (base + srow) * ld) + lc);
          reinterpret_cast<float4 *>(s + (size_t)t * cols)[lc] = v;
        }
      }
    } else {
      const int period = 2 * F;
      for (int t = warp; t < rows_needed; t += kWarps) {
        int m = (first + t) % period;
        if (m < 0) m += period;
        const int srow = (m < F) ? m : (period - 1 - m);
        const float *src = p.in + (base + srow) * ld;
        for (int c = lane; c < cols; c += 32) s[(size_t)t * cols + c] = __ldg(src + c);
      }
    }
    __syncthreads();
    // ---- float64 sums of kCmG-row blocks
    const int nblk_out = (nrows + kCmG - 1) / kCmG;
    const int nblk = nblk_out + nb - 1;  // every block lies inside rows_needed
    if (worker) {
      for (int k = b0; k < nblk; k += nbpt) {
        const float *col = s + (size_t)k * kCmG * cols + c0;
        double a1[VEC], a2[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) a1[v] = a2[v] = 0.0;
#pragma unroll
        for (int r = 0; r < kCmG; ++r) {
          float x[VEC];
          CmVec<VEC>::load(col + r * cols, x);
#pragma unroll
          for (int v = 0; v < VEC; ++v) {
            const double d = (double)x[v];
            a1[v] += d;
            if (STAGE == 2) a2[v] = fma(d, d, a2[v]);
          }
        }
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
          bs1[k * cols + c0 + v] = a1[v];
          if (STAGE == 2) bs2[k * cols + c0 + v] = a2[v];
        }
      }
    }
    __syncthreads();
    // ---- window sums: first row of each block from block sums, then slide down the block
    if (worker) {
      for (int b = b0; b < nblk_out; b += nbpt) {
        double w1[VEC], w2[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) w1[v] = w2[v] = 0.0;
        for (int j = 0; j < nb; ++j) {
#pragma unroll
          for (int v = 0; v < VEC; ++v) {
            w1[v] += bs1[(b + j) * cols + c0 + v];
            if (STAGE == 2) w2[v] += bs2[(b + j) * cols + c0 + v];
          }
        }
        const float *col = s + c0;
        const int tail0 = (b + nb) * kCmG;
        for (int r = 0; r < rem; ++r) {
          float x[VEC];
          CmVec<VEC>::load(col + (size_t)(tail0 + r) * cols, x);
#pragma unroll
          for (int v = 0; v < VEC; ++v) {
            const double d = (double)x[v];
            w1[v] += d;
            if (STAGE == 2) w2[v] = fma(d, d, w2[v]);
          }
        }
        const int q0 = b * kCmG;
        const int q1 = min(q0 + kCmG, nrows);
        float *dst = p.out + (base + r0 + q0) * ld + c0;
        const float *pc = col + (size_t)(q0 + h) * cols;   // centre row
        const float *pin = col + (size_t)(q0 + win) * cols;  // row entering the window
        const float *pout = col + (size_t)q0 * cols;         // row leaving the window
        for (int q = q0; q < q1; ++q) {
          float xc[VEC], o[VEC];
          CmVec<VEC>::load(pc, xc);
#pragma unroll
          for (int v = 0; v < VEC; ++v) {
            if (STAGE == 1) {
              // processing.py:304-307: float32(vec[i] - mean(window))
              o[v] = (float)((double)xc[v] - w1[v] * inv_win);
            } else {
              // processing.py:318-322: ms / (std(window) + eps), population std
              const double mean = w1[v] * inv_win;
              double var = fma(-mean, mean, w2[v] * inv_win);
              var = var < 0.0 ? 0.0 : var;  // select, not fmax: a NaN variance must stay NaN
              o[v] = xc[v] / (sqrtf((float)var) + p.eps);
            }
          }
          CmVec<VEC>::store(dst, o);
          if (q + 1 < q1) {
            float xi[VEC], xo[VEC];
            CmVec<VEC>::load(pin, xi);
            CmVec<VEC>::load(pout, xo);
#pragma unroll
            for (int v = 0; v < VEC; ++v) {
              const double vin = (double)xi[v], vout = (double)xo[v];
              w1[v] += vin - vout;
              if (STAGE == 2) w2[v] += fma(vin, vin, -vout * vout);
            }
          }
          dst += ld; pc += cols; pin += cols; pout += cols;
        }
      }
    }
  }
}

static size_t cm_smem_bytes(int R, int cols, int win) {
  const int rows_ext = R + win - 1;
  const int nblk = R / kCmG + win / kCmG;
  size_t s_bytes = ((size_t)rows_ext * cols * 4 + 15) & ~(size_t)15;
  return s_bytes + 2 * (size_t)nblk * cols * 8;
}

// Rows per tile for (cols, win): multiple of kCmG, as large as fits two CTAs per SM.
int cmvnw_tile_rows(int cols, int win) {
  int best = kCmG;
  for (int r = kCmG; r <= 16 * kCmG; r += kCmG) {
    if (cm_smem_bytes(r, cols, win) <= 112 * 1024) best = r;
  }
  return best;
}

bool cmvnw_supported(int cols, int win) {
  return cols <= kCmThreads && cm_smem_bytes(kCmG, cols, win) <= 200 * 1024;
}

int launch_cmvnw_f32(spb_context *ctx, const float *in, float *scratch_ms, float *out,
                     const int32_t *nframes, const int64_t *frame_offsets, const int2 *tiles,
                     int64_t num_tiles, int tile_rows, int cols, int ld, int win, int varnorm, int first_stage) {
  if (num_tiles <= 0) return SPB_OK;
  CmParams p;
  p.nframes = nframes;
  p.frame_offsets = frame_offsets;
  p.tiles = tiles;
  p.num_tiles = num_tiles;
  p.R = tile_rows;
  p.cols = cols;
  p.ld = ld;
  p.win = win;
  p.eps = 9.313225746154785e-10f;  // 2^-30 (processing.py:292)
  const size_t smem = cm_smem_bytes(tile_rows, cols, win);
  for (int stage = first_stage; stage <= (varnorm ? 2 : 1); ++stage) {
    // two columns per thread need 8-byte aligned rows in global and shared memory
    const bool v2 = (cols % 2 == 0) && (ld % 2 == 0) && ((((size_t)in) & 7) == 0) &&
                    ((((size_t)out) & 7) == 0) && (!scratch_ms || (((size_t)scratch_ms) & 7) == 0);
    auto kern = (stage == 1) ? (v2 ? cmvnw_kernel<1, 2> : cmvnw_kernel<1, 1>)
                             : (v2 ? cmvnw_kernel<2, 2> : cmvnw_kernel<2, 1>);
    int per_sm = 0;
    int rc = configure_kernel(ctx, (const void *)kern, kCmThreads, smem, &per_sm);
    if (rc) return rc;
    if (per_sm < 1) return fail(ctx, SPB_ERR_UNSUPPORTED, "cmvnw: window x columns too large for shared memory");
    long long grid = (long long)ctx->sm_count * per_sm;
    if (grid > num_tiles) grid = num_tiles;
    p.in = (stage == 1) ? in : scratch_ms;
    p.out = (stage == 1) ? (varnorm ? scratch_ms : out) : out;
    const size_t tb = timing_begin(ctx);
    kern<<<(unsigned)grid, kCmThreads, smem, ctx->stream>>>(p);
    timing_end(ctx, stage, tb);
    SPB_LAUNCH_CHECK(ctx, "cmvnw_kernel");
  }
  return SPB_OK;
}

// ====================================================================================
// cmvnw, large-tile variant (the batch hot path).
//
// Same arithmetic as cmvnw_kernel above, restructured for instruction count:
//   * one CTA per SM, 640 threads, tiles of up to ~960 rows (+ win-1 halo rows) so that most
//     utterances are a single tile whose halo is pure reflection (halo factor ~1.3, was 1.9);
//     tiles are balanced per utterance and listed longest-first, CTAs pull them from an
//     atomic counter;
//   * interior rows staged with flat 16-byte loads, only the <= win-1 halo rows pay the
//     reflection arithmetic;
//   * 30-row block sums (win = 10 blocks + 1 row at 301), column pair per thread, COLS as a
//     template constant so every shared-memory offset is an immediate; the slide loop is
//     unrolled by two;
//   * stage 2 uses sqrt.approx / rcp.approx (2 ulp; the reference computes np.std in float32).
// ====================================================================================
constexpr int kBigG = 30;
constexpr int kBigThreads = 640;

struct CmBigParams {
  const float *in;
  float *out;
  const int32_t *nframes;
  const int64_t *frame_offsets;
  const int4 *tiles;   // {utt, first row, rows, 0}
  int num_tiles;
  int *counter;        // dynamic tile scheduler
  int cols, win, ld, rmax;
  float eps;
};

// float -> double with integer instructions: exact for normal numbers; zero and denormals come
// out as +-2^-127-sized values (far below anything these sums resolve); inf / nan stay inf / nan.  F2F.F64.F32 runs on the 16-lane XU pipe, which bounds the
// block-sum phase (2 conversions x 640 threads per row = 80 clocks); doing every other
// conversion on the integer pipe lets the two pipes work in parallel.
__device__ __forceinline__ double f2d_alu(float f) {
  const unsigned u = __float_as_uint(f);
  const unsigned a = u & 0x7fffffffu;
  unsigned hi = ((a >> 3) + 0x38000000u) | (u & 0x80000000u);
  if (a >= 0x7f800000u) hi |= 0x7ff00000u;  // inf / nan keep their class (one predicated LOP3)
  return __hiloint2double((int)hi, (int)(u << 29));
}

__device__ __forceinline__ float2 big_neg(const float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 big_sub(const float2 a, const float2 b) { return __fadd2_rn(a, big_neg(b)); }

template <int STAGE, int COLS>
__global__ void __launch_bounds__(kBigThreads, 1)
cmvnw_big_kernel(const CmBigParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int cols = COLS ? COLS : p.cols;
  const int win = p.win, ld = p.ld;
  const int h = (win - 1) >> 1;
  const int rows_ext_max = p.rmax + win - 1;
  const int nb = win / kBigG, rem = win - nb * kBigG;
  const int nblk_max = (rows_ext_max + kBigG - 1) / kBigG + 1;
  float *s = reinterpret_cast<float *>(smem_raw);
  const size_t s_bytes = ((size_t)rows_ext_max * cols * 4 + 15) & ~(size_t)15;
  double *bs1 = reinterpret_cast<double *>(smem_raw + s_bytes);
  double *bs2 = bs1 + (size_t)nblk_max * cols;
  __shared__ int4 s_desc[2];
  __shared__ long long s_base[2];
  __shared__ __align__(8) uint64_t s_bar;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  constexpr int kWarps = kBigThreads / 32;
  if (tid == 0) {
    mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  uint32_t bar_phase = 0;
  const int ncg = cols >> 1;
  const int nbpt = kBigThreads / ncg;
  // (column pair g, block-thread b0).  40 columns: a row is 160 B = 20 pairs, 32 consecutive pairs straddle two rows
  // 30 apart and 30 x 160 B = 64 (mod 128), so the plain tid % 20 layout paid a bank-conflict wavefront on every
  // other 64-bit request (ncu r02ar, resident kernel: 34 % of the load wavefronts); instead warps 0 .. 15 take pairs
  // 0 .. 15 of two blocks each (a half-warp = 128 contiguous bytes), warps 16 .. 19 the pairs 16 .. 19 of eight blocks
  const bool remap = COLS == 40 && kBigThreads == 640;
  const int g = remap ? (tid < 512 ? (tid & 15) : 16 + ((tid - 512) & 3)) : tid % ncg;
  const int b0 = remap ? (tid < 512 ? (tid >> 4) : ((tid - 512) >> 2)) : tid / ncg;
  const bool worker = b0 < nbpt;
  const int c0 = 2 * g;
  const double inv_win = 1.0 / (double)win;
  const int n4 = cols >> 2;  // cols % 4 == 0 (checked by the launcher)

  // Tile descriptors are fetched one tile ahead by a single thread (atomic ticket + three dependent
  // global loads, ~2 us of latency) while the CTA works on the current tile.
  auto fetch_desc = [&](int slot) {
    const int t = atomicAdd(p.counter, 1);
    int4 d = make_int4(-1, 0, 0, 0);
    long long b = 0;
    if (t < p.num_tiles) {
      const int4 td = p.tiles[t];
      d = make_int4(td.x, td.y, td.z, p.nframes[td.x]);
      b = p.frame_offsets[td.x];
    }
    s_desc[slot] = d;
    s_base[slot] = b;
    if (d.x >= 0 && ld == cols) {
      // pull the next tile's rows (interior and non-reflected halo) into L2 while this tile is
      // being processed: its bulk copies then stream from L2 instead of waiting on HBM
      const int h2 = (win - 1) >> 1;
      const int lo = max(d.y - h2, 0), hi = min(d.y + d.z + h2, d.w);
      const char *src = reinterpret_cast<const char *>(p.in + (b + lo) * (long long)cols);
      const uint32_t total = (uint32_t)(hi - lo) * (uint32_t)cols * 4u;
      for (uint32_t off = 0; off < total; off += 65536u) bulk_prefetch_l2(src + off, min(65536u, total - off));
    }
  };
  if (tid == 0) fetch_desc(0);
  for (int slot = 0;; slot ^= 1) {
    __syncthreads();  // previous tile fully consumed; descriptor `slot` published
    const int4 td = s_desc[slot];
    if (td.x < 0) break;
    if (tid == kBigThreads - 1) fetch_desc(slot ^ 1);  // overlaps with this tile's work
    const int utt = td.x, r0 = td.y, nrows = td.z;
    const int F = td.w;
    const long long base = s_base[slot];
    (void)utt;
    const int rows_needed = nrows + win - 1;
    const int first = r0 - h;
    // ---- stage: ext row e <-> utterance row first + e; interior = [e_lo, e_hi)
    const int e_lo = max(0, -first), e_hi = min(rows_needed, F - first);
    if (e_hi > e_lo) {
      if (ld == cols) {
        // interior rows are one contiguous block in global memory: TMA bulk copies issued by
        // one thread (<= 64 KB each), completion on an mbarrier; the other threads go straight
        // to the halo rows
        if (tid == 0) {
          const char *src = reinterpret_cast<const char *>(p.in + (base + first + e_lo) * (long long)cols);
          char *dst = reinterpret_cast<char *>(s + (size_t)e_lo * cols);
          const uint32_t total = (uint32_t)(e_hi - e_lo) * (uint32_t)cols * 4u;
          // order earlier generic-proxy accesses of this buffer before the async-proxy writes
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          mbar_expect_tx(&s_bar, total);
          for (uint32_t off = 0; off < total; off += 65536u)
            bulk_load(dst + off, src + off, min(65536u, total - off), &s_bar);
        }
      } else {
        for (int i = tid; i < (e_hi - e_lo) * n4; i += kBigThreads) {
          const int r = i / n4, c = i - r * n4;
          reinterpret_cast<float4 *>(s + (size_t)(e_lo + r) * cols)[c] =
              __ldg(reinterpret_cast<const float4 *>(p.in + (base + first + e_lo + r) * (long long)ld) + c);
        }
      }
    }
    {
      // halo rows: [0, e_lo) and [e_hi, rows_needed), symmetric reflection (np.pad 'symmetric')
      const int nh_lo = max(e_lo, 0), nh = nh_lo + max(rows_needed - max(e_hi, e_lo), 0);
      const int rpp = 32 / n4 > 0 ? 32 / n4 : 1;
      const int lr = lane / n4, lc = lane - lr * n4;
      if (n4 <= 32 && lr < rpp) {
        // batches of six rows per thread: all loads of a batch are issued before the first store,
        // so their L2 latencies overlap instead of adding up
        constexpr int kHB = 6;
        for (int k0 = warp * rpp + lr; k0 < nh; k0 += kHB * kWarps * rpp) {
          float4 v[kHB];
          int ee[kHB];
#pragma unroll
          for (int i = 0; i < kHB; ++i) {
            const int k = k0 + i * kWarps * rpp;
            ee[i] = -1;
            if (k < nh) {
              const int e = (k < nh_lo) ? k : max(e_hi, e_lo) + (k - nh_lo);
              const int srow = reflect_row32(first + e, F);
              ee[i] = e;
              v[i] = __ldg(reinterpret_cast<const float4 *>(p.in + (base + srow) * (long long)ld) + lc);
            }
          }
#pragma unroll
          for (int i = 0; i < kHB; ++i)
            if (ee[i] >= 0) reinterpret_cast<float4 *>(s + (size_t)ee[i] * cols)[lc] = v[i];
        }
      } else if (n4 > 32) {
        for (int k = warp; k < nh; k += kWarps) {
          const int e = (k < nh_lo) ? k : max(e_hi, e_lo) + (k - nh_lo);
          const int srow = reflect_row32(first + e, F);
          for (int c = lane; c < n4; c += 32)
            reinterpret_cast<float4 *>(s + (size_t)e * cols)[c] =
                __ldg(reinterpret_cast<const float4 *>(p.in + (base + srow) * (long long)ld) + c);
        }
      }
    }
    if (ld == cols && e_hi > e_lo) {
      mbar_wait(&s_bar, bar_phase);  // bulk copy landed (async-proxy writes visible after the wait)
      bar_phase ^= 1;
    }
    __syncthreads();
    // ---- float64 sums of 30-row blocks
    const int nblk_out = (nrows + kBigG - 1) / kBigG;
    const int nblk = nblk_out + nb - 1;
    if (worker) {
      for (int k = b0; k < nblk; k += nbpt) {
        const float *col = s + (size_t)k * kBigG * cols + c0;
        double a1x = 0.0, a1y = 0.0, a2x = 0.0, a2y = 0.0;
#pragma unroll 10
        for (int r = 0; r < kBigG; ++r) {
          const float2 x = *reinterpret_cast<const float2 *>(col + r * cols);
          const double dx = (double)x.x, dy = f2d_alu(x.y);  // one conversion per pipe (XU / integer)
          a1x += dx; a1y += dy;
          if (STAGE == 2) { a2x = fma(dx, dx, a2x); a2y = fma(dy, dy, a2y); }
        }
        *reinterpret_cast<double2 *>(bs1 + k * cols + c0) = make_double2(a1x, a1y);
        if (STAGE == 2) *reinterpret_cast<double2 *>(bs2 + k * cols + c0) = make_double2(a2x, a2y);
      }
    }
    __syncthreads();
    // ---- window sums: first row of each block from block sums, then slide
    if (worker) {
      for (int b = b0; b < nblk_out; b += nbpt) {
        double w1x = 0.0, w1y = 0.0, w2x = 0.0, w2y = 0.0;
        for (int j = 0; j < nb; ++j) {
          const double2 t1 = *reinterpret_cast<const double2 *>(bs1 + (b + j) * cols + c0);
          w1x += t1.x; w1y += t1.y;
          if (STAGE == 2) {
            const double2 t2 = *reinterpret_cast<const double2 *>(bs2 + (b + j) * cols + c0);
            w2x += t2.x; w2y += t2.y;
          }
        }
        const float *col = s + c0;
        const int tail0 = (b + nb) * kBigG;
        for (int r = 0; r < rem; ++r) {
          const float2 x = *reinterpret_cast<const float2 *>(col + (size_t)(tail0 + r) * cols);
          const double dx = (double)x.x, dy = (double)x.y;
          w1x += dx; w1y += dy;
          if (STAGE == 2) { w2x = fma(dx, dx, w2x); w2y = fma(dy, dy, w2y); }
        }
        const int q0 = b * kBigG;
        const int q1 = min(q0 + kBigG, nrows);
        float *dst = p.out + (base + r0 + q0) * (long long)ld + c0;
        const float *pc = col + (size_t)(q0 + h) * cols;     // centre row
        const float *pin = col + (size_t)(q0 + win) * cols;  // row entering the window
        const float *pout = col + (size_t)q0 * cols;         // row leaving the window
        // The float64 window sums anchor the first row of the block; inside the block
        // (<= 29 steps) the window moves by float32 deltas D = sum(x_in - x_out), which costs no
        // conversions.  Error of D: <= 30 * 2^-24 * |D| on the SUM, i.e. ~1e-7 * |x| on the mean.
        // mean = mh + (ml + D/win) with mh + ml the float64 mean split in two floats, so
        // x - mean is formed as (x - mh) - (ml + D/win): exact to 1 ulp of the RESULT even when
        // |mean| >> |x - mean| (processing.py:304-307 subtracts in float64).
        const float inv_w = (float)inv_win;
        const double m0x = w1x * inv_win, m0y = w1y * inv_win;
        const float mhx = (float)m0x, mhy = (float)m0y;
        const float mlx = (float)(m0x - (double)mhx), mly = (float)(m0y - (double)mhy);
        // packed fp32 (FADD2 / FFMA2 / FMUL2: one instruction per column pair)
        const float2 iw = make_float2(inv_w, inv_w);
        const float2 mh = make_float2(mhx, mhy), ml = make_float2(mlx, mly);
        float2 D = make_float2(0.f, 0.f);
        if (STAGE == 1) {
#pragma unroll 2
          for (int q = q0; q < q1; ++q) {
            const float2 xc = *reinterpret_cast<const float2 *>(pc);
            const float2 xi = *reinterpret_cast<const float2 *>(pin);
            const float2 xo = *reinterpret_cast<const float2 *>(pout);
            *reinterpret_cast<float2 *>(dst) = big_sub(big_sub(xc, mh), __ffma2_rn(D, iw, ml));
            D = __fadd2_rn(D, big_sub(xi, xo));
            dst += ld; pc += cols; pin += cols; pout += cols;
          }
        } else {
          // processing.py:318-322: ms / (std(window) + eps), population std; the input is the
          // mean-subtracted signal, so E[x^2] >> mean^2 and float32 is enough for the variance.
          // Branch-free: x / (s + eps) = x r / (1 + eps r), r = rsqrt(v); v is clamped from below with a
          // select (v <= 0 for constant windows must give x / eps; fmaxf would swallow a NaN variance,
          // which has to poison the row like the reference)
          const float2 e0 = make_float2((float)(w2x * inv_win), (float)(w2y * inv_win));
          const float2 neps = make_float2(p.eps, p.eps), one = make_float2(1.f, 1.f);
          float2 Q = make_float2(0.f, 0.f);
#pragma unroll 2
          for (int q = q0; q < q1; ++q) {
            const float2 xc = *reinterpret_cast<const float2 *>(pc);
            const float2 xi = *reinterpret_cast<const float2 *>(pin);
            const float2 xo = *reinterpret_cast<const float2 *>(pout);
            const float2 m = __fadd2_rn(mh, __ffma2_rn(D, iw, ml));
            const float2 v = __ffma2_rn(make_float2(-m.x, -m.y), m, __ffma2_rn(Q, iw, e0));
            // one MUFU per value where the window is not near-constant (eps r < 1e-3: 1 / (1 + eps r) = 1 - eps r
            // to float32); the exact rsqrt + rcp path otherwise (v <= 1e-12, rounding-negative, NaN)
            float2 ov;
            if (v.x > 1e-12f && v.y > 1e-12f) {
              float rx, ry;
              asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rx) : "f"(v.x));
              asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(ry) : "f"(v.y));
              const float2 r = make_float2(rx, ry);
              ov = __fmul2_rn(__fmul2_rn(xc, r), __ffma2_rn(r, big_neg(neps), one));
            } else {
              const float vx = v.x < 1e-30f ? 1e-30f : v.x, vy = v.y < 1e-30f ? 1e-30f : v.y;
              float rx, ry, cx, cy;
              asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rx) : "f"(vx));
              asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(ry) : "f"(vy));
              const float2 r = make_float2(rx, ry);
              const float2 d = __ffma2_rn(r, neps, one);
              asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(cx) : "f"(d.x));
              asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(cy) : "f"(d.y));
              ov = __fmul2_rn(__fmul2_rn(xc, r), make_float2(cx, cy));
            }
            *reinterpret_cast<float2 *>(dst) = ov;
            D = __fadd2_rn(D, big_sub(xi, xo));
            Q = __fadd2_rn(Q, __ffma2_rn(xi, xi, big_neg(__fmul2_rn(xo, xo))));
            dst += ld; pc += cols; pin += cols; pout += cols;
          }
        }
      }
    }
  }
}

static size_t cm_big_smem_bytes(int rmax, int cols, int win) {
  const int rows_ext = rmax + win - 1;
  const int nblk = (rows_ext + kBigG - 1) / kBigG + 1;
  const size_t s_bytes = ((size_t)rows_ext * cols * 4 + 15) & ~(size_t)15;
  return s_bytes + 2 * (size_t)nblk * cols * 8;
}

// Largest tile (multiple of 30 rows, at most 960) that fits one CTA per SM; 0 = unsupported.
int cmvnw_big_rmax(int cols, int win) {
  if (cols < 4 || (cols & 3) || cols / 2 > kBigThreads || win < kBigG) return 0;
  int best = 0;
  for (int r = kBigG; r <= 960; r += kBigG)
    if (cm_big_smem_bytes(r, cols, win) <= 226 * 1024) best = r;
  return best >= 4 * kBigG ? best : 0;
}

int launch_cmvnw_big(spb_context *ctx, const float *in, float *scratch_ms, float *out,
                     const int32_t *nframes, const int64_t *frame_offsets, const int4 *tiles,
                     int64_t num_tiles, int rmax, int cols, int ld, int win, int varnorm, int first_stage) {
  if (num_tiles <= 0) return SPB_OK;
  int *counters = (int *)scratch(ctx, 7, 256);
  if (!counters) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemsetAsync(counters, 0, 8, ctx->stream));
  CmBigParams p;
  p.nframes = nframes;
  p.frame_offsets = frame_offsets;
  p.tiles = tiles;
  p.num_tiles = (int)num_tiles;
  p.cols = cols;
  p.ld = ld;
  p.win = win;
  p.rmax = rmax;
  p.eps = 9.313225746154785e-10f;  // 2^-30 (processing.py:292)
  const size_t smem = cm_big_smem_bytes(rmax, cols, win);
  for (int stage = first_stage; stage <= (varnorm ? 2 : 1); ++stage) {
    auto kern = (cols == 40) ? (stage == 1 ? cmvnw_big_kernel<1, 40> : cmvnw_big_kernel<2, 40>)
                             : (stage == 1 ? cmvnw_big_kernel<1, 0> : cmvnw_big_kernel<2, 0>);
    int per_sm = 0;
    int rc = configure_kernel(ctx, (const void *)kern, kBigThreads, smem, &per_sm);
    if (rc) return rc;
    if (per_sm < 1) return fail(ctx, SPB_ERR_UNSUPPORTED, "cmvnw: tile does not fit in shared memory");
    long long grid = std::min<long long>(ctx->sm_count, num_tiles);
    p.in = (stage == 1) ? in : scratch_ms;
    p.out = (stage == 1) ? (varnorm ? scratch_ms : out) : out;
    p.counter = counters + (stage - 1);
    const size_t tb = timing_begin(ctx);
    kern<<<(unsigned)grid, kBigThreads, smem, ctx->stream>>>(p);
    timing_end(ctx, stage, tb);
    SPB_LAUNCH_CHECK(ctx, "cmvnw_big_kernel");
  }
  return SPB_OK;
}

// ------------------------------------------------------------------------------------
// cmvnw stage 1 on float64 input (the public processing.cmvnw call hands float64): window means in
// float64 from the float64 values, x - mean in float64, ONE rounding to float32 on the store --
// exactly processing.py:302-307 (mean_subtracted is a float32 array).  Narrowing the input first
// would quantise a large common offset (1e4 +- 1e-2) before it is subtracted.  Stage 2 then runs on
// the float32 rows like the reference.  One thread per (column, chunk of 64 rows): the first window
// is summed directly, the rest slides; not a hot path (this call is bound by PCIe).
// ------------------------------------------------------------------------------------
constexpr int kS1Chunk = 64;
__global__ void __launch_bounds__(128)
cmvnw_stage1_f64_kernel(const double *__restrict__ in, long long rows, int cols, int win, float *__restrict__ out) {
  const int c = blockIdx.y * 32 + threadIdx.x;
  const long long r0 = ((long long)blockIdx.x * 4 + threadIdx.y) * kS1Chunk;
  if (c >= cols || r0 >= rows) return;
  const int F = (int)rows;
  const int h = (win - 1) >> 1;
  const double inv = 1.0 / (double)win;
  double w = 0.0;
  for (int t = -h; t <= h; ++t) w += in[(long long)reflect_row(r0 + t, F) * cols + c];
  const long long r1 = r0 + kS1Chunk < rows ? r0 + kS1Chunk : rows;
  for (long long r = r0; r < r1; ++r) {
    out[r * cols + c] = (float)(in[r * cols + c] - w * inv);
    w += in[(long long)reflect_row(r + h + 1, F) * cols + c] - in[(long long)reflect_row(r - h, F) * cols + c];
    if (w != w || fabs(w) == INFINITY) {
      // a NaN / inf that slid in never slides out of a running sum: sum the next window directly, so that
      // exactly the rows whose window holds it are poisoned (as in the reference)
      w = 0.0;
      for (int t = -h; t <= h; ++t) w += in[(long long)reflect_row(r + 1 + t, F) * cols + c];
    }
  }
}

int launch_cmvnw_stage1_f64(spb_context *ctx, const double *in, int64_t rows, int cols, int win, float *out) {
  if (rows <= 0 || cols <= 0) return SPB_OK;
  const dim3 grid((unsigned)((rows + 4 * kS1Chunk - 1) / (4 * kS1Chunk)), (unsigned)((cols + 31) / 32));
  cmvnw_stage1_f64_kernel<<<grid, dim3(32, 4), 0, ctx->stream>>>(in, rows, cols, win, out);
  SPB_LAUNCH_CHECK(ctx, "cmvnw_stage1_f64_kernel");
  return SPB_OK;
}

// ------------------------------------------------------------------------------------
// cmvn (global), float64 throughout: processing.py:253-271.
//   out = (x - mean0) / (std0(x - mean0) + 2^-30), np.std = sqrt(mean(|c - mean(c)|^2)).
// One CTA per 32 columns, (32 x 8) threads, four sweeps over the rows.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ double cta_colsum(double v, double (*sh)[33]) {
  // threadIdx.x = column lane, threadIdx.y = row group; returns the column total to all rows
  sh[threadIdx.y][threadIdx.x] = v;
  __syncthreads();
  double t = 0.0;
#pragma unroll
  for (int y = 0; y < 8; ++y) t += sh[y][threadIdx.x];
  __syncthreads();
  return t;
}

__global__ void __launch_bounds__(256)
cmvn_kernel(const double *__restrict__ in, double *__restrict__ out, long long rows, int cols,
            int varnorm, double eps) {
  __shared__ double sh[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  const bool ok = c < cols;
  double acc = 0.0;
  if (ok) for (long long r = threadIdx.y; r < rows; r += 8) acc += in[r * cols + c];
  const double mean = cta_colsum(acc, sh) / (double)rows;
  if (!varnorm) {
    if (ok) for (long long r = threadIdx.y; r < rows; r += 8) out[r * cols + c] = in[r * cols + c] - mean;
    return;
  }
  acc = 0.0;
  if (ok) for (long long r = threadIdx.y; r < rows; r += 8) acc += in[r * cols + c] - mean;
  const double m2 = cta_colsum(acc, sh) / (double)rows;
  acc = 0.0;
  if (ok) for (long long r = threadIdx.y; r < rows; r += 8) {
    const double d = (in[r * cols + c] - mean) - m2;
    acc = fma(d, d, acc);
  }
  const double sd = sqrt(cta_colsum(acc, sh) / (double)rows);
  if (ok) for (long long r = threadIdx.y; r < rows; r += 8)
    out[r * cols + c] = (in[r * cols + c] - mean) / (sd + eps);
}

int launch_cmvn_f64(spb_context *ctx, const double *in, int64_t rows, int cols, int varnorm,
                    double *out) {
  if (rows <= 0 || cols <= 0) return SPB_OK;
  dim3 block(32, 8);
  dim3 grid((cols + 31) / 32);
  cmvn_kernel<<<grid, block, 0, ctx->stream>>>(in, out, rows, cols, varnorm, 9.313225746154785e-10);
  SPB_LAUNCH_CHECK(ctx, "cmvn_kernel");
  return SPB_OK;
}

}  // namespace spb
