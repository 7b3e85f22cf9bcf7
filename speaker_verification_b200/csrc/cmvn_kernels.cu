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
#include "spb_internal.cuh"

namespace spb {

constexpr int kCmG = 15;          // rows per block sum
constexpr int kCmThreads = 320;

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
              var = var > 0.0 ? var : 0.0;
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
                     int64_t num_tiles, int tile_rows, int cols, int ld, int win, int varnorm) {
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
  for (int stage = 1; stage <= (varnorm ? 2 : 1); ++stage) {
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
