// K3s: cmvnw in ONE pass (window 301, variance normalisation, 40 columns: the shape of the batch
// path, processing.py:301-322).
//
// The two-pass kernels (cmvn_kernels.cu) read the lmfe rows, write the mean-subtracted rows, read
// those back with a +-150 halo and write the result: 4 x the feature array through HBM plus the
// halo re-reads.  Here one CTA STREAMS an utterance (or a segment of a long one) top to bottom and
// keeps everything it needs in two rings in shared memory:
//
//   X ring (660 rows)  lmfe rows, written by TMA bulk copies one super-step ahead
//   Y ring (525 rows)  mean-subtracted rows (float32, like the reference's intermediate array)
//
// Per super-step of R = 180 rows: (1) the X rows that arrived (and the symmetric-pad rows at the
// utterance ends, copied in shared memory) get float64 sums of 15-row blocks; (2) stage 1 for 180
// rows: window sum of the first row of each 15-row block from 20 block sums + 1 row, then a
// float32 delta slide anchored to it -> Y ring; (3) pad rows / block sums of Y; (4) stage 2 for the
// 180 rows whose +-150 window of Y is complete -> global memory.  Every lmfe row is read from HBM
// once, every output row written once, nothing else touches global memory.
//
// The arithmetic is the one of cmvnw_big_kernel (float64 block sums, float32 deltas over <= 14 rows,
// mean split in two floats, population variance in float32 like np.std of a float32 array) --
// only the anchors sit every 15 rows instead of every 30.
//
// Streams: {utterance, first output row a, end row b}.  An utterance is one stream; utterances
// longer than a few thousand rows are cut into segments, each of which re-derives its 150 + 150
// rows of context from the neighbouring rows.  Utterances shorter than 150 rows (periodic
// reflection) stay with the two-pass kernels.
#include <algorithm>

#include "ptx_helpers.cuh"
#include "spb_internal.cuh"

namespace spb {

constexpr int kStG = 15;                       // rows per block sum
constexpr int kStR = 180;                      // rows per super-step (>= H, multiple of G)
constexpr int kStThreads = 288;                // 8 worker warps + 1 warp whose lane 0 drives TMA and the stream queue
constexpr int kStProducer = 256;
constexpr int kStCols = 40;
constexpr int kStPairs = kStCols / 2;          // a thread owns a column pair ...
constexpr int kStBT = 256 / kStPairs;          // ... of every 12th block

template <int H>
struct StGeom {
  static constexpr int WIN = 2 * H + 1;
  static constexpr int NB = 2 * H / kStG;      // whole blocks per window (+ 1 row)
  static constexpr int CX = 2 * kStR + 2 * H;  // X ring rows
  static constexpr int CY = (kStR + WIN + 2 * kStG + kStG - 1) / kStG * kStG;
  static constexpr int NBX = CX / kStG, NBY = CY / kStG;
  static constexpr size_t smem = (size_t)(CX + CY) * kStCols * 4 + (size_t)(NBX + 2 * NBY) * kStCols * 8;
  static_assert(H % kStG == 0 && kStR % kStG == 0 && kStR >= H, "stream geometry");
  static_assert(smem <= 232448 - 256, "stream kernel: shared memory budget");
};

struct CmStreamParams {
  const float *in;               // [rows][40], addressed by global row
  float *out;
  const int32_t *nframes;
  const int64_t *frame_offsets;
  const int4 *streams;           // {utt, a, b, 0}
  int num_streams;
  int *counter;
  float eps;
};

// everything a stream's schedule depends on, derived from its descriptor
struct StStream {
  int valid;
  int a, T, F;
  long long base;    // first global row of the utterance
  int ylo;           // first real stage-1 row
  int MI;            // real stage-1 rows
  int XE, xr0, xr1;  // X-ext rows; real ones are [xr0, xr1)
  int yr0, yr1, YE;  // Y-ext rows; real ones are [yr0, yr1)
  int NS;            // super-steps
};

template <int H>
__device__ __forceinline__ StStream st_derive(const int4 d, const long long base) {
  StStream s;
  s.valid = d.x >= 0;
  s.a = d.y;
  s.T = d.z - d.y;
  s.F = d.w;
  s.base = base;
  s.ylo = max(0, d.y - H);
  const int yhi = min(d.w, d.z + H);
  s.MI = yhi - s.ylo;
  s.XE = s.MI + 2 * H;
  s.xr0 = max(0, H - s.ylo);
  s.xr1 = min(s.XE, d.w - s.ylo + H);
  s.yr0 = s.ylo - (d.y - H);
  s.yr1 = s.yr0 + s.MI;
  s.YE = s.T + 2 * H;
  s.NS = (s.YE - s.yr0 + kStR - 1) / kStR;
  return s;
}

// X-ext rows complete after the loads of super-step s
template <int H>
__device__ __forceinline__ int st_xfront(const StStream &st, const int s) {
  return s < 0 ? 0 : min((s + 1) * kStR + 2 * H, st.XE);
}
// real X-ext rows super-step s loads: [lo, hi)
template <int H>
__device__ __forceinline__ void st_load_range(const StStream &st, const int s, int &lo, int &hi) {
  lo = max(st.xr0, st_xfront<H>(st, s - 1));
  hi = min(st.xr1, st_xfront<H>(st, s));
}

template <int H>
__device__ __forceinline__ void st_issue_load(const CmStreamParams &p, const StStream &st, const int s, float *X,
                                              uint64_t *bar) {
  using Gm = StGeom<H>;
  int lo, hi;
  st_load_range<H>(st, s, lo, hi);
  if (hi <= lo) return;
  // earlier generic-proxy reads of these ring rows are ordered before the async-proxy writes
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  mbar_expect_tx(bar, (uint32_t)(hi - lo) * (kStCols * 4));
  const float *src = p.in + (st.base + (long long)(lo + st.ylo - H)) * kStCols;
  int x = lo;
  while (x < hi) {
    const int r = x % Gm::CX;
    const int n = min(min(hi - x, Gm::CX - r), 384);  // <= 61 440 B per copy, split at the ring's end
    bulk_load(X + (size_t)r * kStCols, src + (size_t)(x - lo) * kStCols, (uint32_t)n * (kStCols * 4), bar);
    x += n;
  }
}

// ---- packed fp32 helpers (Blackwell FADD2 / FMUL2 / FFMA2: a column pair per instruction)
__device__ __forceinline__ float2 st_neg(const float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 st_sub(const float2 a, const float2 b) { return __fadd2_rn(a, st_neg(b)); }

// float64 sums of one G-row block of a ring (rows contiguous: blocks never straddle the ring's end).
// All loads are issued before the first conversion, three accumulators per sum: with two warps per
// scheduler the kernel lives on instruction-level parallelism.
template <bool SQ>
__device__ __forceinline__ void st_block_sums(const float *col, double2 &s1, double2 &s2) {
  float2 v[kStG];
#pragma unroll
  for (int r = 0; r < kStG; ++r) v[r] = *reinterpret_cast<const float2 *>(col + r * kStCols);
  double ax[3] = {0.0, 0.0, 0.0}, ay[3] = {0.0, 0.0, 0.0}, qx[3] = {0.0, 0.0, 0.0}, qy[3] = {0.0, 0.0, 0.0};
#pragma unroll
  for (int r = 0; r < kStG; ++r) {
    const double dx = (double)v[r].x, dy = (double)v[r].y;
    ax[r % 3] += dx;
    ay[r % 3] += dy;
    if (SQ) {
      qx[r % 3] = fma(dx, dx, qx[r % 3]);
      qy[r % 3] = fma(dy, dy, qy[r % 3]);
    }
  }
  s1 = make_double2((ax[0] + ax[1]) + ax[2], (ay[0] + ay[1]) + ay[2]);
  s2 = make_double2((qx[0] + qx[1]) + qx[2], (qy[0] + qy[1]) + qy[2]);
}

// sum of NB consecutive block sums of a ring of NR blocks starting at block k0 (loads first, four chains)
template <int NB, int NR>
__device__ __forceinline__ double2 st_window_blocks(const double *bs, const int k0) {
  double2 t[NB];
  int k = k0;
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    t[j] = *reinterpret_cast<const double2 *>(bs + (size_t)k * kStCols);
    k = (k + 1 == NR) ? 0 : k + 1;
  }
  double ax[4] = {0.0, 0.0, 0.0, 0.0}, ay[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    ax[j & 3] += t[j].x;
    ay[j & 3] += t[j].y;
  }
  return make_double2((ax[0] + ax[1]) + (ax[2] + ax[3]), (ay[0] + ay[1]) + (ay[2] + ay[3]));
}

// mean of a window sum, split in two floats: x - mean = (x - mh) - (ml + D / win) carries no
// float64 arithmetic in the slide (processing.py:304-307 subtracts in float64)
__device__ __forceinline__ void st_split_mean(const double wx, const double wy, const double inv_win, float2 &mh,
                                              float2 &ml) {
  const double m0x = wx * inv_win, m0y = wy * inv_win;
  mh = make_float2((float)m0x, (float)m0y);
  ml = make_float2((float)(m0x - (double)mh.x), (float)(m0y - (double)mh.y));
}

// stage 2 of one row: ms / (std + eps) (processing.py:318-322), population variance in float32.
// Branch-free (a branch per row keeps the compiler from overlapping the rows of a run):
// x / (s + eps) = x r / (1 + eps r) with r = 1 / s = rsqrt(v).  v is clamped from below with a
// select -- v <= 0 happens for constant windows (rounding) and must give x / eps; fmaxf would swallow
// a NaN variance, which has to poison the row like the reference.
__device__ __forceinline__ float2 st_normalise(const float2 xc, const float2 D, const float2 Q, const float2 mh,
                                               const float2 ml, const float2 e0, const float2 inv_w, const float eps) {
  const float2 m = __fadd2_rn(mh, __ffma2_rn(D, inv_w, ml));
  const float2 v = __ffma2_rn(st_neg(m), m, __ffma2_rn(Q, inv_w, e0));
  const float vx = v.x < 1e-30f ? 1e-30f : v.x, vy = v.y < 1e-30f ? 1e-30f : v.y;
  float rx, ry, cx, cy;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rx) : "f"(vx));
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(ry) : "f"(vy));
  const float2 r = make_float2(rx, ry);
  const float2 d = __ffma2_rn(r, make_float2(eps, eps), make_float2(1.f, 1.f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(cx) : "f"(d.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(cy) : "f"(d.y));
  return __fmul2_rn(__fmul2_rn(xc, r), make_float2(cx, cy));
}

template <int H>
__global__ void __launch_bounds__(kStThreads, 1) cmvnw_stream_kernel(const CmStreamParams p) {
  using Gm = StGeom<H>;
  constexpr int G = kStG, R = kStR, WIN = Gm::WIN, NB = Gm::NB, CX = Gm::CX, CY = Gm::CY;
  constexpr int NBX = Gm::NBX, NBY = Gm::NBY, C = kStCols;
  extern __shared__ __align__(16) unsigned char st_smem[];
  float *X = reinterpret_cast<float *>(st_smem);
  float *Y = X + (size_t)CX * C;
  double *bsX = reinterpret_cast<double *>(Y + (size_t)CY * C);
  double *bsY1 = bsX + (size_t)NBX * C;
  double *bsY2 = bsY1 + (size_t)NBY * C;
  __shared__ int4 s_desc[2];
  __shared__ long long s_base[2];
  __shared__ __align__(8) uint64_t s_bar;

  const int tid = threadIdx.x;
  const int g = tid % kStPairs, bi = tid / kStPairs;
  const bool worker = bi < kStBT;
  const int c0 = 2 * g;
  const float2 inv_w = make_float2(1.0f / (float)WIN, 1.0f / (float)WIN);
  const double inv_win = 1.0 / (double)WIN;

  auto fetch_desc = [&](int slot) {
    const int t = atomicAdd(p.counter, 1);
    int4 d = make_int4(-1, 0, 0, 0);
    long long b = 0;
    if (t < p.num_streams) {
      const int4 sd = p.streams[t];
      d = make_int4(sd.x, sd.y, sd.z, p.nframes[sd.x]);
      b = p.frame_offsets[sd.x];
    }
    s_desc[slot] = d;
    s_base[slot] = b;
  };
  if (tid == kStProducer) {
    mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fetch_desc(0);
    const StStream st0 = st_derive<H>(s_desc[0], s_base[0]);
    if (st0.valid) st_issue_load<H>(p, st0, 0, X, &s_bar);
  }
  uint32_t xphase = 0;

  for (int slot = 0;; slot ^= 1) {
    __syncthreads();  // descriptor `slot` published; the previous stream is finished
    const StStream st = st_derive<H>(s_desc[slot], s_base[slot]);
    if (!st.valid) break;
    if (tid == kStProducer) fetch_desc(slot ^ 1);  // read again by this thread only until the next stream starts
    int yb_done = 0, ob_done = 0;
    for (int s = 0; s < st.NS; ++s) {
      // ---- P0: the X rows of this super-step have landed; ask for the next ones
      int lo, hi;
      st_load_range<H>(st, s, lo, hi);
      if (hi > lo) {
        mbar_wait(&s_bar, xphase);
        xphase ^= 1;
      }
      if (tid == kStProducer && s + 1 < st.NS) st_issue_load<H>(p, st, s + 1, X, &s_bar);
      const int xfp = st_xfront<H>(st, s - 1), xf = st_xfront<H>(st, s);
      // ---- P1: symmetric-pad rows of X in [xfp, xf): copies of rows that are resident
      {
        const int h_hi = min(st.xr0, xf);               // head pads [xfp, h_hi)
        const int t_lo = max(st.xr1, xfp);              // tail pads [t_lo, xf)
        const int nh = max(h_hi - xfp, 0), nt = max(xf - t_lo, 0);
        if (nh + nt > 0) {
          for (int i = tid; i < (nh + nt) * (C / 4); i += kStThreads) {
            const int k = i / (C / 4), c4 = i - k * (C / 4);
            const int x = (k < nh) ? xfp + k : t_lo + (k - nh);
            const int j = x + st.ylo - H;                // utterance row (outside [0, F))
            const int js = (j < 0) ? -1 - j : 2 * st.F - 1 - j;
            const int xs = js - st.ylo + H;
            reinterpret_cast<float4 *>(X + (size_t)(x % CX) * C)[c4] =
                reinterpret_cast<const float4 *>(X + (size_t)(xs % CX) * C)[c4];
          }
          __syncthreads();
        }
      }
      // ---- P2: float64 sums of the X blocks that became complete
      if (worker) {
        for (int blk = xfp / G + bi; blk < xf / G; blk += kStBT) {
          double2 s1, s2;
          st_block_sums<false>(X + (size_t)((blk * G) % CX) * C + c0, s1, s2);
          *reinterpret_cast<double2 *>(bsX + (size_t)(blk % NBX) * C + c0) = s1;
        }
      }
      __syncthreads();
      // ---- P3: stage 1 (processing.py:301-307) for stage-1 rows [sR, min((s+1)R, MI)) -> Y ring.
      // Every run of G rows starts at a multiple of G in its ring (H, R, yr0 are multiples of G), so no
      // pointer wraps inside a run; the row entering the window sits one row further (WIN = 2H + 1),
      // but the update after a block's last row is never used.
      const int m_hi = min((s + 1) * R, st.MI);
      if (worker) {
        for (int mb = s * (R / G) + bi; mb * G < m_hi; mb += kStBT) {
          const int mi0 = mb * G;
          const int nrow = min(G, st.MI - mi0);
          double2 w = st_window_blocks<NB, NBX>(bsX + c0, mb % NBX);
          const float *po = X + (size_t)(mi0 % CX) * C + c0;            // rows leaving the window
          const float *pc = X + (size_t)((mi0 + H) % CX) * C + c0;      // centre rows
          const float *pi = X + (size_t)((mi0 + 2 * H) % CX) * C + c0;  // entering rows, minus one (ring-safe, see above)
          float *py = Y + (size_t)((mi0 + st.yr0) % CY) * C + c0;
          if (nrow == G) {
            // every load of the run first (nothing below may be hoisted over the stores to Y by the compiler)
            float2 xc[G], xi[G], xo[G];
#pragma unroll
            for (int q = 0; q < G; ++q) {
              xc[q] = *reinterpret_cast<const float2 *>(pc + q * C);
              xi[q] = *reinterpret_cast<const float2 *>(pi + q * C);
              if (q + 1 < G) xo[q] = *reinterpret_cast<const float2 *>(po + q * C);
            }
            w.x += (double)xi[0].x;
            w.y += (double)xi[0].y;
            float2 mh, ml;
            st_split_mean(w.x, w.y, inv_win, mh, ml);
            float2 D = make_float2(0.f, 0.f);
            float2 o[G];
#pragma unroll
            for (int q = 0; q < G; ++q) {
              o[q] = st_sub(st_sub(xc[q], mh), __ffma2_rn(D, inv_w, ml));
              *reinterpret_cast<float2 *>(py + q * C) = o[q];
              if (q + 1 < G) D = __fadd2_rn(D, st_sub(xi[q + 1], xo[q]));
            }
            // the block of Y this run filled is complete: its sums are known here
            double ax[3] = {0.0, 0.0, 0.0}, ay[3] = {0.0, 0.0, 0.0}, qx[3] = {0.0, 0.0, 0.0}, qy[3] = {0.0, 0.0, 0.0};
#pragma unroll
            for (int q = 0; q < G; ++q) {
              const double dx = (double)o[q].x, dy = (double)o[q].y;
              ax[q % 3] += dx;
              ay[q % 3] += dy;
              qx[q % 3] = fma(dx, dx, qx[q % 3]);
              qy[q % 3] = fma(dy, dy, qy[q % 3]);
            }
            const int yb = ((mi0 + st.yr0) / G) % NBY;
            *reinterpret_cast<double2 *>(bsY1 + (size_t)yb * C + c0) = make_double2((ax[0] + ax[1]) + ax[2], (ay[0] + ay[1]) + ay[2]);
            *reinterpret_cast<double2 *>(bsY2 + (size_t)yb * C + c0) = make_double2((qx[0] + qx[1]) + qx[2], (qy[0] + qy[1]) + qy[2]);
          } else {
            const float2 v = *reinterpret_cast<const float2 *>(pi);
            w.x += (double)v.x;
            w.y += (double)v.y;
            float2 mh, ml;
            st_split_mean(w.x, w.y, inv_win, mh, ml);
            float2 D = make_float2(0.f, 0.f);
            for (int q = 0; q < nrow; ++q) {
              const float2 xc = *reinterpret_cast<const float2 *>(pc + q * C);
              *reinterpret_cast<float2 *>(py + q * C) = st_sub(st_sub(xc, mh), __ffma2_rn(D, inv_w, ml));
              if (q + 1 < nrow) {
                const float2 xi = *reinterpret_cast<const float2 *>(pi + (q + 1) * C);
                const float2 xo = *reinterpret_cast<const float2 *>(po + q * C);
                D = __fadd2_rn(D, st_sub(xi, xo));
              }
            }
          }
        }
      }
      __syncthreads();
      // stage 1 of this stream is complete: the X ring is free for the first rows of the next stream
      if (tid == kStProducer && m_hi == st.MI && s * R < st.MI) {
        const StStream nx = st_derive<H>(s_desc[slot ^ 1], s_base[slot ^ 1]);
        if (nx.valid) st_issue_load<H>(p, nx, 0, X, &s_bar);
      }
      // ---- P4: symmetric-pad rows of Y (processing.py:314-315 pads the float32 array itself)
      const int cf = min(st.yr0 + (s + 1) * R, st.YE);  // Y-ext rows complete after this super-step
      {
        const int nh = (s == 0) ? st.yr0 : 0;               // head pads [0, yr0), sources [yr0, 2 yr0)
        const int t_lo = max(st.yr1, st.yr0 + s * R);       // tail pads [t_lo, cf)
        const int nt = max(cf - t_lo, 0);
        if (nh + nt > 0) {
          for (int i = tid; i < (nh + nt) * (C / 4); i += kStThreads) {
            const int k = i / (C / 4), c4 = i - k * (C / 4);
            const int y = (k < nh) ? k : t_lo + (k - nh);
            const int ys = (k < nh) ? 2 * st.yr0 - 1 - y : 2 * st.yr1 - 1 - y;
            reinterpret_cast<float4 *>(Y + (size_t)(y % CY) * C)[c4] =
                reinterpret_cast<const float4 *>(Y + (size_t)(ys % CY) * C)[c4];
          }
          __syncthreads();
        }
      }
      // ---- P5: float64 sums (x, x^2) of the complete Y blocks stage 1 did not produce itself: the
      // pad blocks at the two ends and the block that holds the last real rows
      const int yb_hi = cf / G;
      {
        const int full_lo = st.yr0 / G, full_hi = st.yr1 / G;  // blocks [full_lo, full_hi) came out of P3
        const int head_n = max(min(yb_hi, full_lo) - yb_done, 0);
        const int tail_lo = max(yb_done, full_hi);
        const int tail_n = max(yb_hi - tail_lo, 0);
        if (head_n + tail_n > 0) {
          if (worker) {
            for (int k = bi; k < head_n + tail_n; k += kStBT) {
              const int blk = (k < head_n) ? yb_done + k : tail_lo + (k - head_n);
              double2 s1, s2;
              st_block_sums<true>(Y + (size_t)((blk * G) % CY) * C + c0, s1, s2);
              *reinterpret_cast<double2 *>(bsY1 + (size_t)(blk % NBY) * C + c0) = s1;
              *reinterpret_cast<double2 *>(bsY2 + (size_t)(blk % NBY) * C + c0) = s2;
            }
          }
          __syncthreads();
        }
      }
      yb_done = yb_hi;
      // ---- P6: stage 2 (processing.py:318-322) for the output blocks whose windows are complete
      const int ob_hi = (cf == st.YE) ? (st.T + G - 1) / G : max((cf - WIN + 1) / G, 0);
      if (worker) {
        for (int tb = ob_done + bi; tb < ob_hi; tb += kStBT) {
          const int t0 = tb * G;
          const int nrow = min(G, st.T - t0);
          double2 w1 = st_window_blocks<NB, NBY>(bsY1 + c0, tb % NBY);
          double2 w2 = st_window_blocks<NB, NBY>(bsY2 + c0, tb % NBY);
          const float *po = Y + (size_t)(t0 % CY) * C + c0;
          const float *pc = Y + (size_t)((t0 + H) % CY) * C + c0;
          const float *pi = Y + (size_t)((t0 + 2 * H) % CY) * C + c0;
          float *dst = p.out + (st.base + st.a + t0) * (long long)C + c0;
          if (nrow == G) {
            float2 xc[G], xi[G], xo[G];
#pragma unroll
            for (int q = 0; q < G; ++q) {
              xc[q] = *reinterpret_cast<const float2 *>(pc + q * C);
              xi[q] = *reinterpret_cast<const float2 *>(pi + q * C);
              if (q + 1 < G) xo[q] = *reinterpret_cast<const float2 *>(po + q * C);
            }
            {
              const double dx = (double)xi[0].x, dy = (double)xi[0].y;
              w1.x += dx; w1.y += dy;
              w2.x = fma(dx, dx, w2.x); w2.y = fma(dy, dy, w2.y);
            }
            float2 mh, ml;
            st_split_mean(w1.x, w1.y, inv_win, mh, ml);
            const float2 e0 = make_float2((float)(w2.x * inv_win), (float)(w2.y * inv_win));
            float2 D = make_float2(0.f, 0.f), Q = make_float2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < G; ++q) {
              *reinterpret_cast<float2 *>(dst + q * C) = st_normalise(xc[q], D, Q, mh, ml, e0, inv_w, p.eps);
              if (q + 1 < G) {
                D = __fadd2_rn(D, st_sub(xi[q + 1], xo[q]));
                Q = __fadd2_rn(Q, __ffma2_rn(xi[q + 1], xi[q + 1], st_neg(__fmul2_rn(xo[q], xo[q]))));
              }
            }
          } else {
            {
              const float2 v = *reinterpret_cast<const float2 *>(pi);
              const double dx = (double)v.x, dy = (double)v.y;
              w1.x += dx; w1.y += dy;
              w2.x = fma(dx, dx, w2.x); w2.y = fma(dy, dy, w2.y);
            }
            float2 mh, ml;
            st_split_mean(w1.x, w1.y, inv_win, mh, ml);
            const float2 e0 = make_float2((float)(w2.x * inv_win), (float)(w2.y * inv_win));
            float2 D = make_float2(0.f, 0.f), Q = make_float2(0.f, 0.f);
            for (int q = 0; q < nrow; ++q) {
              const float2 xc = *reinterpret_cast<const float2 *>(pc + q * C);
              *reinterpret_cast<float2 *>(dst + q * C) = st_normalise(xc, D, Q, mh, ml, e0, inv_w, p.eps);
              if (q + 1 < nrow) {
                const float2 xi = *reinterpret_cast<const float2 *>(pi + (q + 1) * C);
                const float2 xo = *reinterpret_cast<const float2 *>(po + q * C);
                D = __fadd2_rn(D, st_sub(xi, xo));
                Q = __fadd2_rn(Q, __ffma2_rn(xi, xi, st_neg(__fmul2_rn(xo, xo))));
              }
            }
          }
        }
      }
      ob_done = ob_hi;
      __syncthreads();
    }
  }
}

// ------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------
bool cmvnw_stream_supported(int cols, int win, int varnorm) {
  return cols == kStCols && win == StGeom<150>::WIN && varnorm != 0;
}
int cmvnw_stream_min_rows() { return 150; }

// streams of one utterance (F >= cmvnw_stream_min_rows()): whole, or balanced segments of
// <= ~2 400 rows (multiples of the block) when it is long
void append_streams(std::vector<int4> &v, int utt, int64_t F) {
  if (F < cmvnw_stream_min_rows()) return;
  const int64_t seg_max = 2400;
  if (F <= 3000) {
    v.push_back(make_int4(utt, 0, (int)F, 0));
    return;
  }
  const int64_t n = (F + seg_max - 1) / seg_max;
  int64_t r = (F + n - 1) / n;
  r = (r + kStG - 1) / kStG * kStG;
  for (int64_t a = 0; a < F; a += r) v.push_back(make_int4(utt, (int)a, (int)std::min<int64_t>(a + r, F), 0));
}

int launch_cmvnw_stream(spb_context *ctx, const float *in, float *out, const int32_t *nframes,
                        const int64_t *frame_offsets, const int4 *streams, int64_t num_streams) {
  if (num_streams <= 0) return SPB_OK;
  int *counters = (int *)scratch(ctx, 7, 256);
  if (!counters) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemsetAsync(counters + 2, 0, 4, ctx->stream));
  CmStreamParams p;
  p.in = in;
  p.out = out;
  p.nframes = nframes;
  p.frame_offsets = frame_offsets;
  p.streams = streams;
  p.num_streams = (int)num_streams;
  p.counter = counters + 2;
  p.eps = 9.313225746154785e-10f;  // 2^-30 (processing.py:292)
  auto kern = cmvnw_stream_kernel<150>;
  const size_t smem = StGeom<150>::smem;
  int per_sm = 0;
  int rc = configure_kernel(ctx, (const void *)kern, kStThreads, smem, &per_sm);
  if (rc) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_UNSUPPORTED, "cmvnw stream kernel does not fit in shared memory");
  const long long grid = std::min<long long>(ctx->sm_count, num_streams);
  const size_t tb = timing_begin(ctx);
  kern<<<(unsigned)grid, kStThreads, smem, ctx->stream>>>(p);
  timing_end(ctx, 4, tb);
  SPB_LAUNCH_CHECK(ctx, "cmvnw_stream_kernel");
  return SPB_OK;
}

}  // namespace spb
