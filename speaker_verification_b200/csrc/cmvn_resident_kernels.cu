// K3r: cmvnw with the whole utterance resident in shared memory -- both stages in ONE pass, IN PLACE
// (window 301, variance normalisation, 40 columns; utterances of <= kResMaxRows rows; longer ones stay
// with the two-pass kernels of cmvn_kernels.cu).  processing.py:301-322.
//
// The two-pass kernels spend 3/4 of their time in the per-tile skeleton (tile load, halo, barriers) and
// pay it twice, with the float32 mean-subtracted rows written to and read back from global memory in
// between.  Here a CTA loads an utterance once (TMA bulk copy), fills the +-150 symmetric-pad rows from
// shared memory, and runs
//
//   block sums (float64, 30-row blocks)  ->  stage 1 IN PLACE  ->  pad rows + block sums of the result
//   ->  stage 2  ->  global memory
//
// Stage 1 in place needs care: the thread that slides over output block b reads the centre rows of ext
// block b + 5, the rows leaving the window in ext block b and the rows entering it in ext block b + 10
// (+1), and writes its result over the centre rows -- other threads' inputs.  The 30 rows of a block are
// therefore done in two halves of 15: every thread computes its 15 results into registers, a barrier,
// every thread stores them, a barrier.  That is hazard-free: while the first halves (rows 0..14 of every
// ext block) are being computed nothing has been written; the second halves read, of other blocks, only
// rows 15..29 (leaving rows) and 16..29 (entering rows; the update after a block's last row is never
// used), which the first halves' stores did not touch.  The delta accumulator carries over in a register.
//
// What the phases cost was measured with clock64 (798-row utterances, cycles per utterance): load 2.5 k,
// pad rows 1.3 k, block sums 4.7 k, stage 1 as a 100-thread chain walk 11 k (first version), pad rows 1.2 k,
// block sums 4.7 k, stage 2 11 k -- the XU pipe (16 lanes per SM: F2F.F64 conversions and MUFU) is what
// the arithmetic phases queue on, hence: every second conversion on the integer pipe, ONE MUFU per value
// in stage 2 (rsqrt and a first-order correction; the exact path only where a window is near-constant).
//
// Arithmetic: exactly cmvnw_big_kernel's (float64 sums of 30-row blocks anchor the first row of a block,
// float32 deltas inside it, mean split in two floats, population variance in float32).
#include <algorithm>

#include "ptx_helpers.cuh"
#include "spb_internal.cuh"

namespace spb {

constexpr int kResG = 30;                 // rows per block sum (h = 150 = 5 blocks)
constexpr int kResH = 150, kResWin = 301;
constexpr int kResNB = 10;                // whole blocks per window (+ 1 row)
constexpr int kResCols = 40, kResPairs = 20;
constexpr int kResThreads = 640;
constexpr int kResBT = kResThreads / kResPairs;    // 32 block-threads
constexpr int kResMaxRows = 930;                   // longest resident utterance (9.6 s)
constexpr int kResExtMax = kResMaxRows + 2 * kResH;                 // 1 230 ext rows
constexpr int kResNBlk = (kResExtMax + kResG - 1) / kResG + 1;      // 42 block sums
static_assert((kResMaxRows + kResG - 1) / kResG <= kResBT, "resident cmvnw: one output block per thread");
constexpr size_t kResSmem = (size_t)kResExtMax * kResCols * 4 + 2 * (size_t)kResNBlk * kResCols * 8;
static_assert(kResSmem <= 232448 - 256, "resident cmvnw: shared memory budget");

struct CmResParams {
  const float *in;     // [rows][40], addressed by global row
  float *out;
  const int4 *tiles;   // two per utterance: {utt, F, 0, 0}, {first global row lo, hi, 0, 0}
  int num_tiles;
  int *counter;
  float eps;
};

__device__ __forceinline__ int res_reflect(int t, const int F) {
  // np.pad(..., 'symmetric'): ... x1 x0 | x0 x1 ... x(F-1) | x(F-1) x(F-2) ...
  if (t >= 0 && t < F) return t;
  if (t < 0 && t >= -F) return -1 - t;
  if (t >= F && t < 2 * F) return 2 * F - 1 - t;
  const int period = 2 * F;
  int m = t % period;
  if (m < 0) m += period;
  return (m < F) ? m : (period - 1 - m);
}

// float -> double with integer instructions (exact for normal numbers, inf / nan keep their class; zero and
// denormals become ~2^-127-sized values, far below what these sums resolve): F2F.F64.F32 runs on the XU pipe
__device__ __forceinline__ double res_f2d_alu(const float f) {
  const unsigned u = __float_as_uint(f);
  const unsigned a = u & 0x7fffffffu;
  unsigned hi = ((a >> 3) + 0x38000000u) | (u & 0x80000000u);
  if (a >= 0x7f800000u) hi |= 0x7ff00000u;
  return __hiloint2double((int)hi, (int)(u << 29));
}

__device__ __forceinline__ float2 res_neg(const float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 res_sub(const float2 a, const float2 b) { return __fadd2_rn(a, res_neg(b)); }

// symmetric-pad rows of the ext buffer of a part [lo, lo + N) of an utterance of F rows (ext row e = utterance
// row lo - 150 + e): the halo rows that fall outside [0, F) are copies of rows inside the part (all loads, then
// all stores); halo rows that are real rows of the utterance (the side facing the other part) are left alone
__device__ __forceinline__ void res_fill_pads(float *s, const int lo, const int N, const int F, const int tid) {
  constexpr int n4 = kResCols / 4;
  constexpr int kItems = 2 * kResH * n4;                              // 3 000 16-byte pieces
  constexpr int kPer = (kItems + kResThreads - 1) / kResThreads;      // 5 per thread
  float4 v[kPer];
  int dst[kPer];
#pragma unroll
  for (int j = 0; j < kPer; ++j) {
    const int i = tid + j * kResThreads;
    dst[j] = -1;
    if (i < kItems) {
      const int k = i / n4, c4 = i - k * n4;
      const int e = (k < kResH) ? k : N + k;  // [0, 150) and [N + 150, N + 300)
      const int u = lo - kResH + e;
      if (u < 0 || u >= F) {
        const int src = res_reflect(u, F) - (lo - kResH);
        dst[j] = e * kResCols + c4 * 4;
        v[j] = *reinterpret_cast<const float4 *>(s + (size_t)src * kResCols + c4 * 4);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kPer; ++j)
    if (dst[j] >= 0) *reinterpret_cast<float4 *>(s + dst[j]) = v[j];
}

// thread-block clusters: rank, barrier, a 16-byte load from the peer CTA's shared memory (DSMEM)
__device__ __forceinline__ uint32_t res_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void res_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float4 res_ld_peer(const float *local_ptr, const uint32_t peer_rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_ptr)), "r"(peer_rank));
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(remote) : "memory");
  return v;
}

// float64 sums (and sums of squares) of the 30-row blocks of the ext buffer; rows past `rows` count as absent
template <bool SQ>
__device__ __forceinline__ void res_block_sums(const float *s, const int rows, double *bs1, double *bs2, const int b0,
                                               const int c0) {
  const int nblk = (rows + kResG - 1) / kResG;
  for (int k = b0; k < nblk; k += kResBT) {
    const float *col = s + (size_t)k * kResG * kResCols + c0;
    const int n = min(kResG, rows - k * kResG);
    double ax[2] = {0.0, 0.0}, ay[2] = {0.0, 0.0}, qx[2] = {0.0, 0.0}, qy[2] = {0.0, 0.0};
    if (n == kResG) {
#pragma unroll 10
      for (int r = 0; r < kResG; ++r) {
        const float2 x = *reinterpret_cast<const float2 *>(col + r * kResCols);
        const double dx = (double)x.x, dy = res_f2d_alu(x.y);  // one conversion per pipe (XU / integer)
        ax[r & 1] += dx; ay[r & 1] += dy;
        if (SQ) { qx[r & 1] = fma(dx, dx, qx[r & 1]); qy[r & 1] = fma(dy, dy, qy[r & 1]); }
      }
    } else {
      for (int r = 0; r < n; ++r) {
        const float2 x = *reinterpret_cast<const float2 *>(col + r * kResCols);
        const double dx = (double)x.x, dy = (double)x.y;
        ax[0] += dx; ay[0] += dy;
        if (SQ) { qx[0] = fma(dx, dx, qx[0]); qy[0] = fma(dy, dy, qy[0]); }
      }
    }
    *reinterpret_cast<double2 *>(bs1 + k * kResCols + c0) = make_double2(ax[0] + ax[1], ay[0] + ay[1]);
    if (SQ) *reinterpret_cast<double2 *>(bs2 + k * kResCols + c0) = make_double2(qx[0] + qx[1], qy[0] + qy[1]);
  }
}

// CL = 1: one CTA per utterance (<= kResMaxRows rows), utterances pulled from an atomic queue.
// CL = 2: a CLUSTER of two CTAs per utterance (<= 2 kResMaxRows rows): each CTA keeps one half (a multiple of 30
// rows) with a 150-row halo -- the symmetric pad on its outer side, the neighbour half's rows on the inner side
// (lmfe rows loaded from global memory like its own; the neighbour's mean-subtracted rows fetched from the peer
// CTA's shared memory over DSMEM between two cluster barriers) --, utterances dealt statically to the clusters
// so that both CTAs of a cluster walk the same list.
template <int CL>
__global__ void __launch_bounds__(kResThreads, 1) cmvnw_resident_kernel(const CmResParams p) {
  extern __shared__ __align__(16) unsigned char res_smem[];
  float *s = reinterpret_cast<float *>(res_smem);                                   // [ext rows][40]
  double *bs1 = reinterpret_cast<double *>(res_smem + (size_t)kResExtMax * kResCols * 4);  // [blocks][40]
  double *bs2 = bs1 + (size_t)kResNBlk * kResCols;
  __shared__ int4 s_desc[2];
  __shared__ long long s_base[2];
  __shared__ __align__(8) uint64_t s_bar;
  const int tid = threadIdx.x;
  // (block-thread b0, column pair g) of a thread.  A row is 160 B = 20 pairs, so 32 consecutive pairs straddle two
  // rows 30 apart, and 30 x 160 B = 64 (mod 128): the second row's pairs land on the banks of the first row's last
  // four and every such 64-bit request paid a conflict wavefront (ncu r02ar: 34 % of the load wavefronts).  Instead:
  // warps 0 .. 15 take pairs 0 .. 15 of two blocks each (a half-warp = 128 contiguous bytes), warps 16 .. 19 the
  // pairs 16 .. 19 of eight blocks each.
  static_assert(kResPairs == 20 && kResThreads == 640 && kResBT == 32, "thread layout of the resident cmvnw kernel");
  const int g = tid < 512 ? (tid & 15) : 16 + ((tid - 512) & 3);
  const int b0 = tid < 512 ? (tid >> 4) : ((tid - 512) >> 2);
  const int c0 = 2 * g;
  constexpr int C = kResCols, G = kResG, H = kResH;
  const float2 iw = make_float2(1.0f / (float)kResWin, 1.0f / (float)kResWin);
  const double inv_win = 1.0 / (double)kResWin;
  if (tid == 0) {
    mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  uint32_t bar_phase = 0;
  const uint32_t crank = (CL == 2) ? res_cluster_rank() : 0u;
  auto fetch_desc = [&](int slot, int t) {
    int4 d = make_int4(-1, 0, 0, 0);
    long long b = 0;
    if (t < p.num_tiles) {
      d = __ldg(p.tiles + 2 * (long long)t);
      const int4 bb = __ldg(p.tiles + 2 * (long long)t + 1);
      b = ((long long)(unsigned)bb.x) | ((long long)bb.y << 32);
      // the next utterance's rows into L2 while this one is processed
      const char *src = reinterpret_cast<const char *>(p.in + b * (long long)C);
      const uint32_t total = (uint32_t)d.y * (uint32_t)C * 4u;
      for (uint32_t off = 0; off < total; off += 65536u) bulk_prefetch_l2(src + off, min(65536u, total - off));
    }
    s_desc[slot] = d;
    s_base[slot] = b;
  };
  // CL = 1: tickets from the atomic counter; CL = 2: tile = cluster index + i x clusters (same in both CTAs)
  int next_static = (CL == 2) ? (int)(blockIdx.x >> 1) : 0;
  const int static_step = (CL == 2) ? (int)(gridDim.x >> 1) : 0;
  auto next_ticket = [&]() {
    if (CL == 2) {
      const int t = next_static;
      next_static += static_step;
      return t;
    }
    return atomicAdd(p.counter, 1);
  };
  if (tid == kResThreads - 1) fetch_desc(0, next_ticket());
  for (int slot = 0;; slot ^= 1) {
    __syncthreads();  // previous utterance fully consumed; descriptor `slot` published
    const int4 td = s_desc[slot];
    if (td.x < 0) break;
    if (tid == kResThreads - 1) fetch_desc(slot ^ 1, next_ticket());
    const int F = td.y;
    const long long base = s_base[slot];
    // this CTA's part [lo, lo + N) of the utterance
    const int N0 = (CL == 2) ? ((F / 2 + G - 1) / G) * G : F;
    const int lo = crank ? N0 : 0;
    const int N = crank ? F - N0 : N0;
    const int ext = N + 2 * H;
    // ---- the part's rows and the real halo rows next to them: one contiguous block, TMA bulk copies
    if (tid == 0) {
      const int u_lo = max(0, lo - H), u_hi = min(F, lo + N + H);
      const char *src = reinterpret_cast<const char *>(p.in + (base + u_lo) * (long long)C);
      char *dst = reinterpret_cast<char *>(s + (size_t)(u_lo - (lo - H)) * C);
      const uint32_t total = (uint32_t)(u_hi - u_lo) * (uint32_t)C * 4u;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_expect_tx(&s_bar, total);
      for (uint32_t off = 0; off < total; off += 65536u) bulk_load(dst + off, src + off, min(65536u, total - off), &s_bar);
    }
    mbar_wait_guarded(&s_bar, bar_phase);
    bar_phase ^= 1;
    __syncthreads();
    res_fill_pads(s, lo, N, F, tid);
    __syncthreads();
    res_block_sums<false>(s, ext, bs1, bs2, b0, c0);
    __syncthreads();
    // ---- stage 1 (processing.py:301-307), in place, in two halves of 15 rows (see the file header)
    const int nbo = (N + G - 1) / G;
    {
      constexpr int HG = G / 2;
      const int b = b0;                      // one output block per thread (nbo <= 31 < 32 block-threads)
      const bool act = b < nbo;
      const int q0 = b * G;
      const float *pout = s + (size_t)q0 * C + c0;
      const float *pc = pout + (size_t)H * C;
      const float *pin = pout + (size_t)kResWin * C;
      float2 mh = make_float2(0.f, 0.f), ml = mh, D = mh;
      if (act) {
        double wx = 0.0, wy = 0.0;
#pragma unroll
        for (int i = 0; i < kResNB; ++i) {
          const double2 t = *reinterpret_cast<const double2 *>(bs1 + (b + i) * C + c0);
          wx += t.x; wy += t.y;
        }
        const float2 x = *reinterpret_cast<const float2 *>(pin - C);  // ext row q0 + 300
        wx += (double)x.x; wy += (double)x.y;
        const double m0x = wx * inv_win, m0y = wy * inv_win;
        mh = make_float2((float)m0x, (float)m0y);
        ml = make_float2((float)(m0x - (double)mh.x), (float)(m0y - (double)mh.y));
      }
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float2 o[HG];
        if (act) {
#pragma unroll
          for (int i = 0; i < HG; ++i) {
            const int t = half * HG + i;
            const float2 xc = *reinterpret_cast<const float2 *>(pc + (size_t)t * C);
            const float2 xi = *reinterpret_cast<const float2 *>(pin + (size_t)t * C);
            const float2 xo = *reinterpret_cast<const float2 *>(pout + (size_t)t * C);
            o[i] = res_sub(res_sub(xc, mh), __ffma2_rn(D, iw, ml));
            D = __fadd2_rn(D, res_sub(xi, xo));
          }
        }
        __syncthreads();  // every read of this half is done
        if (act) {
#pragma unroll
          for (int i = 0; i < HG; ++i) {
            const int q = q0 + half * HG + i;
            if (q < N) *reinterpret_cast<float2 *>(s + (size_t)(q + H) * C + c0) = o[i];
          }
        }
        __syncthreads();
      }
    }
    // ---- the float32 mean-subtracted rows are padded themselves (processing.py:314-315), then summed
    if (CL == 2) {
      // the inner halo = the neighbour half's first / last 150 mean-subtracted rows, out of the peer's shared memory
      res_cluster_sync();  // both halves' stage-1 rows are stored (release / acquire at cluster scope)
      const int e_dst = crank ? 0 : N + H;          // my inner halo (ext rows)
      const int e_src = crank ? N0 : H;             // the peer's boundary rows in ITS ext buffer
      for (int i = tid; i < H * (C / 4); i += kResThreads) {
        const int k = i / (C / 4), c4 = i - k * (C / 4);
        reinterpret_cast<float4 *>(s + (size_t)(e_dst + k) * C)[c4] = res_ld_peer(s + (size_t)(e_src + k) * C + c4 * 4, crank ^ 1u);
      }
      res_cluster_sync();  // the peer has read my rows: from here on the two CTAs run free until the next exchange
    }
    res_fill_pads(s, lo, N, F, tid);
    __syncthreads();
    res_block_sums<true>(s, ext, bs1, bs2, b0, c0);
    __syncthreads();
    // ---- stage 2 (processing.py:318-322): one thread per (output block, column pair), rows read only
    for (int b = b0; b < nbo; b += kResBT) {
      double w1x = 0.0, w1y = 0.0, w2x = 0.0, w2y = 0.0;
#pragma unroll
      for (int i = 0; i < kResNB; ++i) {
        const double2 t1 = *reinterpret_cast<const double2 *>(bs1 + (b + i) * C + c0);
        const double2 t2 = *reinterpret_cast<const double2 *>(bs2 + (b + i) * C + c0);
        w1x += t1.x; w1y += t1.y;
        w2x += t2.x; w2y += t2.y;
      }
      const int q0 = b * G, q1 = min(q0 + G, N);
      const float *pout = s + (size_t)q0 * C + c0;
      const float *pc = pout + (size_t)H * C;
      const float *pin = pout + (size_t)kResWin * C;
      {
        const float2 x = *reinterpret_cast<const float2 *>(pin - C);  // ext row q0 + 300: the window's last row
        const double dx = (double)x.x, dy = (double)x.y;
        w1x += dx; w1y += dy;
        w2x = fma(dx, dx, w2x); w2y = fma(dy, dy, w2y);
      }
      const double m0x = w1x * inv_win, m0y = w1y * inv_win;
      const float2 mh = make_float2((float)m0x, (float)m0y);
      const float2 ml = make_float2((float)(m0x - (double)mh.x), (float)(m0y - (double)mh.y));
      const float2 e0 = make_float2((float)(w2x * inv_win), (float)(w2y * inv_win));
      const float2 peps = make_float2(p.eps, p.eps), one = make_float2(1.f, 1.f);
      float2 D = make_float2(0.f, 0.f), Q = make_float2(0.f, 0.f);
      float *dst = p.out + (base + lo + q0) * (long long)C + c0;
#pragma unroll 2
      for (int q = q0; q < q1; ++q) {
        const float2 xc = *reinterpret_cast<const float2 *>(pc);
        const float2 xi = *reinterpret_cast<const float2 *>(pin);
        const float2 xo = *reinterpret_cast<const float2 *>(pout);
        const float2 m = __fadd2_rn(mh, __ffma2_rn(D, iw, ml));
        const float2 v = __ffma2_rn(res_neg(m), m, __ffma2_rn(Q, iw, e0));
        // x / (s + eps) = x r / (1 + eps r), r = 1 / s = rsqrt(v): one MUFU per value and 1 - eps r for the
        // quotient (eps r < 1e-3: the second-order term is below float32).  Near-constant windows
        // (v <= 1e-12, rounding-negative included) and NaN variances take the exact path: v clamped with a
        // select -- fmaxf would swallow the NaN that has to poison the row like the reference
        float2 ov;
        if (v.x > 1e-12f && v.y > 1e-12f) {
          float rx, ry;
          asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rx) : "f"(v.x));
          asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(ry) : "f"(v.y));
          const float2 rr = make_float2(rx, ry);
          ov = __fmul2_rn(__fmul2_rn(xc, rr), __ffma2_rn(rr, res_neg(peps), one));
        } else {
          const float vx = v.x < 1e-30f ? 1e-30f : v.x, vy = v.y < 1e-30f ? 1e-30f : v.y;
          float rx, ry, cx, cy;
          asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rx) : "f"(vx));
          asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(ry) : "f"(vy));
          const float2 rr = make_float2(rx, ry);
          const float2 d = __ffma2_rn(rr, peps, one);
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(cx) : "f"(d.x));
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(cy) : "f"(d.y));
          ov = __fmul2_rn(__fmul2_rn(xc, rr), make_float2(cx, cy));
        }
        *reinterpret_cast<float2 *>(dst) = ov;
        D = __fadd2_rn(D, res_sub(xi, xo));
        Q = __fadd2_rn(Q, __ffma2_rn(xi, xi, res_neg(__fmul2_rn(xo, xo))));
        dst += C; pc += C; pin += C; pout += C;
      }
    }
  }
}

// ------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------
bool cmvnw_resident_supported(int cols, int win, int varnorm) {
  return cols == kResCols && win == kResWin && varnorm != 0;
}
int cmvnw_resident_max_rows() { return kResMaxRows; }

// tiles: two int4 per utterance, {utt, F, 0, 0} and {first global row (lo, hi), 0, 0}.
// cluster = 1: utterances of <= kResMaxRows rows, one CTA each; cluster = 2: utterances of <= 2 kResMaxRows rows
// (and >= 600, so that each half holds its 150 boundary rows), a cluster of two CTAs each.
int launch_cmvnw_resident(spb_context *ctx, const float *in, float *out, const int4 *tiles, int64_t num_tiles, int cluster) {
  if (num_tiles <= 0) return SPB_OK;
  int *counters = (int *)scratch(ctx, 7, 256);
  if (!counters) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemsetAsync(counters + 3, 0, 4, ctx->stream));
  CmResParams p;
  p.in = in;
  p.out = out;
  p.tiles = tiles;
  p.num_tiles = (int)num_tiles;
  p.counter = counters + 3;
  p.eps = 9.313225746154785e-10f;  // 2^-30 (processing.py:292)
  void (*kern)(const CmResParams) = (cluster == 2) ? cmvnw_resident_kernel<2> : cmvnw_resident_kernel<1>;
  int per_sm = 0;
  int rc = configure_kernel(ctx, (const void *)kern, kResThreads, kResSmem, &per_sm);
  if (rc) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_UNSUPPORTED, "resident cmvnw kernel does not fit in shared memory");
  const size_t tb = timing_begin(ctx);
  if (cluster == 2) {
    const int clusters = cmvnw_resident_clusters(ctx);
    if (clusters < 1) return fail(ctx, SPB_ERR_UNSUPPORTED, "resident cmvnw: the device takes no 2-CTA clusters of this kernel");
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)(2 * std::min<long long>(num_tiles, clusters)), 1, 1);
    lc.blockDim = dim3(kResThreads, 1, 1);
    lc.dynamicSmemBytes = kResSmem;
    lc.stream = ctx->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    lc.attrs = at;
    lc.numAttrs = 1;
    SPB_CUDA(ctx, cudaLaunchKernelEx(&lc, kern, p));
  } else {
    const long long grid = std::min<long long>(ctx->sm_count, num_tiles);
    kern<<<(unsigned)grid, kResThreads, kResSmem, ctx->stream>>>(p);
  }
  timing_end(ctx, 5, tb);
  SPB_LAUNCH_CHECK(ctx, "cmvnw_resident_kernel");
  return SPB_OK;
}

// how many 2-CTA clusters of the resident kernel the device runs at once (0: none); asked once
int cmvnw_resident_clusters(spb_context *ctx) {
  static int cached = -1;
  if (cached >= 0) return cached;
  int per_sm = 0;
  if (configure_kernel(ctx, (const void *)cmvnw_resident_kernel<2>, kResThreads, kResSmem, &per_sm) != SPB_OK) return 0;
  cudaLaunchConfig_t qc = {};
  qc.gridDim = dim3((unsigned)(ctx->sm_count & ~1), 1, 1);
  qc.blockDim = dim3(kResThreads, 1, 1);
  qc.dynamicSmemBytes = kResSmem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  qc.attrs = at;
  qc.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, (const void *)cmvnw_resident_kernel<2>, &qc) != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  cached = n;
  return cached;
}

}  // namespace spb
