// Tensor-core variant of the lmfe feature kernel (stack_frames + |rfft|^2/NFFT + mel + log):
// the transform runs on tcgen05 as a hop-block DFT.
//
// Replaces, for the batch paths at L == 2 S (speechpy defaults: 20 ms frames, 10 ms stride):
//   processing.py:93-139   stack_frames        frame f = hops f and f+1 (a hop = S samples)
//   processing.py:158-174  |rfft|^2 / NFFT     X_f[k] = A_f[k] + e^{-2 pi i S k / NFFT} A_{f+1}[k]
//   feature.py:199-217     P . fb^T, zero_handling, log
//
// A_h[k] = sum_{n<S} x[hS+n] e^{-2 pi i k n / NFFT} is ONE dense product per hop, shared by the
// two frames the hop belongs to: rows = hops (the sample arena itself, viewed S samples per
// row), columns = (cos, -sin) of the bins the filterbank reads (speechpy's (K+1) bin mapping
// stops at bin 128; up to 124 bins = 248 columns), K = S = 160.
//
// fp32 accuracy from fp16 tensor cores: x 2^15 = xh + xl, W = Wh + Wl (fp16 each; for int16
// PCM audio the sample split is exact).  Per tile the 20 small products xl.Wh, xh.Wl go first
// into the empty fp32 TMEM accumulator, then the 10 xh.Wh products: TMEM accumulation
// truncates, and only those last 10 additions happen at full magnitude (measured: the
// interleaved order is 3.5x noisier).  |x| >= 2 (or NaN/inf) cannot be split this way: the
// converter raises a flag and the launcher's guarded CUDA-core FFT kernel redoes the launch.
//
// One persistent CTA per SM, 17 warps:
//   warps 0..7   epilogue: quadrant q = warp % 4 owns hop rows 32q..32q+31 (31 frames), the
//                two warps of a quadrant split the filterbank segments; per 16-bin chunk:
//                tcgen05.ld, a branch-free pass (neighbour hop by shuffle, complex MAC with the
//                hop rotation as immediates, |X|^2), then the segment-form mel sums with a
//                flush (lg2, one predicated 16-byte store every fourth filter) where a segment ends
//   warp  8      MMA issuer (one elected thread), M128 N256 K16 kind::f16, two TMEM accumulators
//   warps 9..16  converters: 128-bit sample loads two k-steps ahead (+ L2 bulk prefetch two tiles
//                ahead), split, st.shared into the no-swizzle K-major operand buffers
//                (xh: whole tile, xl: ring of 7 k-steps)
// The DFT matrix (2 x 77.5 KB) stays resident in shared memory; 223 KB of the 227 KB are in use.
#include <cuda.h>
#include <cuda_fp16.h>
#include <math.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <vector>

#include "ptx_helpers.cuh"
#include "spb_internal.cuh"

namespace spb {

constexpr int kTcHop = 160;                       // samples per hop row == K of the product
constexpr int kTcKSteps = kTcHop / 16;            // 10 MMA k-steps
constexpr int kTcRows = 128;                      // hop rows per tile (MMA M)
constexpr int kTcN = 256;                         // MMA N
constexpr int kTcNAlloc = 248;                    // stored columns: (re, im) of up to 124 bins
constexpr int kTcMaxBins = kTcNAlloc / 2;
constexpr int kTcLoSlots = 7;                     // xl ring depth (k-steps)
#ifndef SPB_TC_AHEAD
#define SPB_TC_AHEAD 2   // must divide 10.  Measured 1.119 / 1.126 / 1.140 ms at 1 / 2 / 5: the L2 bulk prefetch covers DRAM, registers matter more
#endif
constexpr int kTcAhead = SPB_TC_AHEAD;            // converter load distance (k-steps)
constexpr int kTcOutPitch = 40;                   // at most this many filters
constexpr uint32_t kTcBLbo = kTcNAlloc * 16;             // 3 968 B between the k-chunks of W
constexpr uint32_t kTcBMat = (kTcHop / 8) * kTcBLbo;     // 79 360 B: [k/8][col][8 halves]
constexpr uint32_t kTcAMat = 2 * kTcRows * 16;           // 4 096 B per k-step: [2][row][8 halves]
constexpr uint32_t kTcOffHi = 2 * kTcBMat;               // xh: 10 k-steps
constexpr uint32_t kTcOffLo = kTcOffHi + kTcKSteps * kTcAMat;
constexpr uint32_t kTcOffBars = kTcOffLo + kTcLoSlots * kTcAMat;
constexpr uint32_t kTcSmem = kTcOffBars + 256;
static_assert(kTcSmem <= 232448, "features_tc: shared memory budget");
constexpr int kTcEpiWarps = 8, kTcMmaWarp = 8, kTcCvtWarp0 = 9, kTcCvtWarps = 8;
constexpr int kTcThreads = (kTcCvtWarp0 + kTcCvtWarps) * 32;
constexpr float kTcInScale = 32768.f;             // 2^15
// D = F32 (bit 4), A = B = F16 (0), K-major both, N >> 3 at bit 17, M >> 4 at bit 24
constexpr uint32_t kTcIdesc = (1u << 4) | ((uint32_t)(kTcN >> 3) << 17) | ((uint32_t)(kTcRows >> 4) << 24);
static constexpr float kTcEps = 2.220446049250313e-16f;  // functions.py:62

// per bin {rise weight, fall weight, cos, sin of the hop rotation}, one table per epilogue half
// (bins outside a half's segments carry zero weights).  Constant bank, indexed by the chunk.
struct TcConst {
  float4 tab[2][128];
};
__constant__ TcConst c_tc;

struct TcHalf {
  uint32_t endmask[4];  // bit b: bin b closes a segment of this half
  int ch_begin, ch_end; // 16-bin chunks [begin, end) this half reads
  int seg0;             // index of the first segment it walks
  int out_lo;           // filters below this index belong to the other half
};

struct TcParams {
  const float *samples;
  const TcGroup *groups;     // 4 per tile
  long long num_tiles;
  const __half *w_hi, *w_lo; // DFT matrix in the shared-memory layout (kTcBMat bytes each)
  TcHalf half[2];
  int M;
  float scale;               // 1 / (NFFT * 2^30)
  float ln_scale;            // ln(scale)
  int mfe;                   // 1: no log
  int vec4;                  // both halves emit multiples of four filters and rows are 16-byte aligned
  float *out;
  int *flag;                 // set when a sample cannot be split (|x| >= 2, NaN, inf)
  const float *zeros;        // 1 KB of zeros: the hop rows past the end of a group read these
  float4 *hopsum;            // optional [tiles * 128]: per hop row {sum x^2, sum x, sum (-1)^n x, 0} for the mfcc energy
};

// K-major operand without swizzle: core matrix = 8 rows x 16 bytes stored contiguously,
// SBO = 128 B between 8-row groups, LBO = distance between the two core matrices of a k-step.
__device__ __forceinline__ uint64_t umma_desc_plain(uint32_t saddr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(128 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// mbarrier wait that lets the hardware park the warp (suspend-time hint) instead of spinning:
// 17 warps polling barriers took ~1/3 of the issue slots of the kernel
__device__ __forceinline__ void mbar_wait_park(uint64_t *bar, uint32_t parity) {
  const uint32_t a = smem_u32(bar);
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "PARK_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
      "@p bra PARK_DONE;\n\t"
      "bra PARK_LOOP;\n\t"
      "PARK_DONE:\n\t"
      "}" ::"r"(a), "r"(parity), "r"(20000u)
      : "memory");
}
// wait of a warp that is normally AHEAD of its producer (the converters): back off between polls
__device__ __forceinline__ void mbar_wait_nap(uint64_t *bar, uint32_t parity) {
  const uint32_t a = smem_u32(bar);
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra NAP_DONE;\n\t"
      "NAP_LOOP:\n\t"
      "nanosleep.u32 96;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra NAP_DONE;\n\t"
      "bra NAP_LOOP;\n\t"
      "NAP_DONE:\n\t"
      "}" ::"r"(a), "r"(parity)
      : "memory");
}
// four scaled samples -> fp16 hi and lo halves (two 32-bit words each)
// bad: bit 15 / 31 set when a half came out as inf or NaN (|x| >= 2, NaN, inf)
__device__ __forceinline__ void split4(const float4 x, uint2 &hi, uint2 &lo, uint32_t &bad) {
  const float a = x.x * kTcInScale, b = x.y * kTcInScale, c = x.z * kTcInScale, d = x.w * kTcInScale;
  const __half2 h0 = __floats2half2_rn(a, b), h1 = __floats2half2_rn(c, d);
  const float2 f0 = __half22float2(h0), f1 = __half22float2(h1);
  const __half2 l0 = __floats2half2_rn(a - f0.x, b - f0.y), l1 = __floats2half2_rn(c - f1.x, d - f1.y);
  hi.x = *reinterpret_cast<const uint32_t *>(&h0);
  hi.y = *reinterpret_cast<const uint32_t *>(&h1);
  lo.x = *reinterpret_cast<const uint32_t *>(&l0);
  lo.y = *reinterpret_cast<const uint32_t *>(&l1);
  bad |= ((hi.x & 0x7fff7fffu) + 0x04000400u) | ((hi.y & 0x7fff7fffu) + 0x04000400u);
}

// cos / sin of 2 pi m / 16: the hop rotation e^{-2 pi i S k / NFFT} = e^{-2 pi i (5k mod 16) / 16}
// at S = 160, NFFT = 512 has period 16 in k, so with kmin mod 16 known at compile time it is an
// immediate operand of the unrolled bin loop.
__device__ constexpr float kCos16[16] = {1.f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f,
                                         0.f, -0.38268343236508977f, -0.70710678118654752f, -0.92387953251128674f,
                                         -1.f, -0.92387953251128674f, -0.70710678118654752f, -0.38268343236508977f,
                                         0.f, 0.38268343236508977f, 0.70710678118654752f, 0.92387953251128674f};
__device__ constexpr float kSin16[16] = {0.f, 0.38268343236508977f, 0.70710678118654752f, 0.92387953251128674f,
                                         1.f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f,
                                         0.f, -0.38268343236508977f, -0.70710678118654752f, -0.92387953251128674f,
                                         -1.f, -0.92387953251128674f, -0.70710678118654752f, -0.38268343236508977f};

// One tile of one epilogue warp: lane = hop row of the quadrant, frame = hops (lane, lane + 1).
// The chunk loop is NOT unrolled: the straight-line version (123 bins x ~40 instructions, twice)
// ran at the instruction-fetch rate (ncu: stall_no_inst on every line).
// KM = kmin mod 16 (rotation as immediates) or -1 (rotation from the table).
// FAST: log output in 16-byte runs (vec4 && !mfe known at compile time, flush without inner branches).
template <int KM, bool FAST>
__device__ __forceinline__ void tc_epilogue_tile(const TcParams &p, const int half, const uint32_t taddr,
                                                 float *pout, const bool row_ok) {
  const int ch_begin = half ? p.half[1].ch_begin : p.half[0].ch_begin;
  const int ch_end = half ? p.half[1].ch_end : p.half[0].ch_end;
  const uint32_t m0 = half ? p.half[1].endmask[0] : p.half[0].endmask[0];
  const uint32_t m1 = half ? p.half[1].endmask[1] : p.half[0].endmask[1];
  const uint32_t m2 = half ? p.half[1].endmask[2] : p.half[0].endmask[2];
  const uint32_t m3 = half ? p.half[1].endmask[3] : p.half[0].endmask[3];
  float rise = 0.f, fall = 0.f, prev_u = 0.f;
  // the first flush of a half only latches U of its first segment (filter j = U[j] + D[j+1])
  int skip = 1;
  // outputs leave in runs of four filters (one 16-byte store per lane; the rows of a warp are
  // 4 M bytes apart, L2 merges the sectors) when the half's filter range allows it
  const bool vec4 = FAST || p.vec4 != 0;
  float o0 = 0.f, o1 = 0.f, o2 = 0.f;
  int cnt = 0, left = 5, q4 = 0;
  const float ln_scale = p.ln_scale, scale = p.scale;
  const bool mfe = !FAST && p.mfe != 0;
#pragma unroll 1
  for (int ch = ch_begin; ch < ch_end; ++ch) {
    uint32_t v[32];
    tmem_ld32(taddr + ch * 32, v);  // load + wait::ld in one asm: v must not be read before the wait
    const uint32_t mw = (ch < 4) ? ((ch < 2) ? m0 : m1) : ((ch < 6) ? m2 : m3);
    const uint32_t mask = mw >> ((ch & 1) * 16);
    const float4 *tb = &c_tc.tab[half][ch * 16];
    // pass A, branch-free so the 16 bins overlap: X = A_own + c A_next, |X|^2
    float pw[16];
    float2 wgt[16];  // weights fetched here too: behind the flush branches every LDC latency is exposed
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      wgt[j] = *reinterpret_cast<const float2 *>(&tb[j]);
      float cr, ci;
      if (KM >= 0) {
        cr = kCos16[(5 * (KM + j)) & 15];
        ci = -kSin16[(5 * (KM + j)) & 15];
      } else {
        cr = tb[j].z;
        ci = tb[j].w;
      }
      const float a_re = __uint_as_float(v[2 * j]), a_im = __uint_as_float(v[2 * j + 1]);
      const float b_re = __shfl_down_sync(0xffffffffu, a_re, 1);
      const float b_im = __shfl_down_sync(0xffffffffu, a_im, 1);
      const float xr = fmaf(-ci, b_im, fmaf(cr, b_re, a_re));
      const float xi = fmaf(ci, b_re, fmaf(cr, b_im, a_im));
      pw[j] = fmaf(xr, xr, xi * xi);
    }
    // pass B: segment sums, a flush where the mask says a segment ends
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      rise = fmaf(wgt[j].x, pw[j], rise);
      fall = fmaf(wgt[j].y, pw[j], fall);
      if ((mask >> j) & 1u) {
        if (FAST) {
          // straight-line flush.  `left` counts down to the next 16-byte store; it starts at 5
          // because the first flush of a half only latches U of its first segment (its value
          // shifts through o2..o0 and is gone before the first store)
          const float a = prev_u + fall;  // unscaled (x 2^30 NFFT): never denormal for real input
          float lg;
          asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(a));
          float r = fmaf(lg, 0.69314718055994531f, ln_scale);
          r = (a == 0.f) ? -36.04365338911715f : r;  // functions.zero_handling (feature.py:217)
          --left;
          if (left == 0 && row_ok) reinterpret_cast<float4 *>(pout)[q4] = make_float4(o0, o1, o2, r);
          q4 += (left == 0) ? 1 : 0;
          left = (left == 0) ? 4 : left;
          o0 = o1; o1 = o2; o2 = r;
        } else
        if (skip == 0) {
          const float a = prev_u + fall;  // unscaled (x 2^30 NFFT): never denormal for real input
          float lg;
          asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(a));
          // functions.zero_handling (feature.py:217): exact zero -> eps
          float r = mfe ? a * scale : fmaf(lg, 0.69314718055994531f, ln_scale);
          if (a == 0.f) r = mfe ? kTcEps : -36.04365338911715f;
          if (vec4) {
            if ((cnt & 3) == 3) {
              if (row_ok) *reinterpret_cast<float4 *>(pout) = make_float4(o0, o1, o2, r);
              pout += 4;
            }
            o0 = o1; o1 = o2; o2 = r;
            ++cnt;
          } else {
            if (row_ok) *pout = r;
            ++pout;
          }
        }
        skip = 0;
        prev_u = rise;
        rise = 0.f;
        fall = 0.f;
      }
    }
  }
}

template <int KM, bool FAST, bool SUMS>
__global__ void __launch_bounds__(kTcThreads, 1) features_tc_kernel(const TcParams p) {
  extern __shared__ __align__(128) unsigned char tc_smem[];
  unsigned char *base = tc_smem;
  uint64_t *bars = reinterpret_cast<uint64_t *>(base + kTcOffBars);
  uint64_t *full = bars;                         // [10] k-step s of the tile converted (8 converter warps)
  uint64_t *hi_empty = bars + 10;                // [10] xh of k-step s consumed (tcgen05.commit)
  uint64_t *lo_empty = bars + 20;                // [3]  xl ring slot consumed
  uint64_t *acc_full = bars + 20 + kTcLoSlots;   // [2]
  uint64_t *acc_empty = acc_full + 2;            // [2] 8 epilogue warps
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + 2);
  static_assert((20 + kTcLoSlots + 4) * 8 + 4 <= 256, "barrier block");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(base);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTcKSteps; ++s) {
      mbar_init(&full[s], kTcCvtWarps);
      mbar_init(&hi_empty[s], 1);
    }
    for (int s = 0; s < kTcLoSlots; ++s) mbar_init(&lo_empty[s], 1);
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], kTcEpiWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kTcMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  {  // resident DFT matrix
    const uint4 *gh = reinterpret_cast<const uint4 *>(p.w_hi);
    const uint4 *gl = reinterpret_cast<const uint4 *>(p.w_lo);
    uint4 *sh = reinterpret_cast<uint4 *>(base);
    uint4 *sl = reinterpret_cast<uint4 *>(base + kTcBMat);
    for (uint32_t i = threadIdx.x; i < kTcBMat / 16; i += kTcThreads) {
      sh[i] = __ldg(gh + i);
      sl[i] = __ldg(gl + i);
    }
    fence_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= kTcCvtWarp0) {
    // ===== converters =====
    const int c = (warp - kTcCvtWarp0) * 32 + lane;  // 0..255
    const int q = c & 3;                             // which 4 samples of the 16 of a k-step
    const int rA = c >> 2, rB = 64 + (c >> 2);       // the two hop rows this thread feeds
    const uint32_t dA = (uint32_t)((q >> 1) * (kTcRows * 16) + rA * 16 + (q & 1) * 8);
    const uint32_t dB = (uint32_t)((q >> 1) * (kTcRows * 16) + rB * 16 + (q & 1) * 8);
    uint32_t bad = 0;
    auto row_ptr = [&](long long t, int r) -> const float * {
      if (t >= p.num_tiles) return p.zeros;
      const TcGroup g = p.groups[t * 4 + (r >> 5)];
      const int i = r & 31;
      return (g.nf > 0 && i <= g.nf) ? p.samples + g.gstart + (long long)i * kTcHop + q * 4 : p.zeros;
    };
    long long t = blockIdx.x;
    const float *pA = row_ptr(t, rA), *pB = row_ptr(t, rB);
    float4 bufA[kTcAhead], bufB[kTcAhead];
#pragma unroll
    for (int s = 0; s < kTcAhead; ++s) {
      bufA[s] = __ldg(reinterpret_cast<const float4 *>(pA + s * 16));
      bufB[s] = __ldg(reinterpret_cast<const float4 *>(pB + s * 16));
    }
    uint32_t it = 0, slot = 0, lo_par = 1;  // tiles produced so far, xl ring slot and its wait parity
    constexpr bool want_sums = SUMS;  // compile-time: the lmfe path pays nothing for the mfcc sums
    for (; t < p.num_tiles; t += gridDim.x, ++it) {
      const float *nA = row_ptr(t + gridDim.x, rA), *nB = row_ptr(t + gridDim.x, rB);
      float a2 = 0.f, a0 = 0.f, aa = 0.f, b2 = 0.f, b0 = 0.f, ba = 0.f;  // hop sums of rows rA, rB (this lane's samples)
      if (c < 4) {
        // pull the samples of the tile after next into L2 with the TMA engine (one 20 KB run per group)
        const long long tp = t + 2 * (long long)gridDim.x;
        if (tp < p.num_tiles) {
          const TcGroup g = p.groups[tp * 4 + c];
          if (g.nf > 0) bulk_prefetch_l2(p.samples + g.gstart, (uint32_t)(g.nf + 1) * (kTcHop * 4));
        }
      }
#pragma unroll
      for (int s = 0; s < kTcKSteps; ++s) {
        const float4 xa = bufA[s % kTcAhead], xb = bufB[s % kTcAhead];
        if (s + kTcAhead < kTcKSteps) {
          bufA[s % kTcAhead] = __ldg(reinterpret_cast<const float4 *>(pA + (s + kTcAhead) * 16));
          bufB[s % kTcAhead] = __ldg(reinterpret_cast<const float4 *>(pB + (s + kTcAhead) * 16));
        } else {
          bufA[s % kTcAhead] = __ldg(reinterpret_cast<const float4 *>(nA + (s + kTcAhead - kTcKSteps) * 16));
          bufB[s % kTcAhead] = __ldg(reinterpret_cast<const float4 *>(nB + (s + kTcAhead - kTcKSteps) * 16));
        }
        if (want_sums) {
          // frame energy of the mfcc path (feature.py:202 by Parseval): the three sums per hop ride along
          a2 = fmaf(xa.x, xa.x, fmaf(xa.y, xa.y, fmaf(xa.z, xa.z, fmaf(xa.w, xa.w, a2))));
          a0 += (xa.x + xa.y) + (xa.z + xa.w);
          aa += (xa.x - xa.y) + (xa.z - xa.w);
          b2 = fmaf(xb.x, xb.x, fmaf(xb.y, xb.y, fmaf(xb.z, xb.z, fmaf(xb.w, xb.w, b2))));
          b0 += (xb.x + xb.y) + (xb.z + xb.w);
          ba += (xb.x - xb.y) + (xb.z - xb.w);
        }
        uint2 ha, la, hb, lb;
        split4(xa, ha, la, bad);
        split4(xb, hb, lb, bad);
        // keep the conversion ahead of the waits (the compiler otherwise sinks it below them and
        // every k-step pays wait + conversion latency in series)
        asm volatile("" ::"r"(ha.x), "r"(ha.y), "r"(la.x), "r"(la.y), "r"(hb.x), "r"(hb.y), "r"(lb.x), "r"(lb.y));
        if (s == 0 || s == kTcKSteps / 2) mbar_wait_nap(&hi_empty[s ? 1 : 0], (it & 1) ^ 1);
        mbar_wait_nap(&lo_empty[slot], lo_par);
        unsigned char *sh = base + kTcOffHi + s * kTcAMat;
        unsigned char *sl = base + kTcOffLo + slot * kTcAMat;
        *reinterpret_cast<uint2 *>(sh + dA) = ha;
        *reinterpret_cast<uint2 *>(sl + dA) = la;
        *reinterpret_cast<uint2 *>(sh + dB) = hb;
        *reinterpret_cast<uint2 *>(sl + dB) = lb;
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[s]);
        if (++slot == kTcLoSlots) { slot = 0; lo_par ^= 1; }
      }
      if (want_sums) {
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {  // the four lanes of a row
          a2 += __shfl_xor_sync(0xffffffffu, a2, o);
          a0 += __shfl_xor_sync(0xffffffffu, a0, o);
          aa += __shfl_xor_sync(0xffffffffu, aa, o);
          b2 += __shfl_xor_sync(0xffffffffu, b2, o);
          b0 += __shfl_xor_sync(0xffffffffu, b0, o);
          ba += __shfl_xor_sync(0xffffffffu, ba, o);
        }
        if (q == 0) {
          p.hopsum[t * kTcRows + rA] = make_float4(a2, a0, aa, 0.f);
          p.hopsum[t * kTcRows + rB] = make_float4(b2, b0, ba, 0.f);
        }
      }
      pA = nA;
      pB = nB;
    }
    if (bad & 0x80008000u) *p.flag = 1;
  } else if (warp == kTcMmaWarp) {
    // ===== MMA issuer =====
    if (elect_one()) {
      const uint32_t b_hi = sbase, b_lo = sbase + kTcBMat;
      uint32_t gs = 0, it = 0;
      for (long long t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++it) {
        const uint32_t acc = it & 1;
        mbar_wait_park(&acc_empty[acc], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * kTcN;
        // small terms first, into the empty accumulator
#pragma unroll
        for (int s = 0; s < kTcKSteps; ++s, ++gs) {
          const uint32_t slot = gs % kTcLoSlots;
          mbar_wait_park(&full[s], it & 1);
          tc_fence_after();
          const uint64_t da_hi = umma_desc_plain(sbase + kTcOffHi + s * kTcAMat, kTcRows * 16);
          const uint64_t da_lo = umma_desc_plain(sbase + kTcOffLo + slot * kTcAMat, kTcRows * 16);
          const uint64_t db_hi = umma_desc_plain(b_hi + s * 2 * kTcBLbo, kTcBLbo);
          const uint64_t db_lo = umma_desc_plain(b_lo + s * 2 * kTcBLbo, kTcBLbo);
          umma_f16(tmem_d, da_lo, db_hi, kTcIdesc, s ? 1u : 0u);
          umma_f16(tmem_d, da_hi, db_lo, kTcIdesc, 1u);
          tc_commit(&lo_empty[slot]);
        }
        // xh k-steps are handed back in two batches (a tcgen05.commit per MMA costs more than it hides)
#pragma unroll
        for (int s = 0; s < kTcKSteps; ++s) {
          const uint64_t da_hi = umma_desc_plain(sbase + kTcOffHi + s * kTcAMat, kTcRows * 16);
          const uint64_t db_hi = umma_desc_plain(b_hi + s * 2 * kTcBLbo, kTcBLbo);
          umma_f16(tmem_d, da_hi, db_hi, kTcIdesc, 1u);
          if (s == kTcKSteps / 2 - 1) tc_commit(&hi_empty[0]);
        }
        tc_commit(&hi_empty[1]);
        tc_commit(&acc_full[acc]);
      }
    }
  } else {
    // ===== epilogue: quadrant q = hop rows 32q..32q+31 = group q of the tile; two warps per quadrant =====
    const int quad = warp & 3, half = warp >> 2;
    const int M = p.M;
    const int col0 = half ? p.half[1].out_lo : 0;
    uint32_t it = 0;
    for (long long t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++it) {
      const TcGroup g = p.groups[t * 4 + quad];
      const uint32_t acc = it & 1;
      mbar_wait_park(&acc_full[acc], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * kTcN;
      tc_epilogue_tile<KM, FAST>(p, half, taddr, p.out + (g.row0 + lane) * M + col0, lane < g.nf);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[acc]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kTcMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

#include "feat_tc2.cuh"

// ------------------------------------------------------------------------------------ host
struct TcConstOwner {
  std::mutex mu;
  const spb_plan *plan = nullptr;
  uint64_t serial = 0;          // plans are heap objects: an address can come back with other tables
  int variant = 0;              // 0: the run-time split tables, 1: the static split of the default bank
  cudaEvent_t last_use = nullptr;
};
static TcConstOwner g_tc_owner[64];
static std::atomic<uint64_t> g_tc_serial{1};

// Plan side: DFT matrix (fp16 hi/lo, shared-memory layout), per-half bin tables and ranges.
// Covered shape: NFFT 512, stride 160, frame 320, a segment-form bank over <= 124 bins and
// <= 40 filters.
int build_tc_plan(spb_context *ctx, spb_plan *pl, const float4 *seg_rd) {
  pl->tc_ok = false;
  const spb_feature_params &q = pl->p;
  if (q.nfft != 512 || q.frame_stride != kTcHop || q.frame_len != 2 * kTcHop) return SPB_OK;
  if (pl->seg_nbins < 1 || pl->seg_nbins > kTcMaxBins || q.num_filters < 1 || q.num_filters > kTcOutPitch) return SPB_OK;
  const int nb = pl->seg_nbins, kmin = pl->seg_kmin, M = q.num_filters;
  const double two_pi = 6.283185307179586476925286766559;
  std::vector<__half> hi((size_t)kTcBMat / 2), lo((size_t)kTcBMat / 2);
  for (int n = 0; n < kTcHop; ++n)
    for (int col = 0; col < kTcNAlloc; ++col) {
      double w = 0.0;
      const int j = col >> 1;
      if (j < nb) {
        const int r = (int)(((long long)(kmin + j) * n) % q.nfft);
        const double ang = two_pi * (double)r / (double)q.nfft;
        w = (col & 1) ? -sin(ang) : cos(ang);
      }
      const __half h = __float2half_rn((float)w);
      const __half l = __float2half_rn((float)(w - (double)__half2float(h)));
      const size_t idx = (size_t)(n >> 3) * (kTcNAlloc * 8) + (size_t)col * 8 + (n & 7);
      hi[idx] = h;
      lo[idx] = l;
    }
  // segments: seg_first[s] / seg_last[s] bins of segment s = 0..M
  std::vector<int> seg_first, seg_last;
  {
    int first = 0;
    for (int j = 0; j < nb; ++j)
      if (seg_rd[j].z != 0.f) {
        seg_first.push_back(first);
        seg_last.push_back(j);
        first = j + 1;
      }
  }
  if ((int)seg_first.size() != M + 1) return SPB_OK;  // not the structure the epilogue walks
  // split the segments between the two warps of a quadrant: half 0 walks segments 0..s1 (filters
  // 0..s1-1), half 1 segments s1..M (filters s1..M-1; bins before segment s1 carry zero weights).
  // Pick the s1 with the best balance, a multiple of four when M is (16-byte output stores).
  const int nch = (nb + 15) / 16;
  const bool want4 = (M % 4) == 0;
  int best_s1 = -1;
  long best_cost = -1;
  for (int s1 = 1; s1 <= M - 1; ++s1) {
    if (want4 && (s1 % 4) != 0) continue;
    const long c0 = (long)(seg_last[s1] / 16 + 1) * 16 * 13 + (long)(s1 + 1) * 10;
    const long c1 = (long)(nch - seg_first[s1] / 16) * 16 * 13 + (long)(M + 1 - s1) * 10;
    const long cost = std::max(c0, c1);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_s1 = s1; }
  }
  pl->tc_vec4 = want4;
  // tables of one split: per half and bin {rise, fall, cos, sin of the hop rotation}, segment-end masks,
  // {first chunk, end chunk, first segment, first filter}
  auto fill_tables = [&](int s1, std::vector<float> &tab, uint32_t (&endmask)[2][4], int (&hf_meta)[2][4],
                         int (&lo_bin)[2], int (&hi_bin)[2]) {
    tab.assign(2 * 128 * 4, 0.f);
    for (int hf = 0; hf < 2; ++hf)
      for (int i = 0; i < 4; ++i) endmask[hf][i] = 0;
    lo_bin[0] = 0; lo_bin[1] = nb;
    hi_bin[0] = nb - 1; hi_bin[1] = nb - 1;
    if (s1 > 0) {
      hi_bin[0] = seg_last[s1];
      lo_bin[1] = seg_first[s1];
      hf_meta[0][0] = 0; hf_meta[0][1] = seg_last[s1] / 16 + 1; hf_meta[0][2] = 0; hf_meta[0][3] = 0;
      hf_meta[1][0] = seg_first[s1] / 16; hf_meta[1][1] = nch; hf_meta[1][2] = s1; hf_meta[1][3] = s1;
    } else {
      hf_meta[0][0] = 0; hf_meta[0][1] = nch; hf_meta[0][2] = 0; hf_meta[0][3] = 0;
      hf_meta[1][0] = 0; hf_meta[1][1] = 0; hf_meta[1][2] = 0; hf_meta[1][3] = M;
    }
    for (int hf = 0; hf < 2; ++hf)
      for (int j = 0; j < nb; ++j) {
        const int r = (int)(((long long)(kmin + j) * q.frame_stride) % q.nfft);
        const double ang = two_pi * (double)r / (double)q.nfft;
        // X_f = A_f + (cos - i sin) A_{f+1}
        float *e = &tab[((size_t)hf * 128 + j) * 4];
        const bool in = (j >= lo_bin[hf] && j <= hi_bin[hf]);
        e[0] = in ? seg_rd[j].x : 0.f;
        e[1] = in ? seg_rd[j].y : 0.f;
        e[2] = (float)cos(ang);
        e[3] = (float)-sin(ang);
        if (in && seg_rd[j].z != 0.f) endmask[hf][j >> 5] |= 1u << (j & 31);
      }
  };
  int lo_bin[2], hi_bin[2];
  fill_tables(best_s1, pl->h_tc_tab, pl->tc_endmask, pl->tc_half, lo_bin, hi_bin);
  // ---- second-generation kernel (feat_tc2.cuh): folded DFT matrices C = cos, S = -sin of
  // th_k (m + 1/2) = 2 pi k (2m + 1) / (2 NFFT), m < 80, fp16 hi + lo, [k/8][col][8 halves] each
  std::vector<__half> w2((size_t)4 * kT2BMat / 2);
  for (int m = 0; m < kT2K; ++m)
    for (int col = 0; col < kT2N; ++col) {
      double c = 0.0, sn = 0.0;
      if (col < nb) {
        const long long r = ((long long)(kmin + col) * (2 * m + 1)) % (2LL * q.nfft);
        const double ang = two_pi * (double)r / (2.0 * (double)q.nfft);
        c = cos(ang);
        sn = -sin(ang);
      }
      const size_t idx = (size_t)(m >> 3) * (kT2N * 8) + (size_t)col * 8 + (m & 7);
      const __half ch = __float2half_rn((float)c), sh = __float2half_rn((float)sn);
      w2[idx] = ch;
      w2[(size_t)kT2BMat / 2 + idx] = __float2half_rn((float)(c - (double)__half2float(ch)));
      w2[(size_t)kT2BMat + idx] = sh;
      w2[(size_t)3 * kT2BMat / 2 + idx] = __float2half_rn((float)(sn - (double)__half2float(sh)));
    }
  // the flagged-filter bit set of the epilogue holds 31 filters per half
  const int n0 = best_s1 > 0 ? best_s1 : M, n1 = best_s1 > 0 ? M - best_s1 : 0;
  pl->tc2_ok = n0 <= 31 && n1 <= 31;
  // the speechpy default bank takes the compile-time epilogue: its own split (kT2Parts parts), checked here
  // against the segments derived from the caller's dense bank; its table is the unmasked one
  pl->tc2_static = false;
  {
    const T2Static &st = kT2StaticHost;
    if (M == st.M && kmin == st.kmin && nb == st.nb) {
      bool same = true;
      for (int pt = 0; pt < kT2Parts && same; ++pt) {
        const int lo = (pt == 0) ? 0 : seg_first[st.sp[pt]];
        const int hi = (pt == kT2Parts - 1) ? nb - 1 : seg_last[st.sp[pt + 1]];
        uint32_t em[4] = {0, 0, 0, 0};
        for (int sg = st.sp[pt]; sg <= st.sp[pt + 1]; ++sg) em[seg_last[sg] >> 5] |= 1u << (seg_last[sg] & 31);
        same = lo == st.lo[pt] && hi == st.hi[pt] && st.out_lo[pt] == st.sp[pt] && st.nout[pt] == st.sp[pt + 1] - st.sp[pt];
        for (int i = 0; i < 4; ++i) same = same && em[i] == st.endmask[pt][i];
      }
      int lo_s[2], hi_s[2];
      fill_tables(0, pl->h_tc_tab_s, pl->tc_endmask_s, pl->tc_half_s, lo_s, hi_s);  // s1 = 0: half 0 spans every bin
      pl->tc2_static = same;
    }
  }
  cudaError_t e;
  if ((e = cudaMalloc(&pl->d_tc_whi, kTcBMat)) != cudaSuccess || (e = cudaMalloc(&pl->d_tc_wlo, kTcBMat)) != cudaSuccess ||
      (e = cudaMalloc(&pl->d_tc2_w, 4 * kT2BMat)) != cudaSuccess)
    return cuda_fail(ctx, e, "cudaMalloc(tc plan)");
  if ((e = cudaMemcpy(pl->d_tc_whi, hi.data(), kTcBMat, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(pl->d_tc_wlo, lo.data(), kTcBMat, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(pl->d_tc2_w, w2.data(), 4 * kT2BMat, cudaMemcpyHostToDevice)) != cudaSuccess)
    return cuda_fail(ctx, e, "cudaMemcpy(tc plan)");
  pl->tc_serial = g_tc_serial.fetch_add(1);
  pl->tc_built = true;
  pl->tc_ok = true;
  tc_plan_apply_preemphasis(pl);
  return SPB_OK;
}

// Pre-emphasis constants of the static epilogue (feat_tc2.cuh, t2_pre_power): per bin H = 1 - c W^k in .z / .w of the
// first table and (P, Q) = (exp(+i th_k 79.5), P W^(320 k)) as the second table; a new serial makes the next launch
// upload them.  cof = 0 restores the rotation constants the table was built with.
void tc_plan_apply_preemphasis(spb_plan *pl) {
  if (!pl->tc_built || !pl->tc2_static || pl->h_tc_tab_s.size() < 2 * 128 * 4) return;
  const double two_pi = 6.283185307179586476925286766559;
  const int nb = pl->seg_nbins, kmin = pl->seg_kmin, nfft = pl->p.nfft, S = pl->p.frame_stride, L = pl->p.frame_len;
  const double c = pl->pre_cof;
  for (int j = 0; j < nb && j < 128; ++j) {
    const int k = kmin + j;
    float *e0 = &pl->h_tc_tab_s[((size_t)0 * 128 + j) * 4], *e1 = &pl->h_tc_tab_s[((size_t)1 * 128 + j) * 4];
    const double rot = two_pi * (double)(((long long)k * S) % nfft) / (double)nfft;
    if (c == 0.0) {
      e0[2] = (float)cos(rot); e0[3] = (float)-sin(rot);
      e1[0] = 0.f; e1[1] = 0.f; e1[2] = (float)cos(rot); e1[3] = (float)-sin(rot);
      continue;
    }
    const double th = two_pi * (double)k / (double)nfft;
    const double ph = th * ((double)S - 1.0) / 2.0;                       // P = exp(+i th (S - 1) / 2): fold centre of a hop
    const double qh = ph - two_pi * (double)(((long long)k * L) % nfft) / (double)nfft;  // Q = P W^(L k)
    e0[2] = (float)(1.0 - c * cos(th));
    e0[3] = (float)(c * sin(th));
    e1[0] = (float)cos(ph); e1[1] = (float)sin(ph);
    e1[2] = (float)cos(qh); e1[3] = (float)sin(qh);
  }
  pl->tc_serial = g_tc_serial.fetch_add(1);
}

bool features_tc_supported(const FeatArgs &a) {
  if (!(a.plan && a.plan->tc_ok && a.tc_groups && a.num_tc_groups > 0 && (a.num_tc_groups % 4) == 0 &&
        (a.mode == FEAT_LMFE || (a.mode == FEAT_MFE && a.energy == nullptr))))
    return false;
  if (a.pre_cof != 0.0) {
    // pre-emphasis: the static epilogue of generation 2 only (log-mel rows of the speechpy default bank)
    const Options &opt = options();
    return a.plan->tc2_ok && a.plan->tc2_static && opt.tc_version != 1 && (opt.tc_epilogue < 0 || opt.tc_epilogue == 2) &&
           a.mode == FEAT_LMFE && a.tc_hopsum == nullptr && (reinterpret_cast<uintptr_t>(a.out) & 15) == 0;
  }
  return true;
}

static int launch_features_tc1(spb_context *ctx, const FeatArgs &a) {
  const spb_plan *pl = a.plan;
  TcParams p;
  p.samples = a.samples;
  p.groups = a.tc_groups;
  p.num_tiles = a.num_tc_groups / 4;
  p.w_hi = (const __half *)pl->d_tc_whi;
  p.w_lo = (const __half *)pl->d_tc_wlo;
  for (int hf = 0; hf < 2; ++hf) {
    for (int i = 0; i < 4; ++i) p.half[hf].endmask[i] = pl->tc_endmask[hf][i];
    p.half[hf].ch_begin = pl->tc_half[hf][0];
    p.half[hf].ch_end = pl->tc_half[hf][1];
    p.half[hf].seg0 = pl->tc_half[hf][2];
    p.half[hf].out_lo = pl->tc_half[hf][3];
  }
  p.M = a.M;
  p.scale = (float)(1.0 / ((double)a.nfft * 1073741824.0));
  p.ln_scale = (float)log(1.0 / ((double)a.nfft * 1073741824.0));
  p.mfe = (a.mode == FEAT_MFE) ? 1 : 0;
  p.vec4 = (pl->tc_vec4 && (reinterpret_cast<uintptr_t>(a.out) & 15) == 0) ? 1 : 0;
  p.out = a.out;
  p.flag = pl->d_tc_flag;
  p.zeros = reinterpret_cast<const float *>(pl->d_tc_flag) + 64;
  p.hopsum = a.tc_hopsum;
  // c_tc is one table per DEVICE, shared by every context (stream) of the process: a new owner
  // uploads only after the previous owner's last launch (event), under a lock
  TcConstOwner &own = g_tc_owner[ctx->device & 63];
  std::lock_guard<std::mutex> lock(own.mu);
  if (!own.last_use) SPB_CUDA(ctx, cudaEventCreateWithFlags(&own.last_use, cudaEventDisableTiming));
  if (own.plan != pl || own.serial != pl->tc_serial || own.variant != 0) {
    if (own.plan) SPB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, own.last_use, 0));
    // the copy is stream-ordered and reads h_tc_tab (pageable) before returning
    SPB_CUDA(ctx, cudaMemcpyToSymbolAsync(c_tc, pl->h_tc_tab.data(), sizeof(TcConst), 0, cudaMemcpyHostToDevice, ctx->stream));
    own.plan = pl;
    own.serial = pl->tc_serial;
    own.variant = 0;
  }
  int per_sm = 0;
  const bool km5 = (pl->seg_kmin & 15) == 5;  // the speechpy default bank starts at bin 5
  const bool fast = p.vec4 && !p.mfe;
  const bool sums = p.hopsum != nullptr;
  auto kern = sums ? (km5 ? (fast ? features_tc_kernel<5, true, true> : features_tc_kernel<5, false, true>)
                          : features_tc_kernel<-1, false, true>)
                   : (km5 ? (fast ? features_tc_kernel<5, true, false> : features_tc_kernel<5, false, false>)
                          : features_tc_kernel<-1, false, false>);
  int rc = configure_kernel(ctx, (const void *)kern, kTcThreads, kTcSmem, &per_sm);
  if (rc) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_INVALID, "features_tc: kernel does not fit on an SM");
  long long grid = ctx->sm_count;
  if (grid > p.num_tiles) grid = p.num_tiles;
  if (grid < 1) return SPB_OK;
  kern<<<(unsigned)grid, kTcThreads, kTcSmem, ctx->stream>>>(p);
  SPB_LAUNCH_CHECK(ctx, "features_tc");
  SPB_CUDA(ctx, cudaEventRecord(own.last_use, ctx->stream));
  return SPB_OK;
}

// generation 2 (feat_tc2.cuh)
static int launch_features_tc2(spb_context *ctx, const FeatArgs &a) {
  const spb_plan *pl = a.plan;
  const Options &opt = options();
  T2Params p;
  p.samples = a.samples;
  p.groups = a.tc_groups;
  p.num_tiles = a.num_tc_groups / 4;
  p.w = (const __half *)pl->d_tc2_w;
  p.M = a.M;
  p.scale = (float)(1.0 / ((double)a.nfft * 268435456.0));       // 2^28: samples are scaled by 2^14
  p.ln_scale = (float)log(1.0 / ((double)a.nfft * 268435456.0));
  p.mfe = (a.mode == FEAT_MFE) ? 1 : 0;
  p.out = a.out;
  p.flag = pl->d_tc_flag;
  p.zeros = reinterpret_cast<const float *>(pl->d_tc_flag) + 64;
  p.hopsum = a.tc_hopsum;
  p.redo = a.redo_tc;
  p.pre_c = (float)a.pre_cof;
  for (int m = 0; m < 40; ++m) {
    float h = 1.f;
    if (a.pre_cof != 0.0 && m < a.M && m < (int)pl->h_band_lo.size()) {
      double hm = 0.0;
      for (int k : {pl->h_band_lo[m], pl->h_band_lo[m] + std::max(pl->h_band_n[m], 1) - 1}) {
        const double th = 6.283185307179586476925286766559 * (double)k / (double)a.nfft;
        hm = std::max(hm, sqrt(1.0 + a.pre_cof * a.pre_cof - 2.0 * a.pre_cof * cos(th)));  // |1 - c W^k|, monotonic in k
      }
      h = (float)((hm + 0.05) * (hm + 0.05));
    }
    p.pre_h[m] = h;
  }
  const bool km5 = (pl->seg_kmin & 15) == 5;  // the speechpy default bank starts at bin 5
  const bool aligned16 = (reinterpret_cast<uintptr_t>(a.out) & 15) == 0;
  // epilogue flavour: 2 static default bank, 1 runs of four filters, 0 generic
  const bool legal1 = !p.mfe && km5 && pl->tc_vec4 && aligned16;
  const bool legal2 = !p.mfe && pl->tc2_static && aligned16;
  int epi = legal2 ? 2 : (legal1 ? 1 : 0);
  if (opt.tc_epilogue == 0 || (opt.tc_epilogue == 1 && legal1)) epi = opt.tc_epilogue;  // cross-checks
  const int variant = (epi == 2) ? 1 : 0;
  const int (*hm)[4] = variant ? pl->tc_half_s : pl->tc_half;
  const uint32_t (*em)[4] = variant ? pl->tc_endmask_s : pl->tc_endmask;
  for (int hf = 0; hf < 2; ++hf) {
    for (int i = 0; i < 4; ++i) p.half[hf].endmask[i] = em[hf][i];
    p.half[hf].ch_begin = hm[hf][0];
    p.half[hf].ch_end = hm[hf][1];
    p.half[hf].seg0 = hm[hf][2];
    p.half[hf].out_lo = hm[hf][3];
  }
  TcConstOwner &own = g_tc_owner[ctx->device & 63];
  std::lock_guard<std::mutex> lock(own.mu);
  if (!own.last_use) SPB_CUDA(ctx, cudaEventCreateWithFlags(&own.last_use, cudaEventDisableTiming));
  if (own.plan != pl || own.serial != pl->tc_serial || own.variant != variant) {
    if (own.plan) SPB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, own.last_use, 0));
    const std::vector<float> &tab = variant ? pl->h_tc_tab_s : pl->h_tc_tab;
    SPB_CUDA(ctx, cudaMemcpyToSymbolAsync(c_tc, tab.data(), sizeof(TcConst), 0, cudaMemcpyHostToDevice, ctx->stream));
    own.plan = pl;
    own.serial = pl->tc_serial;
    own.variant = variant;
  }
  const bool sums = p.hopsum != nullptr;
  if (p.pre_c != 0.f && (epi != 2 || sums)) return fail(ctx, SPB_ERR_UNSUPPORTED, "features_tc2: pre-emphasis needs the static epilogue");
  void (*kern)(const T2Params) = nullptr;
  if (epi == 2 && p.pre_c != 0.f) kern = features_tc2_kernel<5, 2, false, true>;
  else if (epi == 2) kern = sums ? features_tc2_kernel<5, 2, true> : features_tc2_kernel<5, 2, false>;
  else if (epi == 1) kern = sums ? features_tc2_kernel<5, 1, true> : features_tc2_kernel<5, 1, false>;
  else if (km5) kern = sums ? features_tc2_kernel<5, 0, true> : features_tc2_kernel<5, 0, false>;
  else kern = sums ? features_tc2_kernel<-1, 0, true> : features_tc2_kernel<-1, 0, false>;
  int per_sm = 0;
  const int threads = t2_threads(epi);
  int rc = configure_kernel(ctx, (const void *)kern, threads, kT2Smem, &per_sm);
  if (rc) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_INVALID, "features_tc2: kernel does not fit on an SM");
  long long grid = ctx->sm_count;
  if (grid > p.num_tiles) grid = p.num_tiles;
  if (grid < 1) return SPB_OK;
  kern<<<(unsigned)grid, threads, kT2Smem, ctx->stream>>>(p);
  SPB_LAUNCH_CHECK(ctx, "features_tc2");
  SPB_CUDA(ctx, cudaEventRecord(own.last_use, ctx->stream));
  return SPB_OK;
}

int launch_features_tc(spb_context *ctx, const FeatArgs &a) {
  if (options().tc_version != 1 && a.plan->tc2_ok) return launch_features_tc2(ctx, a);
  return launch_features_tc1(ctx, a);
}

// true when launch_features_tc will append to a.redo_tc (generation 2 only)
bool features_tc_has_guard(const FeatArgs &a) { return options().tc_version != 1 && a.plan && a.plan->tc2_ok; }

// ------------------------------------------------------------------------------------
// mfcc (+ delta cube) on top of the tensor-core log-mel rows: frame energy, DCT-II, deltas.
//   feature.py:202      frame_energies = sum_k P[k] over all one-sided bins, via Parseval:
//                       (NFFT sum x^2 + X0^2 + X_{NFFT/2}^2) / (2 NFFT); X0 = sum x, X_{NFFT/2} = sum (-1)^n x
//   feature.py:147-152  dct(type 2, ortho)[:, :C], c0 <- log(energy) when dc_elimination
//   processing.py:223-236 (as written)  D(x)[c] = (x[min(c+1,C-1)] + 2 x[min(c+2,C-1)]) / 10
// One warp per group of the tensor-core table (32 hops -> 31 frames): lane = hop for the three
// sums over its 160 samples (frame = hops lane, lane + 1), lane = frame for the DCT, rows staged
// in shared memory so the group's output block leaves in coalesced stores.
struct PostParams {
  const float *samples;
  const TcGroup *groups;
  long long num_groups;
  const float *lmfe;  // [rows][M], addressed by global row
  const float *dct;   // [C][M]
  const float4 *hopsum;  // [groups * 32] {sum x^2, sum x, sum (-1)^n x} per hop from the tensor-core kernel, or nullptr
  int M, C, dc_elim, delta, nfft;
  float *out;         // [rows][C] or [rows][C][3]
};
constexpr int kPostWarps = 8;

__global__ void __launch_bounds__(kPostWarps * 32) mfcc_post_kernel(const PostParams p) {
  extern __shared__ __align__(16) float post_smem[];
  const int M = p.M, C = p.C;
  const int pitch = C | 1;
  float *s_dct = post_smem;                                   // [C][M]
  float *s_stage = post_smem + ((C * M + 3) & ~3);            // [warps][31][pitch]
  for (int i = threadIdx.x; i < C * M; i += blockDim.x) s_dct[i] = __ldg(p.dct + i);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *st = s_stage + warp * kTcGroupFrames * pitch;
  const int lpitch = M | 1;
  float *sl = s_stage + kPostWarps * kTcGroupFrames * pitch + warp * kTcGroupFrames * lpitch;  // log-mel rows
  const float nf2 = (float)p.nfft, inv2n = 1.0f / (2.f * (float)p.nfft);
  // element i = lane + 32 k of a row-major [rows][W] block walks (row, col) by a fixed step: no division per element
  const int m_f0 = lane / M, m_c0 = lane - m_f0 * M, m_df = 32 / M, m_dc = 32 - m_df * M;
  const int c_f0 = lane / C, c_c0 = lane - c_f0 * C, c_df = 32 / C, c_dc = 32 - c_df * C;
  for (long long gi = (long long)blockIdx.x * kPostWarps + warp; gi < p.num_groups; gi += (long long)gridDim.x * kPostWarps) {
    const TcGroup g = p.groups[gi];
    if (g.nf <= 0) continue;
    // per-hop sums.  Four lanes share a hop (8 hops per pass: a request touches 8 lines, not 32 --
    // with lane = hop the loads ran at the L1 tag rate), quad reduction, then hop h's sums move to lane h
    float s2 = 0.f, s0 = 0.f, sa = 0.f;
    if (p.hopsum != nullptr) {
      const float4 hs = __ldg(p.hopsum + gi * 32 + lane);
      s2 = hs.x; s0 = hs.y; sa = hs.z;
    } else
#pragma unroll
    for (int ps = 0; ps < 4; ++ps) {
      const int h = ps * 8 + (lane >> 2);
      float t2 = 0.f, t0 = 0.f, ta = 0.f;
      if (h <= g.nf) {
        const float4 *q = reinterpret_cast<const float4 *>(p.samples + g.gstart + (long long)h * kTcHop) + (lane & 3);
#pragma unroll
        for (int i = 0; i < kTcHop / 16; ++i) {
          const float4 x = __ldg(q + 4 * i);
          t2 = fmaf(x.x, x.x, fmaf(x.y, x.y, fmaf(x.z, x.z, fmaf(x.w, x.w, t2))));
          t0 += (x.x + x.y) + (x.z + x.w);
          ta += (x.x - x.y) + (x.z - x.w);
        }
      }
#pragma unroll
      for (int o = 1; o <= 2; o <<= 1) {
        t2 += __shfl_xor_sync(0xffffffffu, t2, o);
        t0 += __shfl_xor_sync(0xffffffffu, t0, o);
        ta += __shfl_xor_sync(0xffffffffu, ta, o);
      }
      const int src = (lane & 7) * 4;  // hop ps*8 + (lane & 7) sits in lanes 4 (lane & 7) ..
      const float u2 = __shfl_sync(0xffffffffu, t2, src);
      const float u0 = __shfl_sync(0xffffffffu, t0, src);
      const float ua = __shfl_sync(0xffffffffu, ta, src);
      if ((lane >> 3) == ps) { s2 = u2; s0 = u0; sa = ua; }
    }
    const float n2 = __shfl_down_sync(0xffffffffu, s2, 1);
    const float n0 = __shfl_down_sync(0xffffffffu, s0, 1);
    const float na = __shfl_down_sync(0xffffffffu, sa, 1);
    // the group's log-mel rows are one contiguous block: coalesced into shared memory, then lane = frame
    {
      const float *lsrc = p.lmfe + g.row0 * M;
      const int total = g.nf * M;
      int f = m_f0, m = m_c0;
      // asynchronous 4-byte copies: all of the lane's ~39 loads are in flight at once (as plain
      // loads each one stalled its store: 36 % of the kernel's samples)
      for (int i = lane; i < total; i += 32) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(sl + f * lpitch + m)), "l"(lsrc + i) : "memory");
        f += m_df;
        m += m_dc;
        if (m >= M) { m -= M; ++f; }
      }
      asm volatile("cp.async.wait_all;" ::: "memory");
    }
    __syncwarp();
    if (lane < g.nf) {
      const float x0 = s0 + n0, xn = sa + na;  // the hop length is even: the second hop keeps the signs
      float e = (nf2 * (s2 + n2) + x0 * x0 + xn * xn) * inv2n;
      if (e == 0.f) e = kTcEps;  // feature.py:205
      const float *lrow = sl + lane * lpitch;
      float lm[kTcOutPitch];
#pragma unroll
      for (int m = 0; m < kTcOutPitch; ++m) lm[m] = (m < M) ? lrow[m] : 0.f;
      for (int c = 0; c < C; ++c) {
        const float *drow = s_dct + c * M;
        float acc = 0.f;
        if ((M & 3) == 0) {  // rows of the table are 16-byte aligned: broadcast 128-bit loads
          const float4 *d4 = reinterpret_cast<const float4 *>(drow);
#pragma unroll
          for (int m4 = 0; m4 < kTcOutPitch / 4; ++m4)
            if (m4 * 4 < M) {
              const float4 w = d4[m4];
              acc = fmaf(w.x, lm[4 * m4], acc);
              acc = fmaf(w.y, lm[4 * m4 + 1], acc);
              acc = fmaf(w.z, lm[4 * m4 + 2], acc);
              acc = fmaf(w.w, lm[4 * m4 + 3], acc);
            }
        } else {
#pragma unroll
          for (int m = 0; m < kTcOutPitch; ++m)
            if (m < M) acc = fmaf(drow[m], lm[m], acc);
        }
        if (c == 0 && p.dc_elim) acc = __logf(e);  // feature.py:151-152
        st[lane * pitch + c] = acc;
      }
    }
    __syncwarp();
    if (p.delta) {
      // lane = (frame, coefficient): the three channels x, D(x), D(D(x)) are 12 consecutive bytes
      float *dst = p.out + g.row0 * C * 3;
      const int total = g.nf * C;
      int f = c_f0, c = c_c0;
      for (int i = lane; i < total; i += 32) {
        const float *x = st + f * pitch;
        auto X = [&](int j) { return x[min(j, C - 1)]; };
        const float x1 = X(c + 1), x2 = X(c + 2), x3 = X(c + 3), x4 = X(c + 4);
        // D1(j) = (X(j+1) + 2 X(j+2)) / 10 with j clamped to C-1 (processing.py:223-236 as written)
        const float d_c = (x1 + 2.f * x2) / 10.f;
        const float d_c1 = (c + 1 <= C - 1) ? (x2 + 2.f * x3) / 10.f : (X(C) + 2.f * X(C + 1)) / 10.f;
        const float d_c2 = (c + 2 <= C - 1) ? (x3 + 2.f * x4) / 10.f : (X(C) + 2.f * X(C + 1)) / 10.f;
        float *o3 = dst + 3 * i;
        o3[0] = x[c];
        o3[1] = d_c;
        o3[2] = (d_c1 + 2.f * d_c2) / 10.f;
        f += c_df;
        c += c_dc;
        if (c >= C) { c -= C; ++f; }
      }
    } else {
      float *dst = p.out + g.row0 * C;
      const int total = g.nf * C;
      int f = c_f0, c = c_c0;
      for (int i = lane; i < total; i += 32) {
        dst[i] = st[f * pitch + c];
        f += c_df;
        c += c_dc;
        if (c >= C) { c -= C; ++f; }
      }
    }
    __syncwarp();
  }
}

int launch_mfcc_post(spb_context *ctx, const FeatArgs &a, const float *d_lmfe) {
  if (a.num_tc_groups <= 0) return SPB_OK;
  PostParams p;
  p.samples = a.samples;
  p.groups = a.tc_groups;
  p.num_groups = a.num_tc_groups;
  p.lmfe = d_lmfe;
  p.dct = a.dct;
  p.hopsum = a.tc_hopsum;
  p.M = a.M;
  p.C = a.C;
  p.dc_elim = a.dc_elim;
  p.delta = (a.mode == FEAT_MFCC_DELTA) ? 1 : 0;
  p.nfft = a.nfft;
  p.out = a.out;
  const size_t smem = (size_t)(((a.C * a.M + 3) & ~3) + kPostWarps * kTcGroupFrames * ((a.C | 1) + (a.M | 1))) * sizeof(float);
  int per_sm = 0;
  int rc = configure_kernel(ctx, (const void *)mfcc_post_kernel, kPostWarps * 32, smem, &per_sm);
  if (rc) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_INVALID, "mfcc_post: kernel does not fit on an SM");
  long long grid = (p.num_groups + kPostWarps - 1) / kPostWarps;
  const long long cap = (long long)ctx->sm_count * per_sm;
  if (grid > cap) grid = cap;
  mfcc_post_kernel<<<(unsigned)grid, kPostWarps * 32, smem, ctx->stream>>>(p);
  SPB_LAUNCH_CHECK(ctx, "mfcc_post");
  return SPB_OK;
}

}  // namespace spb
