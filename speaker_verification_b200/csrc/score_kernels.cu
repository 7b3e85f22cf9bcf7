// K4/K5: scoring path -- cosine score matrix on the tcgen05 tensor cores, trial gather,
// row argmax, ROC -> EER / AUC.
//
// Replaces
//   evaluation.py:76-77   one sklearn cosine_similarity(1xd, 1xd) call per (utterance, speaker)
//   evaluation.py:79      argmax over the similarity vector
//   evaluation.py:47-52   roc_curve + roc_auc_score + brentq(1 - x - interp1d(fpr, tpr)(x))
//
// Cosine matrix.  S = normalize(E) . normalize(T)^T (sklearn semantics: zero rows stay zero).
//   prologue  : one warp per row: 1/|x| in fp32, x^ = x/|x|, split x^ = hi + lo with hi rounded
//               to TF32 (cvt.rna) and lo = tf32(x^ - hi); rows padded with zeros to a multiple
//               of 16 columns.  hi/lo live in context scratch (4 x 10 MB at 4874 x 512).
//   GEMM      : persistent, warp-specialised, one CTA per SM.  Per 128 x BN output tile (BN = 256,
//               224 or 192: whichever needs the fewest waves x width) and 16-column k-block, TMA
//               (SWIZZLE_64B) stages A_hi, A_lo (128 x 16) and B_hi, B_lo (BN x 16) in one of four
//               stages and one elected thread issues three groups of two
//               tcgen05.mma.kind::tf32 (M128 N BN K8): lo.hi + hi.lo + hi.hi, all accumulating
//               in the same fp32 TMEM tile -- the 3xTF32 split the 1e-5 tolerance needs
//               (single-pass TF32 is ~7e-4 relative on 512-d embeddings).  TMEM accumulation
//               truncates, so a chain is cut every 8 k-blocks (48 MMAs): chunks alternate between two
//               TMEM accumulators (2 x 256 columns) and are summed by the epilogue warps.
//   epilogue  : 8 warps, tcgen05.ld 32x32b.x32 of every chunk -> fp32 register accumulators
//               (round-to-nearest adds) -> at the end of a tile padded smem transpose ->
//               row-contiguous 128-byte global stores (the row pitch 4n bytes is in general not
//               a multiple of 16, so a TMA store of S is not possible).
#include <cuda.h>
#include <math.h>

#include <mutex>
#include <string>

#include <cub/cub.cuh>

#include "ptx_helpers.cuh"
#include "spb_internal.cuh"

namespace spb {

// K-major operand tile in shared memory, rows of 128 bytes, SWIZZLE_128B (what TMA wrote):
// start address >> 4, LBO = 1 (unused for swizzled K-major), SBO = 1024 B between 8-row
// groups, descriptor version 1 (sm_100), layout type 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t umma_desc_sw128(const void *smem) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_u32(smem) & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// the same with rows of 64 bytes, SWIZZLE_64B: SBO = 512 B between 8-row groups, layout type 4
__device__ __forceinline__ uint64_t umma_desc_sw64(const void *smem) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_u32(smem) & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)4 << 61;
  return d;
}

// ------------------------------------------------------------------------------------
// prologue: normalise + TF32 hi/lo split, one warp per row
// ------------------------------------------------------------------------------------
__device__ __forceinline__ float tf32_round(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

__global__ void __launch_bounds__(256)
normalize_split_kernel(const float *__restrict__ X, long long rows, int d, int dp,
                       float *__restrict__ hi, float *__restrict__ lo) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float *x = X + row * d;
  float ss = 0.f;
  for (int c = lane; c < d; c += 32) {
    const float v = __ldg(x + c);
    ss = fmaf(v, v, ss);
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  // sklearn normalize(): zero-norm rows are divided by 1 (they stay zero)
  const float nrm = sqrtf(ss);
  const float inv = (nrm == 0.f) ? 1.f : 1.f / nrm;
  for (int c = lane; c < dp; c += 32) {
    const float v = (c < d) ? __ldg(x + c) * inv : 0.f;
    const float h = tf32_round(v);
    hi[row * dp + c] = h;
    lo[row * dp + c] = tf32_round(v - h);
  }
}

// ------------------------------------------------------------------------------------
// cosine GEMM
// ------------------------------------------------------------------------------------
// k-block = 16 floats = one 64-byte swizzle row.  Four stages of 16 columns instead of two of 32: the same
// shared memory, but three stages (up to 144 KB) can be in flight behind the one being multiplied --
// with two stages the MMAs of a k-block (1.7 k cycles) did not cover a TMA round trip and the tensor pipe
// sat at 73 % (ncu r02q).
constexpr int kBM = 128, kBK = 16;
// Tile width BN is a template parameter (256, 224 or 192 columns): the launcher picks the one with the
// fewest (waves x BN) -- 4 874 columns are 20 tiles of 256 = 780 tiles = 5.27 waves on 148 SMs (a sixth
// wave with 40 busy SMs), but 22 tiles of 224 = 858 tiles = 5.8 waves of narrower tiles: 12.5 % less time.
constexpr int kAccStride = 256;                // TMEM columns between the two accumulators
constexpr int kStages = 4;
constexpr int kUmmaK = 8;                      // tf32
constexpr int kGemmThreads = 320;              // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int kMaxChunkKB = 8;                 // at most this many k-blocks (48 MMAs) accumulate in TMEM before a drain
constexpr uint32_t kABytes = kBM * kBK * 4;    // 8 KB
constexpr int kEpiPitch = 33;
constexpr uint32_t kEpiBytes = 8 * 32 * kEpiPitch * 4;
template <int BN>
struct GemmGeom {
  static_assert(BN % 32 == 0 && BN > 128 && BN <= 256, "cosine GEMM tile width");
  static constexpr uint32_t kBBytes = BN * kBK * 4;                    // 32 KB at BN = 256
  static constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytes;   // 96 KB at BN = 256
  static constexpr uint32_t kSmem = 1024 /* alignment slack */ + kStages * kStageBytes + kEpiBytes + 128;
  // instruction descriptor: D = F32 (bits 4-5 = 1), A = B = TF32 (bits 7-9, 10-12 = 2), both
  // K-major, N >> 3 at bit 17, M >> 4 at bit 24
  static constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                     ((uint32_t)(kBM >> 4) << 24);
  static constexpr int kChunks1 = (BN - 128) / 32;  // 32-column chunks of the second column half (the first has 4)
};

struct GemmParams {
  float *S;
  long long m, n;
  int kblocks;
  int chunk_kb;  // k-blocks per TMEM accumulation chain (1..kMaxChunkKB)
  int tiles_m, tiles_n;
};

// TMA load of one box into the same shared-memory offset of every CTA in `mask` of the cluster; each
// destination CTA's mbarrier (same offset) receives the bytes
__device__ __forceinline__ void tma_load_2d_mc(const CUtensorMap *map, uint64_t *bar, void *dst, int c0, int c1,
                                               uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}
// tcgen05.commit that arrives on the same-offset mbarrier of every CTA in `mask`
__device__ __forceinline__ void tc_commit_mc(uint64_t *bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// MC: clusters of two CTAs work on two row tiles (tm, tm + 1) of the same column tile tn.  Each CTA loads its
// own A tiles and HALF of the B tiles, multicast into both CTAs' stages: the L2 -> SM operand traffic per CTA
// drops from 128 + BN to 128 + BN / 2 rows per k-block.  Measured on B200 at 4874 x 4874 x 512, both in one
// process: 108.8-109.0 us against 113.8-115.6 us for the plain kernel (spb_set_option("gemm_multicast", 0));
// results are bit-identical.  A stage may be refilled only when BOTH CTAs' MMAs have read it: the MMA
// warps' commits arrive on both CTAs' `empty` barriers (count 2).
template <int BN, bool MC>
__global__ void __launch_bounds__(kGemmThreads, 1)
cosine_gemm_kernel(const __grid_constant__ CUtensorMap map_ehi, const __grid_constant__ CUtensorMap map_elo,
                   const __grid_constant__ CUtensorMap map_thi, const __grid_constant__ CUtensorMap map_tlo,
                   const GemmParams p) {
  using Gm = GemmGeom<BN>;
  constexpr uint32_t kBBytes = Gm::kBBytes, kStageBytes = Gm::kStageBytes, kIdesc = Gm::kIdesc;
  constexpr int kBN = BN;
  const uint32_t crank = MC ? cluster_ctarank() : 0u;
  // work units: MC: (pair of row tiles, column tile), dealt to clusters; else tiles dealt to CTAs
  const int unit_m = MC ? (p.tiles_m + 1) / 2 : p.tiles_m;
  const long long num_tiles = (long long)unit_m * p.tiles_n;
  const long long unit0 = MC ? (long long)(blockIdx.x >> 1) : (long long)blockIdx.x;
  const long long unit_step = MC ? (long long)(gridDim.x >> 1) : (long long)gridDim.x;
  extern __shared__ unsigned char smem_raw[];
  unsigned char *base = reinterpret_cast<unsigned char *>(
      ((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);  // swizzled stages: 1024-byte alignment
  float *s_epi = reinterpret_cast<float *>(base + kStages * kStageBytes);
  uint64_t *bars = reinterpret_cast<uint64_t *>(base + kStages * kStageBytes + kEpiBytes);
  uint64_t *full = bars;                 // [kStages]
  uint64_t *empty = bars + kStages;      // [kStages]
  uint64_t *acc_full = bars + 2 * kStages;       // [2]
  uint64_t *acc_empty = bars + 2 * kStages + 2;  // [2]
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_ehi) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_elo) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_thi) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_tlo) : "memory");
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], MC ? 2 : 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], 8);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    // 512 TMEM columns: two 128 x 256 fp32 accumulators (one CTA per SM, so no contention)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (MC) cluster_sync_all();  // the peer's barriers are initialised before anything arrives on them

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (long long t = unit0; t < num_tiles; t += unit_step) {
        const int tm = MC ? 2 * (int)(t % unit_m) + (int)crank : (int)(t % unit_m), tn = (int)(t / unit_m);
        for (int kb = 0; kb < p.kblocks; ++kb) {
          mbar_wait_guarded(&empty[stage], phase ^ 1);
          unsigned char *st = base + stage * kStageBytes;
          mbar_expect_tx(&full[stage], kStageBytes);
          tma_load_2d(&map_ehi, &full[stage], st, kb * kBK, tm * kBM);
          tma_load_2d(&map_elo, &full[stage], st + kABytes, kb * kBK, tm * kBM);
          if (MC) {
            // this CTA's half of the B rows, into both CTAs
            const uint32_t hb = crank * (kBBytes / 2);
            const int r0 = tn * kBN + (int)crank * (kBN / 2);
            tma_load_2d_mc(&map_thi, &full[stage], st + 2 * kABytes + hb, kb * kBK, r0, (uint16_t)3);
            tma_load_2d_mc(&map_tlo, &full[stage], st + 2 * kABytes + kBBytes + hb, kb * kBK, r0, (uint16_t)3);
          } else {
            tma_load_2d(&map_thi, &full[stage], st + 2 * kABytes, kb * kBK, tn * kBN);
            tma_load_2d(&map_tlo, &full[stage], st + 2 * kABytes + kBBytes, kb * kBK, tn * kBN);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one elected thread =====
    // The tensor core adds into the fp32 TMEM accumulator with truncation (measured: a 192-MMA
    // chain loses ~1e-6 absolute, always towards zero), so a chain is cut after chunk_kb
    // k-blocks: each chunk starts a fresh accumulator (alternating between the two TMEM
    // buffers) and the epilogue warps sum the chunks in registers with round-to-nearest adds.
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      long long cit = 0;  // chunk counter across tiles
      for (long long t = unit0; t < num_tiles; t += unit_step) {
        for (int kb0 = 0; kb0 < p.kblocks; kb0 += p.chunk_kb, ++cit) {
          const uint32_t acc = (uint32_t)(cit & 1);
          const uint32_t acc_phase = (uint32_t)((cit >> 1) & 1);
          mbar_wait_guarded(&acc_empty[acc], acc_phase ^ 1);  // epilogue has drained this accumulator
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + acc * kAccStride;
          const int kb1 = min(kb0 + p.chunk_kb, p.kblocks);
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait_guarded(&full[stage], phase);
            tc_fence_after();
            unsigned char *st = base + stage * kStageBytes;
            const uint64_t a_hi = umma_desc_sw64(st);
            const uint64_t a_lo = umma_desc_sw64(st + kABytes);
            const uint64_t b_hi = umma_desc_sw64(st + 2 * kABytes);
            const uint64_t b_lo = umma_desc_sw64(st + 2 * kABytes + kBBytes);
            // small terms first: lo.hi, hi.lo, then hi.hi
#pragma unroll
            for (int k = 0; k < kBK / kUmmaK; ++k) {
              const uint64_t adv = (uint64_t)((k * kUmmaK * 4) >> 4);  // 32 bytes along K per step
              umma_tf32(tmem_d, a_lo + adv, b_hi + adv, kIdesc, (kb != kb0 || k != 0) ? 1u : 0u);
            }
#pragma unroll
            for (int k = 0; k < kBK / kUmmaK; ++k) {
              const uint64_t adv = (uint64_t)((k * kUmmaK * 4) >> 4);
              umma_tf32(tmem_d, a_hi + adv, b_lo + adv, kIdesc, 1u);
            }
#pragma unroll
            for (int k = 0; k < kBK / kUmmaK; ++k) {
              const uint64_t adv = (uint64_t)((k * kUmmaK * 4) >> 4);
              umma_tf32(tmem_d, a_hi + adv, b_hi + adv, kIdesc, 1u);
            }
            // frees the smem stage when these MMAs retire (MC: in both CTAs, each of which writes into it)
            if (MC) tc_commit_mc(&empty[stage], (uint16_t)3);
            else tc_commit(&empty[stage]);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
          tc_commit(&acc_full[acc]);   // chunk complete -> epilogue
        }
      }
    }
  } else {
    // ===== epilogue: warps 2..9.  TMEM lane quadrant = warp % 4, column half = (warp-2)/4 =====
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    float *my = s_epi + (warp - 2) * 32 * kEpiPitch;
    const int nch = (half == 0) ? 4 : Gm::kChunks1;
    long long cit = 0;
    for (long long t = unit0; t < num_tiles; t += unit_step) {
      const int tm = MC ? 2 * (int)(t % unit_m) + (int)crank : (int)(t % unit_m), tn = (int)(t / unit_m);
      float accr[128];
#pragma unroll
      for (int j = 0; j < 128; ++j) accr[j] = 0.f;
      for (int kb0 = 0; kb0 < p.kblocks; kb0 += p.chunk_kb, ++cit) {
        const uint32_t acc = (uint32_t)(cit & 1);
        const uint32_t acc_phase = (uint32_t)((cit >> 1) & 1);
        mbar_wait_guarded(&acc_full[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * kAccStride + half * 128;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          if (ch < nch) {  // warp-uniform: the second column half is narrower when BN < 256
            uint32_t v[32];
            tmem_ld32(taddr + ch * 32, v);
#pragma unroll
            for (int j = 0; j < 32; ++j) accr[ch * 32 + j] += __uint_as_float(v[j]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_empty[acc]);
      }
      // write-out: lane = row of the quadrant; transpose through smem so stores run along columns
      const long long row_base = (long long)tm * kBM + quad * 32;
      const long long col_base = (long long)tn * kBN + half * 128;
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        if (ch >= nch) break;
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 32; ++j) my[lane * kEpiPitch + j] = accr[ch * 32 + j];
        __syncwarp();
        const long long col = col_base + ch * 32 + lane;
        if (col < p.n) {
#pragma unroll 8
          for (int r = 0; r < 32; ++r) {
            const long long row = row_base + r;
            if (row < p.m) p.S[row * p.n + col] = my[r * kEpiPitch + lane];
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (MC) cluster_sync_all();  // no CTA retires while its peer may still write into its stages / barriers
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ------------------------------------------------------------------------------------
// host side of the cosine GEMM
// ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn(spb_context *ctx) {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void *p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    cudaGetLastError();
    fail(ctx, SPB_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    return nullptr;
  }
  fn = (EncodeTiledFn)p;
  return fn;
}

// (rows, dp) row-major fp32, box = (kBK columns, box_rows), SWIZZLE_64B, OOB rows read as 0
static int make_operand_map(spb_context *ctx, CUtensorMap *map, const float *ptr, long long rows, int dp,
                            int box_rows) {
  EncodeTiledFn enc = get_encode_fn(ctx);
  if (!enc) return SPB_ERR_CUDA;
  cuuint64_t dims[2] = {(cuuint64_t)dp, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)dp * 4};
  cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)ptr, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ctx, SPB_ERR_CUDA, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  return SPB_OK;
}

// floats of hi/lo scratch a (m, d) x (n, d) problem needs: [E_hi | E_lo | T_hi | T_lo], each block 1 KB aligned
static inline size_t split_align(size_t x) { return (x + 255) & ~(size_t)255; }
static size_t cosine_scratch_floats(int64_t m, int64_t n, int d) {
  const int dp = (d + kBK - 1) / kBK * kBK;
  return 2 * split_align((size_t)m * dp) + 2 * split_align((size_t)n * dp);
}

// prologue + GEMM with the hi/lo split in `buf` (cosine_scratch_floats(m, n, d) floats).  timed: the
// GEMM launch is bracketed by the context's timing events when timing is on (those are created on demand).
static int cosine_scores_in(spb_context *ctx, const float *d_E, int64_t m, const float *d_T, int64_t n, int d,
                            float *d_S, float *buf, bool timed) {
  const int dp = (d + kBK - 1) / kBK * kBK;
  const size_t e_elems = (size_t)m * dp, t_elems = (size_t)n * dp;
  auto al = split_align;
  float *e_hi = buf, *e_lo = e_hi + al(e_elems), *t_hi = e_lo + al(e_elems), *t_lo = t_hi + al(t_elems);
  const int rows_per_block = 256 / 32;
  normalize_split_kernel<<<(unsigned)((m + rows_per_block - 1) / rows_per_block), 256, 0, ctx->stream>>>(
      d_E, m, d, dp, e_hi, e_lo);
  SPB_LAUNCH_CHECK(ctx, "normalize_split_kernel(E)");
  normalize_split_kernel<<<(unsigned)((n + rows_per_block - 1) / rows_per_block), 256, 0, ctx->stream>>>(
      d_T, n, d, dp, t_hi, t_lo);
  SPB_LAUNCH_CHECK(ctx, "normalize_split_kernel(T)");
  const int tiles_m = (int)((m + kBM - 1) / kBM);
  typedef void (*GemmKern)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const GemmParams);
  struct Variant { int bn; GemmKern plain, mc; size_t smem; };
  const Variant variants[3] = {
      {256, cosine_gemm_kernel<256, false>, cosine_gemm_kernel<256, true>, GemmGeom<256>::kSmem},
      {224, cosine_gemm_kernel<224, false>, cosine_gemm_kernel<224, true>, GemmGeom<224>::kSmem},
      {192, cosine_gemm_kernel<192, false>, cosine_gemm_kernel<192, true>, GemmGeom<192>::kSmem}};
  // 2-CTA clusters with the T tiles multicast when there are at least two row tiles and the device takes
  // the clusters; how many fit is asked once per kernel
  bool mc = tiles_m >= 2 && options().gemm_multicast != 0;
  int clusters = 0;
  if (mc) {
    static int cached_clusters[3] = {-1, -1, -1};
    for (int v = 0; v < 3 && mc; ++v) {
      if (cached_clusters[v] < 0) {
        int per_sm = 0;
        int rc0 = configure_kernel(ctx, (const void *)variants[v].mc, kGemmThreads, variants[v].smem, &per_sm);
        if (rc0) return rc0;
        cudaLaunchConfig_t qc = {};
        qc.gridDim = dim3((unsigned)(ctx->sm_count & ~1), 1, 1);
        qc.blockDim = dim3(kGemmThreads, 1, 1);
        qc.dynamicSmemBytes = variants[v].smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        qc.attrs = at;
        qc.numAttrs = 1;
        int n_cl = 0;
        if (cudaOccupancyMaxActiveClusters(&n_cl, (const void *)variants[v].mc, &qc) != cudaSuccess) { cudaGetLastError(); n_cl = 0; }
        cached_clusters[v] = n_cl;
      }
      mc = cached_clusters[v] >= 8;
    }
    if (mc) clusters = std::min(std::min(cached_clusters[0], cached_clusters[1]), cached_clusters[2]);
  }
  // tile width: fewest waves x width over the persistent grid (ties go to the wider tile)
  int vi = 0;
  {
    long long best = -1;
    for (int v = 0; v < 3; ++v) {
      const int cand = variants[v].bn;
      const long long units = (long long)(mc ? (tiles_m + 1) / 2 : tiles_m) * ((n + cand - 1) / cand);
      const long long workers = mc ? clusters : ctx->sm_count;
      const long long cost = ((units + workers - 1) / workers) * cand;
      if (best < 0 || cost < best) { best = cost; vi = v; }
    }
  }
  const int bn = variants[vi].bn;
  CUtensorMap m_ehi, m_elo, m_thi, m_tlo;
  int rc;
  const int b_box = mc ? bn / 2 : bn;
  if ((rc = make_operand_map(ctx, &m_ehi, e_hi, m, dp, kBM)) || (rc = make_operand_map(ctx, &m_elo, e_lo, m, dp, kBM)) ||
      (rc = make_operand_map(ctx, &m_thi, t_hi, n, dp, b_box)) || (rc = make_operand_map(ctx, &m_tlo, t_lo, n, dp, b_box)))
    return rc;
  GemmParams p;
  p.S = d_S;
  p.m = m;
  p.n = n;
  p.kblocks = dp / kBK;
  // short chains for small d (precision), up to 8 k-blocks (48 MMAs) for large d (drain bandwidth)
  p.chunk_kb = std::max(1, std::min(kMaxChunkKB, p.kblocks / 4));
  p.tiles_m = tiles_m;
  p.tiles_n = (int)((n + bn - 1) / bn);
  const GemmKern kern = mc ? variants[vi].mc : variants[vi].plain;
  const size_t smem = variants[vi].smem;
  int per_sm = 0;
  if ((rc = configure_kernel(ctx, (const void *)kern, kGemmThreads, smem, &per_sm))) return rc;
  if (per_sm < 1) return fail(ctx, SPB_ERR_INVALID, "cosine_gemm_kernel does not fit on an SM");
  const size_t tb = timed ? timing_begin(ctx) : (size_t)-1;
  if (mc) {
    const long long units = (long long)((tiles_m + 1) / 2) * p.tiles_n;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)(2 * std::min<long long>(units, clusters)), 1, 1);
    lc.blockDim = dim3(kGemmThreads, 1, 1);
    lc.dynamicSmemBytes = smem;
    lc.stream = ctx->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    lc.attrs = at;
    lc.numAttrs = 1;
    SPB_CUDA(ctx, cudaLaunchKernelEx(&lc, kern, m_ehi, m_elo, m_thi, m_tlo, p));
  } else {
    const long long grid = std::min<long long>((long long)p.tiles_m * p.tiles_n, ctx->sm_count);
    kern<<<(unsigned)grid, kGemmThreads, smem, ctx->stream>>>(m_ehi, m_elo, m_thi, m_tlo, p);
  }
  if (timed) timing_end(ctx, 3, tb);
  SPB_LAUNCH_CHECK(ctx, "cosine_gemm_kernel");
  return SPB_OK;
}

int launch_cosine_scores(spb_context *ctx, const float *d_E, int64_t m, const float *d_T, int64_t n, int d,
                         float *d_S) {
  if (m == 0 || n == 0) return SPB_OK;
  // slot 6 of the context's grow-only scratch
  float *buf = (float *)scratch(ctx, 6, cosine_scratch_floats(m, n, d) * 4);
  if (!buf) return SPB_ERR_CUDA;
  return cosine_scores_in(ctx, d_E, m, d_T, n, d, d_S, buf, true);
}

// ------------------------------------------------------------------------------------
// Allocation-free variant (spb_cosine_scores_dev_noalloc).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops, harnesses that forbid it).  The hi/lo split lives in a file-scope __device__
// array of kNoAllocFloats floats (8 MiB) that every GPU which loads the library holds until exit.
// In host code `g_noalloc_split` is only a stand-in for the symbol: the device address is resolved
// once per device with cudaGetSymbolAddress (that device current) and cached.
// ------------------------------------------------------------------------------------
constexpr size_t kNoAllocFloats = (size_t)2 << 20;  // 2 Mi floats = 8 MiB
static_assert(kNoAllocFloats * sizeof(float) == SPB_NOALLOC_SCRATCH_BYTES, "header and library disagree on the limit");
static __device__ __align__(256) float g_noalloc_split[kNoAllocFloats];

static std::mutex g_noalloc_mu;
static float *g_noalloc_ptr[64] = {};

static int noalloc_scratch(spb_context *ctx, float **out) {
  if (ctx->device < 0 || ctx->device >= 64) return fail(ctx, SPB_ERR_INVALID, "noalloc: device index out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  float *ptr = g_noalloc_ptr[ctx->device];
  if (ptr == nullptr) {
    // set-up, once per device, with that device current (the caller made it so)
    void *sym = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&sym, g_noalloc_split);
    if (e != cudaSuccess) return cuda_fail(ctx, e, "cudaGetSymbolAddress(g_noalloc_split)");
    if (sym == nullptr) return fail(ctx, SPB_ERR_CUDA, "noalloc: the static scratch has no device address");
    ptr = g_noalloc_ptr[ctx->device] = static_cast<float *>(sym);
  }
  *out = ptr;
  return SPB_OK;
}

// ------------------------------------------------------------------------------------
// trial gather and row argmax
// ------------------------------------------------------------------------------------
__global__ void trial_gather_kernel(const float *__restrict__ S, long long m, long long n,
                                    const int32_t *__restrict__ ti, const int32_t *__restrict__ tj,
                                    long long nt, float *__restrict__ out, int *__restrict__ bad) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nt) return;
  const long long i = ti[t], j = tj[t];
  if (i < 0 || i >= m || j < 0 || j >= n) {
    *bad = 1;
    out[t] = 0.f;
    return;
  }
  out[t] = __ldg(S + i * n + j);
}

int launch_trial_gather(spb_context *ctx, const float *d_S, int64_t m, int64_t n, const int32_t *ti,
                        const int32_t *tj, int64_t nt, float *out) {
  if (nt <= 0) return SPB_OK;
  int *d_bad = (int *)scratch(ctx, 7, 64);
  if (!d_bad) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemsetAsync(d_bad, 0, 4, ctx->stream));
  trial_gather_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, ctx->stream>>>(d_S, m, n, ti, tj, nt, out, d_bad);
  SPB_LAUNCH_CHECK(ctx, "trial_gather_kernel");
  int bad = 0;
  SPB_CUDA(ctx, cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (bad) return fail(ctx, SPB_ERR_INVALID, "trial index out of range");
  return SPB_OK;
}

// first index of the row maximum (np.argmax, evaluation.py:79); one warp per row
__global__ void __launch_bounds__(256)
argmax_rows_kernel(const float *__restrict__ S, long long m, long long n, int32_t *__restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= m) return;
  const float *s = S + row * n;
  float best = -INFINITY;
  long long bi = n;  // larger than any valid index
  for (long long c = lane; c < n; c += 32) {
    const float v = __ldg(s + c);
    if (v > best || (bi == n)) { best = v; bi = c; }
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (oi < n && (bi == n || ov > best || (ov == best && oi < bi))) { best = ov; bi = oi; }
  }
  if (lane == 0) out[row] = (int32_t)bi;
}

int launch_argmax_rows(spb_context *ctx, const float *d_S, int64_t m, int64_t n, int32_t *out) {
  if (m <= 0) return SPB_OK;
  if (n <= 0) return fail(ctx, SPB_ERR_INVALID, "attempt to get argmax of an empty sequence");
  argmax_rows_kernel<<<(unsigned)((m + 7) / 8), 256, 0, ctx->stream>>>(d_S, m, n, out);
  SPB_LAUNCH_CHECK(ctx, "argmax_rows_kernel");
  return SPB_OK;
}

// ------------------------------------------------------------------------------------
// EER / AUC (evaluation.py:47-52)
//   sort (score, label) by descending score; ROC vertices sit at the last element of every
//   run of equal scores (sklearn _binary_clf_curve); tp = inclusive sum of labels,
//   fp = index + 1 - tp.  drop_intermediate only removes collinear vertices, so the
//   piecewise-linear curve -- and with it brentq's root and the trapezoid area -- is the same
//   without it.  All counting is integer; g = 1 - fpr - tpr is compared through
//   P*N - fp*P - tp*N (exact in int64).  The vertex before the first g <= 0 vertex comes
//   from a max-scan of (group start index, tp before the group) packed in 64 bits.
// ------------------------------------------------------------------------------------
struct MaxU64 {
  __device__ __forceinline__ unsigned long long operator()(unsigned long long a, unsigned long long b) const {
    return a > b ? a : b;
  }
};

__global__ void group_start_kernel(const float *__restrict__ s, const int32_t *__restrict__ lab,
                                   const int32_t *__restrict__ tp, long long n,
                                   unsigned long long *__restrict__ marks) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const bool start = (i == 0) || (s[i - 1] != s[i]);
  const unsigned long long tp_before = (unsigned long long)(tp[i] - (lab[i] != 0 ? 1 : 0));
  marks[i] = start ? (((unsigned long long)i << 32) | tp_before) : 0ull;
}

struct EerAccum {
  unsigned long long first_v;   // smallest group-end index with g <= 0
  unsigned long long auc2;      // sum (fp_i - fp_prev) * (tp_i + tp_prev)  == 2 * P * N * AUC
};

__global__ void __launch_bounds__(256)
roc_vertex_kernel(const float *__restrict__ s, const int32_t *__restrict__ tp,
                  const unsigned long long *__restrict__ gs, long long n, long long P, long long N,
                  EerAccum *__restrict__ acc) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long area = 0;
  if (i < n) {
    const bool end = (i == n - 1) || (s[i] != s[i + 1]);
    if (end) {
      const long long tpi = tp[i], fpi = i + 1 - tpi;
      const long long start = (long long)(gs[i] >> 32), tpp = (long long)(gs[i] & 0xffffffffull);
      const long long fpp = start - tpp;
      area = (unsigned long long)(fpi - fpp) * (unsigned long long)(tpi + tpp);
      const long long g = P * N - fpi * P - tpi * N;
      if (g <= 0) atomicMin(&acc->first_v, (unsigned long long)i);
    }
  }
  typedef cub::BlockReduce<unsigned long long, 256> Reduce;
  __shared__ typename Reduce::TempStorage tmp;
  const unsigned long long tot = Reduce(tmp).Sum(area);
  if (threadIdx.x == 0 && tot) atomicAdd(&acc->auc2, tot);
}

__global__ void eer_solve_kernel(const int32_t *__restrict__ tp, const unsigned long long *__restrict__ gs,
                                 long long P, long long N, const EerAccum *__restrict__ acc,
                                 double *__restrict__ out) {
  const long long v = (long long)acc->first_v;
  const long long tpv = tp[v], fpv = v + 1 - tpv;
  const long long start = (long long)(gs[v] >> 32), tpu = (long long)(gs[v] & 0xffffffffull);
  const long long fpu = start - tpu;
  const double fpr_v = (double)fpv / (double)N, tpr_v = (double)tpv / (double)P;
  const double fpr_u = (double)fpu / (double)N, tpr_u = (double)tpu / (double)P;
  double eer;
  if (P * N - fpv * P - tpv * N == 0) {
    eer = fpr_v;
  } else {
    const double dx = fpr_v - fpr_u, dy = tpr_v - tpr_u;
    eer = (fpv == fpu) ? fpr_u : fpr_u + (1.0 - fpr_u - tpr_u) * dx / (dx + dy);
  }
  out[0] = eer;
  out[1] = (double)acc->auc2 / (2.0 * (double)P * (double)N);
}

int run_eer(spb_context *ctx, const float *d_scores, const int32_t *d_labels, int64_t n, double *eer,
            double *auc) {
  if (n <= 0) return fail(ctx, SPB_ERR_INVALID, "eer: no trials");
  if (n >= ((int64_t)1 << 31) - 1) return fail(ctx, SPB_ERR_INVALID, "eer: too many trials");
  const int in = (int)n;
  size_t sort_bytes = 0, sum_bytes = 0, max_bytes = 0;
  cub::DeviceRadixSort::SortPairsDescending(nullptr, sort_bytes, (const float *)nullptr, (float *)nullptr,
                                            (const int32_t *)nullptr, (int32_t *)nullptr, in, 0, 32, ctx->stream);
  cub::DeviceScan::InclusiveSum(nullptr, sum_bytes, (const int32_t *)nullptr, (int32_t *)nullptr, in, ctx->stream);
  cub::DeviceScan::InclusiveScan(nullptr, max_bytes, (const unsigned long long *)nullptr,
                                 (unsigned long long *)nullptr, MaxU64(), in, ctx->stream);
  const size_t tmp_bytes = std::max(sort_bytes, std::max(sum_bytes, max_bytes));
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t off_s = 0, off_l = off_s + al((size_t)n * 4), off_tp = off_l + al((size_t)n * 4),
               off_marks = off_tp + al((size_t)n * 4), off_gs = off_marks + al((size_t)n * 8),
               off_acc = off_gs + al((size_t)n * 8), off_out = off_acc + 256, off_tmp = off_out + 256;
  unsigned char *buf = (unsigned char *)scratch(ctx, 7, off_tmp + al(tmp_bytes));
  if (!buf) return SPB_ERR_CUDA;
  float *s = (float *)(buf + off_s);
  int32_t *lab = (int32_t *)(buf + off_l), *tp = (int32_t *)(buf + off_tp);
  unsigned long long *marks = (unsigned long long *)(buf + off_marks), *gs = (unsigned long long *)(buf + off_gs);
  EerAccum *acc = (EerAccum *)(buf + off_acc);
  double *d_out = (double *)(buf + off_out);
  void *tmp = buf + off_tmp;
  size_t tb = tmp_bytes;
  SPB_CUDA(ctx, cub::DeviceRadixSort::SortPairsDescending(tmp, tb, d_scores, s, d_labels, lab, in, 0, 32, ctx->stream));
  ctx->launches += 1;
  tb = tmp_bytes;
  SPB_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp, tb, lab, tp, in, ctx->stream));
  ctx->launches += 1;
  int32_t P32 = 0;
  SPB_CUDA(ctx, cudaMemcpyAsync(&P32, tp + (n - 1), 4, cudaMemcpyDeviceToHost, ctx->stream));
  const unsigned blocks = (unsigned)((n + 255) / 256);
  group_start_kernel<<<blocks, 256, 0, ctx->stream>>>(s, lab, tp, n, marks);
  SPB_LAUNCH_CHECK(ctx, "group_start_kernel");
  tb = tmp_bytes;
  SPB_CUDA(ctx, cub::DeviceScan::InclusiveScan(tmp, tb, marks, gs, MaxU64(), in, ctx->stream));
  ctx->launches += 1;
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const long long P = P32, N = n - P32;
  // sklearn: roc_auc_score raises ValueError when only one class is present
  if (P == 0 || N == 0) return fail(ctx, SPB_ERR_INVALID, "Only one class present in y_true. ROC AUC score is not defined in that case.");
  const EerAccum init = {~0ull, 0ull};
  SPB_CUDA(ctx, cudaMemcpyAsync(acc, &init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
  roc_vertex_kernel<<<blocks, 256, 0, ctx->stream>>>(s, tp, gs, n, P, N, acc);
  SPB_LAUNCH_CHECK(ctx, "roc_vertex_kernel");
  eer_solve_kernel<<<1, 1, 0, ctx->stream>>>(tp, gs, P, N, acc, d_out);
  SPB_LAUNCH_CHECK(ctx, "eer_solve_kernel");
  double h_out[2] = {0, 0};
  SPB_CUDA(ctx, cudaMemcpyAsync(h_out, d_out, 16, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (eer) *eer = h_out[0];
  if (auc) *auc = h_out[1];
  return SPB_OK;
}

// ------------------------------------------------------------------ ROC curve export
// sklearn.metrics.roc_curve(label, score, pos_label=1) as evaluation.py:48 calls it
// (drop_intermediate=True): vertices at the last element of every run of equal scores in
// descending order; a vertex is dropped when the second differences of BOTH count sequences
// are zero (collinear), first and last are kept; then (fpr, tpr, threshold) = (0, 0, +inf) is
// prepended.  Counts are integers, so the kept set is exact; fpr = fp / N, tpr = tp / P in
// float64 like sklearn's float division.
__global__ void roc_end_flags_kernel(const float *__restrict__ s, long long n, unsigned char *__restrict__ flags) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flags[i] = (i == n - 1) || (s[i] != s[i + 1]);
}

__global__ void roc_keep_kernel(const int32_t *__restrict__ vidx, const int32_t *__restrict__ tp, int V,
                                unsigned char *__restrict__ keep) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= V) return;
  if (v == 0 || v == V - 1) {
    keep[v] = 1;
    return;
  }
  const long long i0 = vidx[v - 1], i1 = vidx[v], i2 = vidx[v + 1];
  const long long t0 = tp[i0], t1 = tp[i1], t2 = tp[i2];
  const long long f0 = i0 + 1 - t0, f1 = i1 + 1 - t1, f2 = i2 + 1 - t2;
  keep[v] = ((f2 - 2 * f1 + f0) != 0) || ((t2 - 2 * t1 + t0) != 0);
}

__global__ void roc_emit_kernel(const int32_t *__restrict__ kidx, const float *__restrict__ s,
                                const int32_t *__restrict__ tp, int Kc, double P, double N,
                                double *__restrict__ fpr, double *__restrict__ tpr, double *__restrict__ thr) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > Kc) return;
  if (k == 0) {
    fpr[0] = 0.0;
    tpr[0] = 0.0;
    thr[0] = __longlong_as_double(0x7ff0000000000000ll);  // +inf (sklearn >= 1.3)
    return;
  }
  const long long i = kidx[k - 1];
  const long long t = tp[i], f = i + 1 - t;
  fpr[k] = (double)f / N;
  tpr[k] = (double)t / P;
  thr[k] = (double)s[i];
}

int run_roc_curve(spb_context *ctx, const float *d_scores, const int32_t *d_labels, int64_t n,
                  double *h_fpr, double *h_tpr, double *h_thr, int64_t capacity, int64_t *count) {
  if (n <= 0) return fail(ctx, SPB_ERR_INVALID, "roc_curve: no trials");
  if (n >= ((int64_t)1 << 31) - 1) return fail(ctx, SPB_ERR_INVALID, "roc_curve: too many trials");
  const int in = (int)n;
  size_t sort_bytes = 0, sum_bytes = 0, sel_bytes = 0;
  cub::DeviceRadixSort::SortPairsDescending(nullptr, sort_bytes, (const float *)nullptr, (float *)nullptr,
                                            (const int32_t *)nullptr, (int32_t *)nullptr, in, 0, 32, ctx->stream);
  cub::DeviceScan::InclusiveSum(nullptr, sum_bytes, (const int32_t *)nullptr, (int32_t *)nullptr, in, ctx->stream);
  cub::DeviceSelect::Flagged(nullptr, sel_bytes, cub::CountingInputIterator<int32_t>(0), (const unsigned char *)nullptr,
                             (int32_t *)nullptr, (int32_t *)nullptr, in, ctx->stream);
  const size_t tmp_bytes = std::max(sort_bytes, std::max(sum_bytes, sel_bytes));
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t n4 = al((size_t)n * 4), n8 = al(((size_t)n + 1) * 8);
  const size_t off_s = 0, off_l = off_s + n4, off_tp = off_l + n4, off_vidx = off_tp + n4, off_kidx = off_vidx + n4,
               off_flags = off_kidx + n4, off_cnt = off_flags + al((size_t)n), off_fpr = off_cnt + 256,
               off_tpr = off_fpr + n8, off_thr = off_tpr + n8, off_tmp = off_thr + n8;
  unsigned char *buf = (unsigned char *)scratch(ctx, 7, off_tmp + al(tmp_bytes));
  if (!buf) return SPB_ERR_CUDA;
  float *s = (float *)(buf + off_s);
  int32_t *lab = (int32_t *)(buf + off_l), *tp = (int32_t *)(buf + off_tp);
  int32_t *vidx = (int32_t *)(buf + off_vidx), *kidx = (int32_t *)(buf + off_kidx);
  unsigned char *flags = buf + off_flags;
  int32_t *d_cnt = (int32_t *)(buf + off_cnt);
  double *fpr = (double *)(buf + off_fpr), *tpr = (double *)(buf + off_tpr), *thr = (double *)(buf + off_thr);
  void *tmp = buf + off_tmp;
  size_t tb = tmp_bytes;
  SPB_CUDA(ctx, cub::DeviceRadixSort::SortPairsDescending(tmp, tb, d_scores, s, d_labels, lab, in, 0, 32, ctx->stream));
  tb = tmp_bytes;
  SPB_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp, tb, lab, tp, in, ctx->stream));
  ctx->launches += 2;
  const unsigned blocks = (unsigned)((n + 255) / 256);
  roc_end_flags_kernel<<<blocks, 256, 0, ctx->stream>>>(s, n, flags);
  SPB_LAUNCH_CHECK(ctx, "roc_end_flags_kernel");
  tb = tmp_bytes;
  SPB_CUDA(ctx, cub::DeviceSelect::Flagged(tmp, tb, cub::CountingInputIterator<int32_t>(0), flags, vidx, d_cnt, in, ctx->stream));
  ctx->launches += 1;
  int32_t h[2] = {0, 0};  // V, P
  SPB_CUDA(ctx, cudaMemcpyAsync(&h[0], d_cnt, 4, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaMemcpyAsync(&h[1], tp + (n - 1), 4, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const int V = h[0];
  const long long P = h[1], N = n - P;
  if (P == 0 || N == 0) return fail(ctx, SPB_ERR_INVALID, "Only one class present in y_true. ROC AUC score is not defined in that case.");
  roc_keep_kernel<<<(unsigned)((V + 255) / 256), 256, 0, ctx->stream>>>(vidx, tp, V, flags);
  SPB_LAUNCH_CHECK(ctx, "roc_keep_kernel");
  tb = tmp_bytes;
  SPB_CUDA(ctx, cub::DeviceSelect::Flagged(tmp, tb, vidx, flags, kidx, d_cnt, V, ctx->stream));
  ctx->launches += 1;
  int32_t Kc = 0;
  SPB_CUDA(ctx, cudaMemcpyAsync(&Kc, d_cnt, 4, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (count) *count = (int64_t)Kc + 1;
  if ((int64_t)Kc + 1 > capacity) return fail(ctx, SPB_ERR_INVALID, "roc_curve: output capacity too small (n + 1 always suffices)");
  roc_emit_kernel<<<(unsigned)((Kc + 1 + 255) / 256), 256, 0, ctx->stream>>>(kidx, s, tp, Kc, (double)P, (double)N, fpr, tpr, thr);
  SPB_LAUNCH_CHECK(ctx, "roc_emit_kernel");
  const size_t ob = ((size_t)Kc + 1) * 8;
  if (h_fpr) SPB_CUDA(ctx, cudaMemcpyAsync(h_fpr, fpr, ob, cudaMemcpyDeviceToHost, ctx->stream));
  if (h_tpr) SPB_CUDA(ctx, cudaMemcpyAsync(h_tpr, tpr, ob, cudaMemcpyDeviceToHost, ctx->stream));
  if (h_thr) SPB_CUDA(ctx, cudaMemcpyAsync(h_thr, thr, ob, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

}  // namespace spb

using namespace spb;

#define SPB_REQ(ctx, cond, msg) \
  do {                          \
    if (!(cond)) return fail(ctx, SPB_ERR_INVALID, msg); \
  } while (0)

static int check_scores_args(spb_context *ctx, const void *E, int64_t m, const void *T, int64_t n, int32_t d,
                             const void *S) {
  SPB_REQ(ctx, ctx != nullptr, "cosine scores: NULL context");
  SPB_REQ(ctx, m >= 0 && n >= 0 && d >= 1, "cosine scores: bad shape");
  SPB_REQ(ctx, d <= 65536, "cosine scores: embedding dimension too large");
  if (m > 0 && n > 0) SPB_REQ(ctx, E && T && S, "cosine scores: NULL buffer");
  return SPB_OK;
}

extern "C" {

int spb_cosine_scores_dev(spb_context *ctx, const float *E, int64_t m, const float *T, int64_t n, int32_t d,
                          float *S) {
  int rc = check_scores_args(ctx, E, m, T, n, d, S);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return launch_cosine_scores(ctx, E, m, T, n, d, S);
}

int spb_cosine_scores_dev_noalloc(spb_context *ctx, const float *E, int64_t m, const float *T, int64_t n, int32_t d,
                                  float *S) {
  int rc = check_scores_args(ctx, E, m, T, n, d, S);
  if (rc) return rc;
  if (m == 0 || n == 0) return SPB_OK;
  const size_t need = cosine_scratch_floats(m, n, d);
  if (need > kNoAllocFloats)
    return fail(ctx, SPB_ERR_INVALID,
                "spb_cosine_scores_dev_noalloc: (m + n) x d needs " + std::to_string(need * 4) +
                    " bytes of split scratch, the static scratch holds " + std::to_string(kNoAllocFloats * 4) +
                    " (use spb_cosine_scores_dev)");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  float *buf = nullptr;
  if ((rc = noalloc_scratch(ctx, &buf))) return rc;
  // every element of the split that the GEMM reads is rewritten by the prologue of this call (zero
  // padding of the columns included; rows past m / n are zero-filled by TMA): nothing to clear
  return cosine_scores_in(ctx, E, m, T, n, d, S, buf, false);
}

int spb_cosine_scores(spb_context *ctx, const float *E, int64_t m, const float *T, int64_t n, int32_t d,
                      float *S) {
  int rc = check_scores_args(ctx, E, m, T, n, d, S);
  if (rc) return rc;
  if (m == 0 || n == 0) return SPB_OK;
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  float *d_E = (float *)scratch(ctx, 0, (size_t)m * d * 4);
  float *d_T = (float *)scratch(ctx, 5, (size_t)n * d * 4);
  float *d_S = (float *)scratch(ctx, 1, (size_t)m * n * 4);
  if (!d_E || !d_T || !d_S) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemcpyAsync(d_E, E, (size_t)m * d * 4, cudaMemcpyHostToDevice, ctx->stream));
  SPB_CUDA(ctx, cudaMemcpyAsync(d_T, T, (size_t)n * d * 4, cudaMemcpyHostToDevice, ctx->stream));
  rc = launch_cosine_scores(ctx, d_E, m, d_T, n, d, d_S);
  if (rc) return rc;
  SPB_CUDA(ctx, cudaMemcpyAsync(S, d_S, (size_t)m * n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  SPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SPB_OK;
}

int spb_trial_scores_dev(spb_context *ctx, const float *S, int64_t m, int64_t n, const int32_t *ti,
                         const int32_t *tj, int64_t nt, float *out) {
  SPB_REQ(ctx, ctx != nullptr, "trial scores: NULL context");
  SPB_REQ(ctx, m >= 0 && n >= 0 && nt >= 0, "trial scores: bad shape");
  if (nt == 0) return SPB_OK;
  SPB_REQ(ctx, S && ti && tj && out, "trial scores: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return launch_trial_gather(ctx, S, m, n, ti, tj, nt, out);
}

int spb_argmax_rows_dev(spb_context *ctx, const float *S, int64_t m, int64_t n, int32_t *out) {
  SPB_REQ(ctx, ctx != nullptr, "argmax: NULL context");
  SPB_REQ(ctx, m >= 0 && n >= 0, "argmax: bad shape");
  if (m == 0) return SPB_OK;
  SPB_REQ(ctx, S && out, "argmax: NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return launch_argmax_rows(ctx, S, m, n, out);
}

int spb_eer_dev(spb_context *ctx, const float *s, const int32_t *l, int64_t n, double *eer, double *auc) {
  SPB_REQ(ctx, ctx != nullptr, "eer: NULL context");
  SPB_REQ(ctx, n > 0 && s && l, "eer: no trials or NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return run_eer(ctx, s, l, n, eer, auc);
}

int spb_eer(spb_context *ctx, const float *s, const int32_t *l, int64_t n, double *eer, double *auc) {
  SPB_REQ(ctx, ctx != nullptr, "eer: NULL context");
  SPB_REQ(ctx, n > 0 && s && l, "eer: no trials or NULL buffer");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  float *d_s = (float *)scratch(ctx, 0, (size_t)n * 4);
  int32_t *d_l = (int32_t *)scratch(ctx, 1, (size_t)n * 4);
  if (!d_s || !d_l) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemcpyAsync(d_s, s, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  SPB_CUDA(ctx, cudaMemcpyAsync(d_l, l, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  return run_eer(ctx, d_s, d_l, n, eer, auc);
}

int spb_roc_curve_dev(spb_context *ctx, const float *s, const int32_t *l, int64_t n, double *fpr, double *tpr,
                      double *thresholds, int64_t capacity, int64_t *count) {
  SPB_REQ(ctx, ctx != nullptr, "roc_curve: NULL context");
  SPB_REQ(ctx, n > 0 && s && l, "roc_curve: no trials or NULL buffer");
  SPB_REQ(ctx, capacity >= 0 && count, "roc_curve: bad capacity or NULL count");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return run_roc_curve(ctx, s, l, n, fpr, tpr, thresholds, capacity, count);
}

int spb_roc_curve(spb_context *ctx, const float *s, const int32_t *l, int64_t n, double *fpr, double *tpr,
                  double *thresholds, int64_t capacity, int64_t *count) {
  SPB_REQ(ctx, ctx != nullptr, "roc_curve: NULL context");
  SPB_REQ(ctx, n > 0 && s && l, "roc_curve: no trials or NULL buffer");
  SPB_REQ(ctx, capacity >= 0 && count, "roc_curve: bad capacity or NULL count");
  SPB_CUDA(ctx, cudaSetDevice(ctx->device));
  float *d_s = (float *)scratch(ctx, 0, (size_t)n * 4);
  int32_t *d_l = (int32_t *)scratch(ctx, 1, (size_t)n * 4);
  if (!d_s || !d_l) return SPB_ERR_CUDA;
  SPB_CUDA(ctx, cudaMemcpyAsync(d_s, s, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  SPB_CUDA(ctx, cudaMemcpyAsync(d_l, l, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  return run_roc_curve(ctx, d_s, d_l, n, fpr, tpr, thresholds, capacity, count);
}

}  // extern "C"
