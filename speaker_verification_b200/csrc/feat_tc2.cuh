// features_tc2_kernel: the lmfe kernel of the batch paths (second generation of the hop-block DFT on
// tcgen05; included at the end of feat_tc_kernels.cu, whose tables / constant bank it shares).
//
// Same contract as features_tc_kernel (frame = 2 hops, X_f = A_f + e^{-2 pi i S k / NFFT} A_{f+1}),
// but every hop is folded about its centre before it reaches the tensor cores:
//
//   n = 79.5 +- (m + 1/2), m = 0..79:   s[m] = x[80+m] + x[79-m],   d[m] = x[80+m] - x[79-m]
//   A_h[k] = e^{-i 79.5 th_k} ( sum_m s[m] cos(th_k (m+1/2))  -  i sum_m d[m] sin(th_k (m+1/2)) )
//
// The common phase drops out of |X_f|^2, so the product per hop is TWO real K = 80 contractions
// (s . C into the "real" accumulator columns, d . (-S) into the "imaginary" ones) instead of one
// K = 160 contraction against (cos, -sin): half the tensor work and half the resident matrix
// (4 x 20 KB: C_hi, C_lo, S_hi, S_lo), which frees shared memory for a 9-slot ring of whole
// k-steps of the A operand (s_hi, s_lo, d_hi, d_lo; 16 KB per slot), i.e. the converters run up to
// 1.8 tiles ahead of the MMAs and no role waits on another inside a tile.
//
// fp32 accuracy from fp16 tensor cores as before: s 2^14 = s_hi + s_lo (exact for 16-bit PCM:
// 17 bits = 11 + 6), W = W_hi + W_lo; per tile the 20 small products first, then the 10 large ones
// (TMEM accumulation truncates).  |s 2^14| >= 65520 (|x| >= 2), NaN, inf raise the flag and the
// guarded CUDA-core FFT kernel redoes the launch.
//
// Epilogue: packed fp32 (FFMA2 / FMUL2) over bin PAIRS with the hop rotation and the mel weights
// as uniform-register operands, one (rise, fall) FFMA2 per bin for the segment-form mel sums.
// Three flavours of the segment walk:
//   EPI 2  the speechpy default bank (40 filters, 16 kHz, NFFT 512): segment ends, the split of
//          the filters between the two warps of a quadrant and the store schedule are
//          compile-time constants (no per-bin tests, no counters)
//   EPI 1  any bank whose halves hold multiples of four filters: log output in 16-byte runs
//   EPI 0  everything else (mfe, odd filter counts)
//
// Exactness guard (north_star: |log-mel error| <= 1e-4 on every value): the converters track
// max |s|, |d| per hop; a filter whose energy is below tau x the frame's level bound
// 80 (max_f^2 + max_{f+1}^2) >= sum x^2 cannot be trusted in fp32 (absolute error ~1e-7 of the
// level) and is appended to a list that redo_list_kernel (redo_kernels.cu) recomputes in float64.
// (Included inside namespace spb.)
#pragma once

constexpr int kT2K = 80;                        // K of each product (folded hop)
constexpr int kT2KSteps = kT2K / 16;            // 5 MMA k-steps
constexpr int kT2N = 128;                       // MMA N: bins (cos block) / bins (sin block)
constexpr int kT2MaxBins = 128;
#ifndef SPB_T2_SLOTS
#define SPB_T2_SLOTS 7
#endif
#ifndef SPB_T2_INTERLEAVE
#define SPB_T2_INTERLEAVE 0
#endif
#ifndef SPB_T2_PF
#define SPB_T2_PF 2  // L2 prefetch distance of the converters, in tiles of this CTA
#endif
#ifndef SPB_T2_STAGE
#define SPB_T2_STAGE 2
#endif
constexpr int kT2Slots = SPB_T2_SLOTS;          // A-operand ring depth (k-steps)
constexpr int kT2Stage = SPB_T2_STAGE;          // raw-sample staging depth (k-steps of cp.async in flight per converter)
constexpr uint32_t kT2BLbo = kT2N * 16;                 // 2 048 B between the k-chunks of a W matrix
constexpr uint32_t kT2BMat = (kT2K / 8) * kT2BLbo;      // 20 480 B: [k/8][col][8 halves]
// A-operand ring slot (one k-step): the first k-chunk column ([row][8 halves], 2 048 B) of s_hi | s_lo | d_hi | d_lo, 64 B of
// padding, then their second columns.  The padding puts the two columns of an array 64 bytes apart modulo 128, so that a
// half-warp of converters -- four rows x (two columns x two 8-byte halves) -- stores without bank conflicts.
constexpr uint32_t kT2APart = kTcRows * 16;             // 2 048 B: one column of one array; array k starts at k * kT2APart
constexpr uint32_t kT2ALbo = 4 * kT2APart + 64;         // 8 256 B between the two k-chunk columns of an array
constexpr uint32_t kT2SlotBytes = 8 * kT2APart + 64;    // 16 448 B
constexpr uint32_t kT2OffA = 4 * kT2BMat;               // C_hi | C_lo | S_hi | S_lo, then the ring
constexpr uint32_t kT2OffStage = kT2OffA + kT2Slots * kT2SlotBytes;       // [stage][4 loads][256 converters][16 B]
constexpr uint32_t kT2StageBytes = 4 * kTcCvtWarps * 32 * 16;              // 16 KB per k-step
constexpr uint32_t kT2OffBars = kT2OffStage + kT2Stage * kT2StageBytes;
constexpr uint32_t kT2OffLvl = kT2OffBars + 256;        // [4][128] per-hop level (max |s|, |d|), by tile & 3
constexpr uint32_t kT2Smem = kT2OffLvl + 4 * kTcRows * 4;
static_assert(kT2Smem <= 232448, "features_tc2: shared memory budget");
static_assert(kT2Stage >= 1 && kT2Stage <= kT2KSteps, "features_tc2: staging depth (at most one tile ahead)");
static_assert((2 * kT2Slots + 4) * 8 + 4 <= 256, "features_tc2: barrier block");
constexpr float kT2InScale = 16384.f;           // 2^14: s, d reach 2 |x|
constexpr uint32_t kT2Idesc = (1u << 4) | ((uint32_t)(kT2N >> 3) << 17) | ((uint32_t)(kTcRows >> 4) << 24);

// ------------------------------------------------------------------------------------------
// The speechpy default bank at compile time.  Edges = feature.filterbanks(40, 257, 16000, 0|300,
// 8000) (feature.py:55-82); build_tc_plan compares the runtime tables it derives
// from the caller's dense bank with these and only then selects the static epilogue.
// Bins are relative to kmin; segment s = bins (edge[s], edge[s+1]], the last one ends at bin 127.
// ------------------------------------------------------------------------------------------
#ifndef SPB_T2_PARTS
#define SPB_T2_PARTS 2
#endif
// warps per TMEM lane quadrant in the static epilogue (each walks a run of segments).  Measured on B200
// (10.8 audio-h): 2 parts 0.898 ms; 3 parts (21 warps, 80 registers per thread) 0.964 ms.
constexpr int kT2Parts = SPB_T2_PARTS;
static_assert(kT2Parts == 2 || kT2Parts == 3, "features_tc2: static epilogue parts");
struct T2Static {
  int M, kmin, nb;
  int sp[kT2Parts + 1];          // part p walks segments sp[p] .. sp[p+1] and emits filters sp[p] .. sp[p+1]-1
  int lo[kT2Parts], hi[kT2Parts];        // bins each part walks
  int out_lo[kT2Parts], nout[kT2Parts];  // filters each part emits
  uint32_t endmask[kT2Parts][4];  // bit b: bin b closes a segment of this part
};

constexpr T2Static make_t2_static() {
  constexpr int edges[42] = {4,  5,  6,  7,  8,  9,  10, 12, 13, 14, 16, 17, 19, 20,  22,  24,  26,  28,  30,  32, 35,
                             37, 40, 42, 45, 49, 52, 55, 59, 63, 67, 71, 75, 80, 85, 90, 96, 102, 108, 114, 121, 128};
  T2Static t{};
  t.M = 40;
  t.kmin = 5;
  t.nb = 123;
  // balance of bins x 7 + flushes x 11 + TMEM loads x 12 instructions per warp: two parts 0..25 | 26..39
  // (702 | 750), three parts 0..19 | 20..31 | 32..39 (498 | 506 | 511)
  if (kT2Parts == 2) {
    t.sp[0] = 0; t.sp[1] = 26; t.sp[kT2Parts] = 40;
  } else {
    t.sp[0] = 0; t.sp[1] = 20; t.sp[kT2Parts - 1] = 32; t.sp[kT2Parts] = 40;
  }
  int first[41] = {}, last[41] = {};
  for (int s = 0; s <= 40; ++s) {
    first[s] = edges[s] + 1 - t.kmin;
    last[s] = (edges[s + 1] < 127 ? edges[s + 1] : 127) - t.kmin;
  }
  for (int p = 0; p < kT2Parts; ++p) {
    t.lo[p] = (p == 0) ? 0 : first[t.sp[p]];
    t.hi[p] = (p == kT2Parts - 1) ? t.nb - 1 : last[t.sp[p + 1]];
    t.out_lo[p] = t.sp[p];
    t.nout[p] = t.sp[p + 1] - t.sp[p];
    for (int i = 0; i < 4; ++i) t.endmask[p][i] = 0;
    for (int s = t.sp[p]; s <= t.sp[p + 1]; ++s) t.endmask[p][last[s] >> 5] |= 1u << (last[s] & 31);
  }
  return t;
}
constexpr T2Static kT2StaticHost = make_t2_static();
__device__ constexpr T2Static kT2S = make_t2_static();

struct T2Params {
  const float *samples;
  const TcGroup *groups;     // 4 per tile
  long long num_tiles;
  const __half *w;           // C_hi | C_lo | S_hi | S_lo in the shared-memory layout (4 x kT2BMat bytes)
  TcHalf half[2];
  int M;
  float scale;               // 1 / (NFFT * 2^28)
  float ln_scale;            // ln(scale)
  int mfe;                   // 1: no log
  float *out;
  int *flag;                 // set when a sample cannot be split (|x| >= 2, NaN, inf)
  const float *zeros;        // >= 640 B of zeros: the hop rows past the end of a group read these
  float4 *hopsum;            // optional [tiles * 128]: per hop row {sum x^2, sum x, sum (-1)^n x, 0} for the mfcc energy
  RedoList redo;             // low-energy filters to recompute exactly (entries == nullptr: off)
  float pre_c;               // pre-emphasis coefficient (PRE kernels; processing.py:45-58, shift = 1)
  float pre_h[40];           // PRE: per filter (max |H_k| over its bins + 0.05)^2 -- what pre-emphasis does to the error of X,
                             // hence to the guard's threshold of that filter (0.05: rounding of the two boundary terms)
};


// ------------------------------------------------------------------------------------------
// mbarrier waits with a watchdog: a protocol bug (or a fault elsewhere in the CTA) must end as a
// launch error, not as a hung GPU.  The clock is read every 4096 polls only.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ bool t2_try_wait(const uint32_t bar_addr, const uint32_t parity, const uint32_t hint_ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(bar_addr), "r"(parity), "r"(hint_ns)
      : "memory");
  return ok != 0;
}
__device__ __noinline__ void t2_watchdog(unsigned long long &t0) {
  unsigned long long now;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
  if (t0 == 0) t0 = now;
  else if (now - t0 > 4000000000ull) __trap();  // 4 s
}
// the hardware may park the warp for up to the hint
__device__ __forceinline__ void t2_wait_park(uint64_t *bar, const uint32_t parity) {
  const uint32_t a = smem_u32(bar);
  uint32_t n = 0;
  unsigned long long t0 = 0;
  while (!t2_try_wait(a, parity, 20000u)) {
    if ((++n & 4095u) == 0) t2_watchdog(t0);
  }
}
// wait of a warp that is normally AHEAD of its producer (the converters): back off between polls
__device__ __forceinline__ void t2_wait_nap(uint64_t *bar, const uint32_t parity) {
  const uint32_t a = smem_u32(bar);
  if (t2_try_wait(a, parity, 0u)) return;
  uint32_t n = 0;
  unsigned long long t0 = 0;
  do {
    __nanosleep(96);
    if ((++n & 4095u) == 0) t2_watchdog(t0);
  } while (!t2_try_wait(a, parity, 0u));
}

// ------------------------------------------------------------------------------------------
// converter arithmetic: four forward samples f = x[80+m .. 80+m+3] and the four mirrored ones
// b = x[76-m .. 79-m] -> s[i] = f[i] + b[3-i], d[i] = f[i] - b[3-i] (scaled by 2^14, one rounding:
// fma(b, 2^14, f 2^14)), each split into fp16 hi + lo.  mx: running max of the |hi| bit patterns
// (two 16-bit lanes; >= 0x7c00 means inf / NaN).  Packed fp32 throughout; the mirrored pair is
// the LO_HI operand swizzle of FFMA2, not a move.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t t2_pack_h2(const float lo, const float hi) {
  const __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t *>(&h);
}

struct T2Fold {
  float2 s01, s23, d01, d23;  // scaled s, d of the four m values
};

__device__ __forceinline__ T2Fold t2_fold_split(const float4 f, const float4 b, uint2 &shi, uint2 &slo, uint2 &dhi,
                                                uint2 &dlo, uint32_t &mx) {
  const float2 sc = make_float2(kT2InScale, kT2InScale), nsc = make_float2(-kT2InScale, -kT2InScale);
  const float2 a01 = __fmul2_rn(make_float2(f.x, f.y), sc), a23 = __fmul2_rn(make_float2(f.z, f.w), sc);
  T2Fold r;
  r.s01 = __ffma2_rn(make_float2(b.w, b.z), sc, a01);
  r.s23 = __ffma2_rn(make_float2(b.y, b.x), sc, a23);
  r.d01 = __ffma2_rn(make_float2(b.w, b.z), nsc, a01);
  r.d23 = __ffma2_rn(make_float2(b.y, b.x), nsc, a23);
  const __half2 hs0 = __floats2half2_rn(r.s01.x, r.s01.y), hs1 = __floats2half2_rn(r.s23.x, r.s23.y);
  const __half2 hd0 = __floats2half2_rn(r.d01.x, r.d01.y), hd1 = __floats2half2_rn(r.d23.x, r.d23.y);
  const float2 fs0 = __half22float2(hs0), fs1 = __half22float2(hs1);
  const float2 fd0 = __half22float2(hd0), fd1 = __half22float2(hd1);
  shi.x = *reinterpret_cast<const uint32_t *>(&hs0);
  shi.y = *reinterpret_cast<const uint32_t *>(&hs1);
  dhi.x = *reinterpret_cast<const uint32_t *>(&hd0);
  dhi.y = *reinterpret_cast<const uint32_t *>(&hd1);
  const float2 ls0 = __fadd2_rn(r.s01, make_float2(-fs0.x, -fs0.y)), ls1 = __fadd2_rn(r.s23, make_float2(-fs1.x, -fs1.y));
  const float2 ld0 = __fadd2_rn(r.d01, make_float2(-fd0.x, -fd0.y)), ld1 = __fadd2_rn(r.d23, make_float2(-fd1.x, -fd1.y));
  slo.x = t2_pack_h2(ls0.x, ls0.y);
  slo.y = t2_pack_h2(ls1.x, ls1.y);
  dlo.x = t2_pack_h2(ld0.x, ld0.y);
  dlo.y = t2_pack_h2(ld1.x, ld1.y);
  mx = __vimax3_s16x2(mx, shi.x & 0x7fff7fffu, shi.y & 0x7fff7fffu);
  mx = __vimax3_s16x2(mx, dhi.x & 0x7fff7fffu, dhi.y & 0x7fff7fffu);
  return r;
}

// per-hop sums of the mfcc energy path from the folded values (scaled by 2^14):
// sum x = sum s, sum (-1)^n x = sum (-1)^m d (n = 80+m and 79-m have opposite parities),
// sum x^2 = (sum s^2 + d^2) / 2
__device__ __forceinline__ void t2_hop_sums(const T2Fold &v, float &q2, float &q0, float &qa) {
  q2 = fmaf(v.s01.x, v.s01.x, fmaf(v.s01.y, v.s01.y, fmaf(v.s23.x, v.s23.x, fmaf(v.s23.y, v.s23.y, q2))));
  q2 = fmaf(v.d01.x, v.d01.x, fmaf(v.d01.y, v.d01.y, fmaf(v.d23.x, v.d23.x, fmaf(v.d23.y, v.d23.y, q2))));
  q0 += (v.s01.x + v.s01.y) + (v.s23.x + v.s23.y);
  qa += (v.d01.x - v.d01.y) + (v.d23.x - v.d23.y);
}

// fp16 bit pattern (positive) -> float
__device__ __forceinline__ float t2_level_of(const uint32_t bits16) {
  return __half2float(__ushort_as_half((unsigned short)bits16));
}

// log-mel value of one filter from its unscaled energy a (x 2^28 NFFT): feature.py:216-217,256
__device__ __forceinline__ float t2_log_value(const float a, const float ln_scale) {
  float lg;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(a));
  const float r = fmaf(lg, 0.69314718055994531f, ln_scale);
  return (a == 0.f) ? -36.04365338911715f : r;  // functions.zero_handling: exact zero -> eps
}

// hop rotation e^{-2 pi i 160 k / 512} = (cos, -sin)(2 pi (5k mod 16) / 16) of the bin pairs (j, j+1), j even, of a
// 16-bin chunk when kmin = 5 (mod 16)
struct T2Rot {
  float2 cr[8], ci[8];
};
__constant__ T2Rot c_rot5 = {{{-0.9238795325112868f, 0.7071067811865474f}, {0.38268343236508984f, -1.0f}, {0.38268343236509f, 0.7071067811865476f}, {-0.9238795325112867f, 0.f}, {0.9238795325112867f, -0.7071067811865475f}, {-0.38268343236509034f, 1.0f}, {-0.3826834323650897f, -0.7071067811865477f}, {0.9238795325112865f, 0.f}},
                             {{0.38268343236508967f, 0.7071067811865477f}, {-0.9238795325112867f, 0.f}, {0.9238795325112866f, -0.7071067811865475f}, {-0.3826834323650899f, 1.0f}, {-0.3826834323650898f, -0.7071067811865476f}, {0.9238795325112865f, 0.f}, {-0.9238795325112867f, 0.7071067811865475f}, {0.3826834323650904f, -1.0f}}};

// Pre-emphasis in the spectral domain (static epilogue, PRE kernels).  With y[n] = x[n] - c x[n-1] the 512-point
// transform of a 320-sample frame that starts at sample s is
//     Y[k] = (1 - c W^k) X[k] - c x[s-1] + c x[s+319] W^(320 k),      W = exp(-2 pi i / 512),
// (shift the sum by one sample: the frame gains x[s-1] in front and loses x[s+319] at the end).  The epilogue has
// X[k] up to the phase of the folded transform -- X' = P_k X, P_k = exp(+i th_k 79.5): the hop blocks are folded
// about their centres --, so Y' = H_k X' + P_k (-c x[s-1]) + Q_k (c x[s+319]) with Q_k = P_k W^(320 k).  H_k sits in
// .z / .w of the static variant's table (the compile-time rotation leaves them free), (P_k, Q_k) in the table of the
// second half, which the static epilogue does not use.  Two samples per frame, eight packed operations per bin pair.
// (xr, xi) = X' of the bin pair (b, b + 1) -> |Y|^2; pa = -c x[s-1], pe = c x[s+319] in the accumulator's units
__device__ __forceinline__ float2 t2_pre_power(const float2 xr, const float2 xi, const float4 *h, const float4 *pq, const float pa,
                                               const float pe) {
  const float2 hr = make_float2(h[0].z, h[1].z), hi = make_float2(h[0].w, h[1].w);
  const float2 a2 = make_float2(pa, pa), e2 = make_float2(pe, pe);
  const float2 tr = __ffma2_rn(e2, make_float2(pq[0].z, pq[1].z), __fmul2_rn(a2, make_float2(pq[0].x, pq[1].x)));
  const float2 ti = __ffma2_rn(e2, make_float2(pq[0].w, pq[1].w), __fmul2_rn(a2, make_float2(pq[0].y, pq[1].y)));
  float2 yr = __ffma2_rn(hr, xr, tr);
  yr = __ffma2_rn(make_float2(-hi.x, -hi.y), xi, yr);
  float2 yi = __ffma2_rn(hr, xi, ti);
  yi = __ffma2_rn(hi, xr, yi);
  return __ffma2_rn(yi, yi, __fmul2_rn(yr, yr));
}

// |A_own + rot A_next|^2 of one bin pair (j, j + 1) of a 16-bin chunk.  KM = kmin mod 16 (rotation
// as compile-time constants, period 16 in k at S = 160, NFFT = 512) or -1 (rotation from the table).
template <int KM, bool PRE = false>
__device__ __forceinline__ float2 t2_pair_power(const uint32_t (&vr)[16], const uint32_t (&vi)[16], const int j,
                                                const float4 *tb, const float pre_a = 0.f, const float pre_e = 0.f) {
  float2 cr, ci;
  if (KM == 5) {
    // from the constant bank, not as immediates: ptxas then parks the 16 pairs in uniform registers
    // for the whole epilogue instead of re-materialising two UMOVs in front of every FFMA2
    cr = c_rot5.cr[j >> 1];
    ci = c_rot5.ci[j >> 1];
  } else if (KM >= 0) {
    cr = make_float2(kCos16[(5 * (KM + j)) & 15], kCos16[(5 * (KM + j + 1)) & 15]);
    ci = make_float2(-kSin16[(5 * (KM + j)) & 15], -kSin16[(5 * (KM + j + 1)) & 15]);
  } else {
    cr = make_float2(tb[j].z, tb[j + 1].z);
    ci = make_float2(tb[j].w, tb[j + 1].w);
  }
  const float2 a_re = make_float2(__uint_as_float(vr[j]), __uint_as_float(vr[j + 1]));
  const float2 a_im = make_float2(__uint_as_float(vi[j]), __uint_as_float(vi[j + 1]));
  float2 b_re, b_im;
  b_re.x = __shfl_down_sync(0xffffffffu, a_re.x, 1);
  b_re.y = __shfl_down_sync(0xffffffffu, a_re.y, 1);
  b_im.x = __shfl_down_sync(0xffffffffu, a_im.x, 1);
  b_im.y = __shfl_down_sync(0xffffffffu, a_im.y, 1);
  float2 xr = __ffma2_rn(cr, b_re, a_re);
  xr = __ffma2_rn(make_float2(-ci.x, -ci.y), b_im, xr);
  float2 xi = __ffma2_rn(cr, b_im, a_im);
  xi = __ffma2_rn(ci, b_re, xi);
  if (PRE) return t2_pre_power(xr, xi, tb + j, tb + 128 + j, pre_a, pre_e);  // tab[1] follows tab[0]
  return __ffma2_rn(xi, xi, __fmul2_rn(xr, xr));
}


// ------------------------------------------------------------------------------------------
// Epilogue of one tile for one warp, run-time bank (EPI 0 / 1).  lane = hop row of the quadrant,
// frame = hops (lane, lane + 1).  HALF is a template parameter so that every quantity the segment
// walk branches on (chunk range, end masks) comes from kernel parameters and the loop counter only:
// ptxas then keeps them in uniform registers and the per-bin branch is a BRA.U without
// reconvergence bookkeeping.  The chunk loop is NOT unrolled (instruction cache).
// thr: tau x level of this lane's frame (negative for rows that are not frames: never flagged).
// ------------------------------------------------------------------------------------------
template <int KM, int EPI, int HALF>
__device__ __forceinline__ void t2_epilogue_dyn(const T2Params &p, const uint32_t taddr, float *pout, const bool row_ok,
                                                const float thr, const long long soff, const long long oidx0) {
  constexpr bool FAST = (EPI == 1);
  const int ch_begin = p.half[HALF].ch_begin, ch_end = p.half[HALF].ch_end;
  const uint32_t m0 = p.half[HALF].endmask[0], m1 = p.half[HALF].endmask[1];
  const uint32_t m2 = p.half[HALF].endmask[2], m3 = p.half[HALF].endmask[3];
  float2 rf = make_float2(0.f, 0.f);  // (rise, fall) sums of the current segment
  float prev_u = 0.f;
  // FAST: outputs leave in runs of four filters (one 16-byte store per lane).  `left` counts down to
  // the next store; it starts at 5 because the first flush of a half only latches U of its first
  // segment (filter j = U[j] + D[j+1]); that value shifts through o2..o0 and is gone before the
  // first store.
  float o0 = 0.f, o1 = 0.f, o2 = 0.f;
  int left = 5, q4 = 0;
  int skip = 1;                              // generic flush: the latch-only first flush
  uint32_t bit = 0x80000000u, fbits = 0u;    // flagged filters; bit 31 absorbs the first flush
  const float ln_scale = p.ln_scale, scale = p.scale;
  const bool mfe = !FAST && p.mfe != 0;
#pragma unroll 1
  for (int ch = ch_begin; ch < ch_end; ++ch) {
    uint32_t vr[16], vi[16];
    tmem_ld16x2(taddr + ch * 16, taddr + kT2N + ch * 16, vr, vi);
    const uint32_t mw = (ch < 4) ? ((ch < 2) ? m0 : m1) : ((ch < 6) ? m2 : m3);
    const uint32_t mask = mw >> ((ch & 1) * 16);
    const float4 *tb = &c_tc.tab[HALF][ch * 16];
    float pw[16];
#pragma unroll
    for (int j = 0; j < 16; j += 2) {
      const float2 pp = t2_pair_power<KM>(vr, vi, j, tb);
      pw[j] = pp.x;
      pw[j + 1] = pp.y;
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float2 w = *reinterpret_cast<const float2 *>(&tb[j]);
      rf = __ffma2_rn(w, make_float2(pw[j], pw[j]), rf);
      if ((mask >> j) & 1u) {
        const float a = prev_u + rf.y;  // unscaled (x 2^28 NFFT): never denormal for real input
        fbits |= (a < thr) ? bit : 0u;
        bit = __funnelshift_l(bit, bit, 1);
        if (FAST) {
          const float r = t2_log_value(a, ln_scale);
          --left;
          if (left == 0 && row_ok) reinterpret_cast<float4 *>(pout)[q4] = make_float4(o0, o1, o2, r);
          q4 += (left == 0) ? 1 : 0;
          left = (left == 0) ? 4 : left;
          o0 = o1; o1 = o2; o2 = r;
        } else if (skip == 0) {
          float r = mfe ? a * scale : t2_log_value(a, ln_scale);
          if (mfe && a == 0.f) r = kTcEps;  // functions.zero_handling (feature.py:217)
          if (row_ok) *pout = r;
          ++pout;
        }
        skip = 0;
        prev_u = rf.x;
        rf = make_float2(0.f, 0.f);
      }
    }
  }
  fbits &= 0x7fffffffu;
  if (fbits != 0u && p.redo.entries != nullptr) {
    while (fbits) {
      const int f = __ffs((int)fbits) - 1;
      fbits &= fbits - 1;
      redo_push(p.redo, soff, oidx0 + f);
    }
  }
}

// ------------------------------------------------------------------------------------------
// Epilogue for the speechpy default bank (EPI 2): everything the walk depends on is a compile-time
// constant (kT2S), the code is straight-line over the bins this half needs (~650 instructions per
// half), the filter index of every flush is static: results sit in four named registers until their
// 16-byte (or, at the split 26 | 14, 8-byte) store.
// prow: row base (filter 0) of this lane's output row.
// ------------------------------------------------------------------------------------------
// filters g0..m of a row (a run inside one group of eight, ov[e & 7] = filter e) with the widest stores
// their alignment allows: a whole group is ONE 32-byte store = one full sector of the row (rows are
// 160 B apart: 16-byte stores from lanes that each own a row filled every sector in two visits)
template <int HALF>
__device__ __forceinline__ void t2_store_run(float *prow, const float (&ov)[8], const int g0, const int m, const bool row_ok) {
  if (!row_ok) return;
  if (m - g0 + 1 == 8) {
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(prow + g0), "f"(ov[0]), "f"(ov[1]),
                 "f"(ov[2]), "f"(ov[3]), "f"(ov[4]), "f"(ov[5]), "f"(ov[6]), "f"(ov[7])
                 : "memory");
    return;
  }
  // a shorter run (g0 and m are compile-time values at every call site): the widest aligned stores, chosen while
  // the loop unrolls, so that ov[] is only ever indexed statically and stays in registers
  int skip = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int e = g0 + k;
    if (e > m) continue;
    if (skip > 0) {
      --skip;
      continue;
    }
    if ((e & 3) == 0 && e + 3 <= m) {
      *reinterpret_cast<float4 *>(prow + e) = make_float4(ov[e & 7], ov[(e + 1) & 7], ov[(e + 2) & 7], ov[(e + 3) & 7]);
      skip = 3;
    } else if ((e & 1) == 0 && e + 1 <= m) {
      *reinterpret_cast<float2 *>(prow + e) = make_float2(ov[e & 7], ov[(e + 1) & 7]);
      skip = 1;
    } else {
      prow[e] = ov[e & 7];
    }
  }
}

template <int HALF, bool PRE>
__device__ __forceinline__ void t2_epilogue_static(const T2Params &p, const uint32_t taddr, float *prow, const bool row_ok,
                                                   const float thr, const long long soff, const long long oidx_row,
                                                   const float pre_a, const float pre_e) {
  constexpr int LO = kT2S.lo[HALF], HI = kT2S.hi[HALF];
  constexpr int OUT_LO = kT2S.out_lo[HALF], NOUT = kT2S.nout[HALF];
  constexpr int KM = kT2S.kmin & 15;
  float2 rf = make_float2(0.f, 0.f);
  float prev_u = 0.f;
  float ov[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  uint32_t fbits = 0u;
  int nfl = 0;  // flushes so far: a compile-time value once the loops below are unrolled
  const float ln_scale = p.ln_scale;
#pragma unroll
  for (int ch = 0; ch < 8; ++ch) {
    if (ch * 16 + 15 < LO || ch * 16 > HI) continue;
    uint32_t vr[16], vi[16];
    tmem_ld16x2(taddr + ch * 16, taddr + kT2N + ch * 16, vr, vi);
    const float4 *tb = &c_tc.tab[0][ch * 16];  // the static variant's table is not masked per part
    float pw[16];
#pragma unroll
    for (int j = 0; j < 16; j += 2) {
      const int b = ch * 16 + j;
      if (b + 1 < LO || b > HI) continue;
      const float2 pp = t2_pair_power<KM, PRE>(vr, vi, j, tb, pre_a, pre_e);
      pw[j] = pp.x;
      pw[j + 1] = pp.y;
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int b = ch * 16 + j;
      if (b < LO || b > HI) continue;
      const float2 w = *reinterpret_cast<const float2 *>(&tb[j]);
      rf = __ffma2_rn(w, make_float2(pw[j], pw[j]), rf);
      if ((kT2S.endmask[HALF][b >> 5] >> (b & 31)) & 1u) {
        if (nfl >= 1) {
          const int m = OUT_LO + nfl - 1;  // filter this flush completes
          const float a = prev_u + rf.y;
          fbits |= (a < (PRE ? thr * p.pre_h[m] : thr)) ? (1u << (nfl - 1)) : 0u;
          ov[m & 7] = t2_log_value(a, ln_scale);
          if ((m & 7) == 7 || nfl == NOUT) {
            const int g0 = ((m & ~7) > OUT_LO) ? (m & ~7) : OUT_LO;
            t2_store_run<HALF>(prow, ov, g0, m, row_ok);
          }
        }
        ++nfl;
        prev_u = rf.x;
        rf = make_float2(0.f, 0.f);
      }
    }
  }
  if (fbits != 0u && p.redo.entries != nullptr) {
    while (fbits) {
      const int f = __ffs((int)fbits) - 1;
      fbits &= fbits - 1;
      redo_push(p.redo, soff, oidx_row + OUT_LO + f);
    }
  }
}


// ------------------------------------------------------------------------------------------
// The kernel.  One persistent CTA per SM, 17 warps (21 with the static epilogue: three warps per quadrant):
//   warps 0..7   epilogue (quadrant = warp % 4 owns hop rows 32q..32q+31 = group q of the tile,
//   (0..11)      part = warp / 4 owns a range of filters)
//   warp  8 (12) MMA issuer (one elected thread): per k-step s_lo.C_hi, s_hi.C_lo -> real columns,
//                d_lo.S_hi, d_hi.S_lo -> imaginary columns; then per k-step s_hi.C_hi, d_hi.S_hi and
//                a tcgen05.commit that hands the ring slot back; M128 N128 K16 kind::f16
//   the next 8   converters: lane = (k-chunk, row of 8, 8-byte half) so that a half-warp's
//                st.shared.v2 covers 128 contiguous bytes of the no-swizzle K-major operand
//                (conflict-free); 128-bit sample loads one k-step ahead + L2 bulk prefetch two
//                tiles ahead
// ------------------------------------------------------------------------------------------
// warps of a CTA: 4 x parts epilogue warps, the MMA warp, 8 converter warps
#ifndef SPB_T2_CVT3
#define SPB_T2_CVT3 4
#endif
__host__ __device__ constexpr int t2_epi_warps(const int epi) { return 4 * (epi == 2 ? kT2Parts : 2); }
// converter warps: 8 (two hop rows per thread), or SPB_T2_CVT3 beside a three-part static epilogue
__host__ __device__ constexpr int t2_cvt_warps(const int epi) { return (epi == 2 && kT2Parts == 3) ? SPB_T2_CVT3 : kTcCvtWarps; }
__host__ __device__ constexpr int t2_threads(const int epi) { return (t2_epi_warps(epi) + 1 + t2_cvt_warps(epi)) * 32; }

template <int KM, int EPI, bool SUMS, bool PRE = false>
__global__ void __launch_bounds__(t2_threads(EPI), 1) features_tc2_kernel(const T2Params p) {
  static_assert(!PRE || (EPI == 2 && !SUMS), "pre-emphasis: static epilogue, log-mel output only");
  constexpr int kEpiW = t2_epi_warps(EPI), kMmaW = kEpiW, kCvt0 = kEpiW + 1, kCvtW = t2_cvt_warps(EPI);
  constexpr int kThreads = t2_threads(EPI);
  extern __shared__ __align__(128) unsigned char tc_smem[];
  unsigned char *base = tc_smem;
  uint64_t *bars = reinterpret_cast<uint64_t *>(base + kT2OffBars);
  uint64_t *full = bars;                          // [slots] k-step converted (8 converter warps)
  uint64_t *empty = bars + kT2Slots;              // [slots] k-step consumed (tcgen05.commit)
  uint64_t *acc_full = bars + 2 * kT2Slots;       // [2]
  uint64_t *acc_empty = acc_full + 2;             // [2] 8 epilogue warps
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + 2);
  float *lvl = reinterpret_cast<float *>(base + kT2OffLvl);  // [4][128]

  // the shuffle tells ptxas that the warp index is warp-uniform: role branches become uniform branches and
  // uniform registers (memory descriptors, rotation constants) stay live across them
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(base);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kT2Slots; ++s) {
      mbar_init(&full[s], kCvtW);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], kEpiW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kMmaW) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  {  // resident DFT matrices
    const uint4 *g = reinterpret_cast<const uint4 *>(p.w);
    uint4 *s = reinterpret_cast<uint4 *>(base);
    for (uint32_t i = threadIdx.x; i < 4 * kT2BMat / 16; i += kThreads) s[i] = __ldg(g + i);
    fence_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= kCvt0) {
    // ===== converters =====
    // lane = (row of 8, k-chunk, 8-byte half): the four lanes of a row are adjacent, so their four 16-byte cp.async
    // pieces of a k-step are 64 contiguous bytes of the hop and reach shared memory as ONE write wavefront (with the
    // row's pieces 16 lanes apart every 32-byte sector arrived on its own: ncu r02ar counted 16 wavefronts per
    // LDGSTS.128 where 4 is the ideal -- half of all the kernel's shared-memory wavefronts); a thread feeds RPT
    // hop rows (r0 + 8 CW j)
    constexpr int CW = kCvtW, RPT = kTcRows / (CW * 8), NLD = 2 * RPT;
    constexpr uint32_t kPlane = CW * 32 * 16;      // one 16-byte load of every converter thread
    static_assert(NLD * kPlane == kT2StageBytes, "features_tc2: staging layout");
    const int cw = warp - kCvt0;
    const int kc = (lane >> 1) & 1, r8 = lane >> 2, hf = lane & 1;
    const int qq = 2 * kc + hf;                    // which 4 of the 16 m values of a k-step
    const int r0 = cw * 8 + r8;
    const uint32_t d0 = (uint32_t)(kc * kT2ALbo + r0 * 16 + hf * 8);
    const int c = cw * 32 + lane;
    auto row_ptr = [&](long long t, int r) -> const float * {
      if (t >= p.num_tiles) return p.zeros;
      const TcGroup g = p.groups[t * 4 + (r >> 5)];
      const int i = r & 31;
      return (g.nf > 0 && i <= g.nf) ? p.samples + g.gstart + (long long)i * kTcHop : p.zeros;
    };
    // forward samples of k-step s: x[80 + 16 s + 4 qq ..], mirrored: x[76 - 16 s - 4 qq ..]
    const int foff = kT2K + 4 * qq, boff = kT2K - 4 - 4 * qq;
    // Raw samples reach the converter through cp.async (LDGSTS) into a private staging area: kT2Stage
    // k-steps (2 RPT x 16 B each) per thread are in flight at any time without holding registers -- with
    // plain loads one k-step ahead the converters sat on the load latency (ncu r02l: long_scoreboard
    // 6.5 warps per issue).  stage[d][load][c]: 16 B per converter thread, a warp covers 512 contiguous
    // bytes on the way in and on the way out; a thread reads back only what it copied itself, so
    // cp.async.wait_group is all the synchronisation there is.
    const uint32_t stg = sbase + kT2OffStage + (uint32_t)c * 16u;
    const float *pR[RPT], *nR[RPT];
    auto stage_issue = [&](const uint32_t d, const float *const (&q)[RPT], const int st) {
      const uint32_t a = stg + d * kT2StageBytes;
#pragma unroll
      for (int j = 0; j < RPT; ++j) {
        cp_async16(a + (2 * j) * kPlane, q[j] + foff + 16 * st);
        cp_async16(a + (2 * j + 1) * kPlane, q[j] + boff - 16 * st);
      }
      cp_async_commit();
    };
    long long t = blockIdx.x;
#pragma unroll
    for (int j = 0; j < RPT; ++j) pR[j] = row_ptr(t, r0 + j * CW * 8);
#pragma unroll
    for (int d = 0; d < kT2Stage; ++d) stage_issue((uint32_t)d, pR, d);
    uint32_t it = 0, slot = 0, par = 1;  // tiles produced so far, ring slot and its wait parity
    uint32_t sd = 0;                     // staging slot of the k-step about to be converted
    uint32_t bad = 0;
    for (; t < p.num_tiles; t += gridDim.x, ++it) {
#pragma unroll
      for (int j = 0; j < RPT; ++j) nR[j] = row_ptr(t + gridDim.x, r0 + j * CW * 8);
      uint32_t mx[RPT];
      float q2[RPT], q0[RPT], qa[RPT];
#pragma unroll
      for (int j = 0; j < RPT; ++j) { mx[j] = 0u; q2[j] = 0.f; q0[j] = 0.f; qa[j] = 0.f; }
      if (c < 4) {
        // pull the samples of the tile after next into L2 with the TMA engine (one run per group)
        const long long tp = t + SPB_T2_PF * (long long)gridDim.x;
        if (tp < p.num_tiles) {
          const TcGroup g = p.groups[tp * 4 + c];
          if (g.nf > 0) bulk_prefetch_l2(p.samples + g.gstart, (uint32_t)(g.nf + 1) * (kTcHop * 4));
        }
      }
#pragma unroll
      for (int s = 0; s < kT2KSteps; ++s) {
        cp_async_wait<kT2Stage - 1>();  // the oldest k-step in flight has landed
        const uint32_t a = stg + sd * kT2StageBytes;
        float4 cf[RPT], cb[RPT];
#pragma unroll
        for (int j = 0; j < RPT; ++j) {
          cf[j] = lds128(a + (2 * j) * kPlane);
          cb[j] = lds128(a + (2 * j + 1) * kPlane);
        }
        uint2 sh[RPT], sl_[RPT], dh[RPT], dl[RPT];
#pragma unroll
        for (int j = 0; j < RPT; ++j) {
          const T2Fold v = t2_fold_split(cf[j], cb[j], sh[j], sl_[j], dh[j], dl[j], mx[j]);
          if (SUMS) t2_hop_sums(v, q2[j], q0[j], qa[j]);
          // keep the conversion ahead of the wait (the compiler otherwise sinks it below)
          asm volatile("" ::"r"(sh[j].x), "r"(sh[j].y), "r"(sl_[j].x), "r"(sl_[j].y), "r"(dh[j].x), "r"(dh[j].y),
                       "r"(dl[j].x), "r"(dl[j].y));
        }
        // the staged values are in registers (consumed above): refill the slot with the k-step kT2Stage ahead,
        // which lies in this tile or the next (past the last tile row_ptr hands out the zero block)
        if (s + kT2Stage < kT2KSteps) stage_issue(sd, pR, s + kT2Stage);
        else stage_issue(sd, nR, s + kT2Stage - kT2KSteps);
        sd = (sd + 1 == kT2Stage) ? 0u : sd + 1;
        if (s == kT2KSteps - 1) {
          // per-hop level for the epilogue's exactness guard (and the unsplittable-sample flag): max over
          // the four lanes of a row, written BEFORE the arrive of the tile's last k-step (the epilogue
          // reads it after acc_full, which is ordered after every full[] arrive of the tile)
#pragma unroll
          for (int j = 0; j < RPT; ++j) {
            mx[j] = __vmaxu2(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 1));
            mx[j] = __vmaxu2(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 2));
            const uint32_t m = max(mx[j] & 0xffffu, mx[j] >> 16);
            bad |= (m >= 0x7c00u) ? 1u : 0u;
            if ((lane & 3) == 0) lvl[(it & 3) * kTcRows + r0 + j * CW * 8] = t2_level_of(m);
          }
        }
        t2_wait_nap(&empty[slot], par);
        unsigned char *sl = base + kT2OffA + slot * kT2SlotBytes;
#pragma unroll
        for (int j = 0; j < RPT; ++j) {
          const uint32_t d = d0 + (uint32_t)(j * CW * 8 * 16);
          *reinterpret_cast<uint2 *>(sl + d) = sh[j];
          *reinterpret_cast<uint2 *>(sl + kT2APart + d) = sl_[j];
          *reinterpret_cast<uint2 *>(sl + 2 * kT2APart + d) = dh[j];
          *reinterpret_cast<uint2 *>(sl + 3 * kT2APart + d) = dl[j];
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[slot]);
        if (++slot == kT2Slots) { slot = 0; par ^= 1; }
      }
      if (SUMS) {
        // the four lanes of a row: lane bits 0 and 1
        constexpr float u1 = 1.0f / kT2InScale, u2 = 0.5f / (kT2InScale * kT2InScale);
#pragma unroll
        for (int j = 0; j < RPT; ++j) {
          q2[j] += __shfl_xor_sync(0xffffffffu, q2[j], 1); q0[j] += __shfl_xor_sync(0xffffffffu, q0[j], 1); qa[j] += __shfl_xor_sync(0xffffffffu, qa[j], 1);
          q2[j] += __shfl_xor_sync(0xffffffffu, q2[j], 2); q0[j] += __shfl_xor_sync(0xffffffffu, q0[j], 2); qa[j] += __shfl_xor_sync(0xffffffffu, qa[j], 2);
          if ((lane & 3) == 0) p.hopsum[t * kTcRows + r0 + j * CW * 8] = make_float4(q2[j] * u2, q0[j] * u1, qa[j] * u1, 0.f);
        }
      }
#pragma unroll
      for (int j = 0; j < RPT; ++j) pR[j] = nR[j];
    }
    cp_async_wait<0>();  // nothing may still be landing in shared memory when the CTA retires
    if (bad) *p.flag = 1;
  } else if (warp == kMmaW) {
    // ===== MMA issuer =====
    if (elect_one()) {
      const uint32_t w_chi = sbase, w_clo = sbase + kT2BMat, w_shi = sbase + 2 * kT2BMat, w_slo = sbase + 3 * kT2BMat;
      uint32_t slot = 0, par = 0, it = 0;
      for (long long t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++it) {
        const uint32_t acc = it & 1;
        t2_wait_park(&acc_empty[acc], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tm_re = tmem_base + acc * (2 * kT2N), tm_im = tm_re + kT2N;
        const uint32_t slot0 = slot;
        // small terms first, into the empty accumulators
#pragma unroll
        for (int s = 0; s < kT2KSteps; ++s) {
          t2_wait_park(&full[slot], par);
          tc_fence_after();
          const uint32_t a = sbase + kT2OffA + slot * kT2SlotBytes;
          const uint64_t d_shi = umma_desc_plain(a, kT2ALbo), d_slo = umma_desc_plain(a + kT2APart, kT2ALbo);
          const uint64_t d_dhi = umma_desc_plain(a + 2 * kT2APart, kT2ALbo);
          const uint64_t d_dlo = umma_desc_plain(a + 3 * kT2APart, kT2ALbo);
          const uint32_t ko = s * 2 * kT2BLbo;
          umma_f16(tm_re, d_slo, umma_desc_plain(w_chi + ko, kT2BLbo), kT2Idesc, s ? 1u : 0u);
          umma_f16(tm_re, d_shi, umma_desc_plain(w_clo + ko, kT2BLbo), kT2Idesc, 1u);
          umma_f16(tm_im, d_dlo, umma_desc_plain(w_shi + ko, kT2BLbo), kT2Idesc, s ? 1u : 0u);
          umma_f16(tm_im, d_dhi, umma_desc_plain(w_slo + ko, kT2BLbo), kT2Idesc, 1u);
#if SPB_T2_INTERLEAVE
          // experiment: the large terms of the k-step right away, the ring slot handed back at once
          umma_f16(tm_re, d_shi, umma_desc_plain(w_chi + ko, kT2BLbo), kT2Idesc, 1u);
          umma_f16(tm_im, d_dhi, umma_desc_plain(w_shi + ko, kT2BLbo), kT2Idesc, 1u);
          tc_commit(&empty[slot]);
#endif
          if (++slot == kT2Slots) { slot = 0; par ^= 1; }
        }
        // then the large ones; every k-step hands its ring slot back
        uint32_t sl2 = slot0;
#pragma unroll
        for (int s = 0; s < (SPB_T2_INTERLEAVE ? 0 : kT2KSteps); ++s) {
          const uint32_t a = sbase + kT2OffA + sl2 * kT2SlotBytes;
          const uint32_t ko = s * 2 * kT2BLbo;
          umma_f16(tm_re, umma_desc_plain(a, kT2ALbo), umma_desc_plain(w_chi + ko, kT2BLbo), kT2Idesc, 1u);
          umma_f16(tm_im, umma_desc_plain(a + 2 * kT2APart, kT2ALbo), umma_desc_plain(w_shi + ko, kT2BLbo), kT2Idesc, 1u);
          tc_commit(&empty[sl2]);
          if (++sl2 == kT2Slots) sl2 = 0;
        }
        tc_commit(&acc_full[acc]);
      }
    }
  } else {
    // ===== epilogue =====
    const int quad = warp & 3, half = warp >> 2;
    const int M = p.M;
    uint32_t it = 0;
    for (long long t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++it) {
      const TcGroup g = p.groups[t * 4 + quad];
      const uint32_t acc = it & 1;
      // pre-emphasis: the sample in front of this lane's frame (the utterance's last one for its first frame: np.roll)
      // and the frame's last sample, fetched (L2 hits: the converters read them a moment ago) before the wait
      float pre_a = 0.f, pre_e = 0.f;
      if (PRE && lane < g.nf) {
        const long long s0 = g.gstart + (long long)lane * kTcHop;
        const float xm = __ldg(p.samples + (lane == 0 ? g.prior : s0 - 1));
        const float xe = __ldg(p.samples + s0 + (2 * kTcHop - 1));
        pre_a = -p.pre_c * kT2InScale * xm;
        pre_e = p.pre_c * kT2InScale * xe;
      }
      t2_wait_park(&acc_full[acc], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * (2 * kT2N);
      const bool row_ok = lane < g.nf;
      // level bound of this lane's frame: sum x^2 <= 80 (max_f^2 + max_{f+1}^2), in the units of the accumulator
      const float l0 = lvl[(it & 3) * kTcRows + quad * 32 + lane];
      const float l1 = lvl[(it & 3) * kTcRows + min(quad * 32 + lane + 1, kTcRows - 1)];
      const float thr = row_ok ? p.redo.tau * (80.f * fmaf(l0, l0, l1 * l1)) : -1.f;
      const long long soff = g.gstart + (long long)lane * kTcHop;
      const long long row = g.row0 + lane;
      if (EPI == 2) {
        if (half == 0) t2_epilogue_static<0, PRE>(p, taddr, p.out + row * M, row_ok, thr, soff, row * M, pre_a, pre_e);
        else if (half == 1) t2_epilogue_static<1, PRE>(p, taddr, p.out + row * M, row_ok, thr, soff, row * M, pre_a, pre_e);
        else t2_epilogue_static<kT2Parts - 1, PRE>(p, taddr, p.out + row * M, row_ok, thr, soff, row * M, pre_a, pre_e);
      } else {
        if (half == 0) t2_epilogue_dyn<KM, EPI, 0>(p, taddr, p.out + row * M, row_ok, thr, soff, row * M);
        else t2_epilogue_dyn<KM, EPI, 1>(p, taddr, p.out + row * M + p.half[1].out_lo, row_ok, thr, soff,
                                         row * M + p.half[1].out_lo);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[acc]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaW) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}
