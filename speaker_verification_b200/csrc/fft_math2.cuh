// Two-frame packed variant of the register FFT blocks in fft_math.cuh (device only).
//
// Blackwell (sm_100) has packed fp32 instructions (FADD2 / FMUL2 / FFMA2: add.f32x2,
// mul.f32x2, fma.rn.f32x2) that perform two fp32 operations per issue slot at the same lane
// throughput as the scalar forms (measured with tools/ubench_f32x2.cu: 121-127 lane-ops per
// clock per SM either way).  The FFT kernel is issue-bound, so each lane transforms TWO frames
// at once: every scalar of fft_math.cuh becomes a float2 holding (frame A, frame B), every
// FADD/FMUL/FFMA becomes its packed form; twiddles stay scalar (the packed instructions take
// a broadcast scalar or an immediate as an operand).  Arithmetic per frame is identical to the
// scalar code, operation for operation.
#pragma once

#include <cuda_runtime.h>

namespace spb {

struct C2 {      // one complex value of each of two frames
  float2 re;     // (re_A, re_B)
  float2 im;     // (im_A, im_B)
};

__device__ __forceinline__ float2 p_make(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 p_bc(float s) { return make_float2(s, s); }
__device__ __forceinline__ float2 p_add(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 p_sub(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }
__device__ __forceinline__ float2 p_mul(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 p_muls(float2 a, float s) { return __fmul2_rn(a, p_bc(s)); }
__device__ __forceinline__ float2 p_fma(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 p_fmas(float2 a, float s, float2 c) { return __ffma2_rn(a, p_bc(s), c); }
__device__ __forceinline__ float2 p_neg(float2 a) { return make_float2(-a.x, -a.y); }

__device__ __forceinline__ C2 c2_make(float2 re, float2 im) { C2 r; r.re = re; r.im = im; return r; }
__device__ __forceinline__ C2 c2_zero() { return c2_make(p_bc(0.f), p_bc(0.f)); }
__device__ __forceinline__ C2 c2_add(C2 a, C2 b) { return c2_make(p_add(a.re, b.re), p_add(a.im, b.im)); }
__device__ __forceinline__ C2 c2_sub(C2 a, C2 b) { return c2_make(p_sub(a.re, b.re), p_sub(a.im, b.im)); }
// a * w, w a scalar twiddle shared by both frames: same operation order as c_mul
// (fmaf(a.x, w.x, -a.y * w.y), fmaf(a.x, w.y, a.y * w.x))
__device__ __forceinline__ C2 c2_mul(C2 a, float2 w) {
  return c2_make(p_fmas(a.re, w.x, p_muls(a.im, -w.y)), p_fmas(a.re, w.y, p_muls(a.im, w.x)));
}
// a * (-i)
__device__ __forceinline__ C2 c2_mul_mi(C2 a) { return c2_make(a.im, p_neg(a.re)); }

template <int NIN>
__device__ __forceinline__ void dft4_2(C2 &a, C2 &b, C2 &c, C2 &d) {
  if (NIN >= 4) {
    C2 s0 = c2_add(a, c), s1 = c2_sub(a, c), s2 = c2_add(b, d), s3 = c2_sub(b, d);
    a = c2_add(s0, s2);
    c = c2_sub(s0, s2);
    b = c2_make(p_add(s1.re, s3.im), p_sub(s1.im, s3.re));  // s1 - i*s3
    d = c2_make(p_sub(s1.re, s3.im), p_add(s1.im, s3.re));  // s1 + i*s3
  } else if (NIN == 3) {  // d == 0
    C2 s0 = c2_add(a, c), s1 = c2_sub(a, c);
    C2 bb = b;
    a = c2_add(s0, bb);
    c = c2_sub(s0, bb);
    b = c2_make(p_add(s1.re, bb.im), p_sub(s1.im, bb.re));
    d = c2_make(p_sub(s1.re, bb.im), p_add(s1.im, bb.re));
  } else if (NIN == 2) {  // c == d == 0
    C2 aa = a, bb = b;
    a = c2_add(aa, bb);
    c = c2_sub(aa, bb);
    b = c2_make(p_add(aa.re, bb.im), p_sub(aa.im, bb.re));
    d = c2_make(p_sub(aa.re, bb.im), p_add(aa.im, bb.re));
  } else if (NIN == 1) {
    b = a; c = a; d = a;
  } else {
    a = b = c = d = c2_zero();
  }
}

// 16-point forward DFT of two frames in registers (mirror of fft16<NZ> in fft_math.cuh).
template <int NZ>
__device__ __forceinline__ void fft16_2(C2 (&v)[16]) {
  constexpr float C1 = 0.92387953251128674f;  // cos(pi/8)
  constexpr float S1 = 0.38268343236508977f;  // sin(pi/8)
  constexpr float R2 = 0.70710678118654752f;  // sqrt(1/2)
#define SPB_NIN(r) ((NZ - (r) + 3) / 4 < 0 ? 0 : ((NZ - (r) + 3) / 4 > 4 ? 4 : (NZ - (r) + 3) / 4))
  dft4_2<SPB_NIN(0)>(v[0], v[4], v[8], v[12]);
  dft4_2<SPB_NIN(1)>(v[1], v[5], v[9], v[13]);
  dft4_2<SPB_NIN(2)>(v[2], v[6], v[10], v[14]);
  dft4_2<SPB_NIN(3)>(v[3], v[7], v[11], v[15]);
#undef SPB_NIN
  // twiddles W16^(r*q)
  v[5] = c2_mul(v[5], make_float2(C1, -S1));
  v[9] = c2_make(p_muls(p_add(v[9].re, v[9].im), R2), p_muls(p_sub(v[9].im, v[9].re), R2));       // W^2 = R2(1 - i)
  v[13] = c2_mul(v[13], make_float2(S1, -C1));
  v[6] = c2_make(p_muls(p_add(v[6].re, v[6].im), R2), p_muls(p_sub(v[6].im, v[6].re), R2));
  v[10] = c2_mul_mi(v[10]);
  v[14] = c2_make(p_muls(p_sub(v[14].im, v[14].re), R2), p_muls(p_add(v[14].re, v[14].im), -R2));  // W^6 = R2(-1 - i)
  v[7] = c2_mul(v[7], make_float2(S1, -C1));
  v[11] = c2_make(p_muls(p_sub(v[11].im, v[11].re), R2), p_muls(p_add(v[11].re, v[11].im), -R2));
  v[15] = c2_mul(v[15], make_float2(-C1, S1));
  dft4_2<4>(v[0], v[1], v[2], v[3]);
  dft4_2<4>(v[4], v[5], v[6], v[7]);
  dft4_2<4>(v[8], v[9], v[10], v[11]);
  dft4_2<4>(v[12], v[13], v[14], v[15]);
  C2 t;
#define SPB_SWAP(i, j) t = v[i]; v[i] = v[j]; v[j] = t;
  SPB_SWAP(1, 4) SPB_SWAP(2, 8) SPB_SWAP(3, 12) SPB_SWAP(6, 9) SPB_SWAP(7, 13) SPB_SWAP(11, 14)
#undef SPB_SWAP
}

}  // namespace spb
