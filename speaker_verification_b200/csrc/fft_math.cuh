// Register-resident FFT building blocks, written as __host__ __device__ so the
// exact same code is exercised on the CPU by tests/host_fft_check.cu.
//
// Convention: forward DFT, X[k] = sum_n x[n] exp(-2*pi*i*n*k/N) (numpy rfft,
// speech_feature_extraction/speechpy/processing.py:158).
#pragma once

#include <cuda_runtime.h>

#ifdef __CUDACC__
#define SPB_HD __host__ __device__ __forceinline__
#else
#define SPB_HD inline
#endif

namespace spb {

SPB_HD float2 c_make(float x, float y) { float2 r; r.x = x; r.y = y; return r; }
SPB_HD float2 c_add(float2 a, float2 b) { return c_make(a.x + b.x, a.y + b.y); }
SPB_HD float2 c_sub(float2 a, float2 b) { return c_make(a.x - b.x, a.y - b.y); }
// a * w
SPB_HD float2 c_mul(float2 a, float2 w) {
  return c_make(fmaf(a.x, w.x, -a.y * w.y), fmaf(a.x, w.y, a.y * w.x));
}
// a * (-i)
SPB_HD float2 c_mul_mi(float2 a) { return c_make(a.y, -a.x); }

// 4-point forward DFT, NIN = how many leading inputs can be non-zero.
// (y0,y1,y2,y3) = DFT4(a,b,c,d) written back into a,b,c,d.
template <int NIN>
SPB_HD void dft4(float2 &a, float2 &b, float2 &c, float2 &d) {
  if (NIN >= 4) {
    float2 s0 = c_add(a, c), s1 = c_sub(a, c), s2 = c_add(b, d), s3 = c_sub(b, d);
    a = c_add(s0, s2);
    c = c_sub(s0, s2);
    b = c_make(s1.x + s3.y, s1.y - s3.x);  // s1 - i*s3
    d = c_make(s1.x - s3.y, s1.y + s3.x);  // s1 + i*s3
  } else if (NIN == 3) {  // d == 0
    float2 s0 = c_add(a, c), s1 = c_sub(a, c);
    float2 bb = b;
    a = c_add(s0, bb);
    c = c_sub(s0, bb);
    b = c_make(s1.x + bb.y, s1.y - bb.x);
    d = c_make(s1.x - bb.y, s1.y + bb.x);
  } else if (NIN == 2) {  // c == d == 0
    float2 aa = a, bb = b;
    a = c_add(aa, bb);
    c = c_sub(aa, bb);
    b = c_make(aa.x + bb.y, aa.y - bb.x);
    d = c_make(aa.x - bb.y, aa.y + bb.x);
  } else if (NIN == 1) {  // only a
    b = a; c = a; d = a;
  } else {
    a = b = c = d = c_make(0.f, 0.f);
  }
}

// 16-point forward DFT in registers, natural order in, natural order out.
// NZ: inputs v[NZ..15] are known to be zero (input pruning for zero-padded frames).
template <int NZ>
SPB_HD void fft16(float2 (&v)[16]) {
  constexpr float C1 = 0.92387953251128674f;  // cos(pi/8)
  constexpr float S1 = 0.38268343236508977f;  // sin(pi/8)
  constexpr float R2 = 0.70710678118654752f;  // sqrt(1/2)
  // stage a: j = r + 4*jh ; t[r][q] = sum_jh v[r+4jh] W4^(jh q)  -> stored at v[r + 4q]
#define SPB_NIN(r) ((NZ - (r) + 3) / 4 < 0 ? 0 : ((NZ - (r) + 3) / 4 > 4 ? 4 : (NZ - (r) + 3) / 4))
  dft4<SPB_NIN(0)>(v[0], v[4], v[8], v[12]);
  dft4<SPB_NIN(1)>(v[1], v[5], v[9], v[13]);
  dft4<SPB_NIN(2)>(v[2], v[6], v[10], v[14]);
  dft4<SPB_NIN(3)>(v[3], v[7], v[11], v[15]);
#undef SPB_NIN
  // twiddles W16^(r*q), W16 = exp(-i*pi/8)
  // r=1: q=1 -> W^1, q=2 -> W^2, q=3 -> W^3
  v[5] = c_mul(v[5], c_make(C1, -S1));
  v[9] = c_make((v[9].x + v[9].y) * R2, (v[9].y - v[9].x) * R2);        // W^2 = R2(1 - i)
  v[13] = c_mul(v[13], c_make(S1, -C1));
  // r=2: q=1 -> W^2, q=2 -> W^4 = -i, q=3 -> W^6 = R2(-1 - i)
  v[6] = c_make((v[6].x + v[6].y) * R2, (v[6].y - v[6].x) * R2);
  v[10] = c_mul_mi(v[10]);
  v[14] = c_make((v[14].y - v[14].x) * R2, -(v[14].x + v[14].y) * R2);
  // r=3: q=1 -> W^3, q=2 -> W^6, q=3 -> W^9 = (-C1, +S1)
  v[7] = c_mul(v[7], c_make(S1, -C1));
  v[11] = c_make((v[11].y - v[11].x) * R2, -(v[11].x + v[11].y) * R2);
  v[15] = c_mul(v[15], c_make(-C1, S1));
  // stage b: out[q + 4p] = sum_r t[r][q] W4^(r p); t[r][q] is at v[r + 4q]
  dft4<4>(v[0], v[1], v[2], v[3]);      // q = 0 -> k = 0,4,8,12 at v[0..3]
  dft4<4>(v[4], v[5], v[6], v[7]);      // q = 1 -> k = 1,5,9,13
  dft4<4>(v[8], v[9], v[10], v[11]);    // q = 2
  dft4<4>(v[12], v[13], v[14], v[15]);  // q = 3
  // v[p + 4q] holds out[q + 4p]  -> transpose the 4x4 index grid into natural order
  float2 t;
#define SPB_SWAP(i, j) t = v[i]; v[i] = v[j]; v[j] = t;
  SPB_SWAP(1, 4) SPB_SWAP(2, 8) SPB_SWAP(3, 12) SPB_SWAP(6, 9) SPB_SWAP(7, 13) SPB_SWAP(11, 14)
#undef SPB_SWAP
}

// ---- layout of the 16x16 exchange buffer (float2 units) used by the 512-point kernel.
// Element (k1, n2) lives at row k1 (16 float2 = 128 B), 16-byte chunk (n2>>1) ^ (k1&7).
// Writers (fixed k1, lanes n2 = 0..15) cover one 128 B row; readers (lane = k1) fetch
// logical chunk c with one 128-bit load and the xor spreads the 8 lanes of a
// quarter-warp over all 8 chunk positions -> no bank conflicts either way.
SPB_HD int exch_store_index(int k1, int n2) {
  return k1 * 16 + ((((n2 >> 1) ^ (k1 & 7)) << 1) | (n2 & 1));
}
SPB_HD int exch_load_chunk_index(int k1, int c) {  // float2 index of the pair (n2 = 2c, 2c+1)
  return k1 * 16 + ((c ^ (k1 & 7)) << 1);
}

}  // namespace spb
