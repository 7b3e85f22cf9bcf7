// CPU emulation of the 16-lane 512-point real-FFT scheme of features_fast512
// (speaker_verification_b200/csrc/feat_kernels.cu) using the very same
// __host__ __device__ building blocks from fft_math.cuh.  Checks against a naive
// float64 DFT.  Build: nvcc -I speaker_verification_b200/csrc tests/host/host_fft_check.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fft_math.cuh"

using namespace spb;

template <int NJ>
static double run_case(int L, unsigned seed) {
  std::vector<float> x(512, 0.f);
  srand(seed);
  for (int i = 0; i < L; ++i) x[i] = (float)rand() / RAND_MAX - 0.5f;
  // reference power spectrum (unscaled |X|^2), float64
  std::vector<double> pref(257);
  const double two_pi = 6.283185307179586;
  for (int k = 0; k <= 256; ++k) {
    double re = 0, im = 0;
    for (int n = 0; n < L; ++n) {
      re += x[n] * cos(two_pi * n * k / 512.0);
      im -= x[n] * sin(two_pi * n * k / 512.0);
    }
    pref[k] = re * re + im * im;
  }
  // tables as api.cu builds them
  float2 twA[16][16], twB[16][9];
  for (int l = 0; l < 16; ++l)
    for (int k1 = 0; k1 < 16; ++k1) {
      double a = -two_pi * ((l * k1) % 256) / 256.0;
      twA[l][k1] = c_make((float)cos(a), (float)sin(a));
    }
  for (int k1 = 0; k1 < 16; ++k1)
    for (int k2 = 0; k2 < 9; ++k2) {
      double a = -two_pi * (k1 + 16 * k2) / 512.0;
      twB[k1][k2] = c_make((float)cos(a), (float)sin(a));
    }
  float2 exch[256];
  float2 Z[16][16];  // per lane after step 2
  // step 1 on every lane
  for (int l = 0; l < 16; ++l) {
    float2 v[16];
    for (int j = 0; j < 16; ++j) {
      int m = 16 * j + l;
      float re = (j < NJ && 2 * m < L) ? x[2 * m] : 0.f;
      float im = (j < NJ && 2 * m + 1 < L) ? x[2 * m + 1] : 0.f;
      v[j] = c_make(re, im);
    }
    fft16<NJ>(v);
    for (int k1 = 1; k1 < 16; ++k1) v[k1] = c_mul(v[k1], twA[l][k1]);
    for (int k1 = 0; k1 < 16; ++k1) exch[exch_store_index(k1, l)] = v[k1];
  }
  // step 2
  for (int l = 0; l < 16; ++l) {
    float2 v[16];
    for (int c = 0; c < 8; ++c) {
      v[2 * c] = exch[exch_load_chunk_index(l, c)];
      v[2 * c + 1] = exch[exch_load_chunk_index(l, c) + 1];
    }
    fft16<16>(v);
    for (int k2 = 0; k2 < 16; ++k2) Z[l][k2] = v[k2];
  }
  // untangle
  std::vector<double> p(257, -1.0);
  double sumsq = 0;
  for (int i = 0; i < L; ++i) sumsq += (double)x[i] * x[i];
  for (int l = 0; l < 16; ++l) {
    int partner = (16 - l) & 15;
    for (int k2 = 0; k2 < 8; ++k2) {
      int t = 7 - k2;
      float2 zp = (partner == 0) ? Z[partner][(9 + t) & 15] : Z[partner][8 + t];
      float2 zk = Z[l][k2];
      float2 A = c_make(zk.x + zp.x, zk.y - zp.y);
      float2 B = c_make(zk.y + zp.y, zp.x - zk.x);
      float2 WB = c_mul(B, twB[l][k2]);
      float xr = A.x + WB.x, xi = A.y + WB.y;
      float yr = A.x - WB.x, yi = A.y - WB.y;
      p[l + 16 * k2] = 0.25f * (xr * xr + xi * xi);
      p[256 - (l + 16 * k2)] = 0.25f * (yr * yr + yi * yi);
    }
    if (l == 0) p[128] = Z[0][8].x * Z[0][8].x + Z[0][8].y * Z[0][8].y;
  }
  double maxref = 0, maxerr = 0;
  for (int k = 0; k <= 256; ++k) {
    if (p[k] < 0) { printf("bin %d never written\n", k); return 1e9; }
    maxref = fmax(maxref, pref[k]);
    maxerr = fmax(maxerr, fabs(p[k] - pref[k]));
  }
  // Parseval energy
  double e_ref = 0;
  for (int k = 0; k <= 256; ++k) e_ref += pref[k] / 512.0;
  double x0 = Z[0][0].x + Z[0][0].y, x256 = Z[0][0].x - Z[0][0].y;
  double e = (512.0 * sumsq + x0 * x0 + x256 * x256) / (2.0 * 512.0);
  double rel_e = fabs(e - e_ref) / e_ref;
  printf("L=%d NJ=%d  max|dP|/max P = %.3e   energy rel err = %.3e\n", L, NJ, maxerr / maxref, rel_e);
  return fmax(maxerr / maxref, rel_e);
}

int main() {
  // exchange layout must be a bijection
  int seen[256] = {0};
  for (int k1 = 0; k1 < 16; ++k1)
    for (int n2 = 0; n2 < 16; ++n2) seen[exch_store_index(k1, n2)]++;
  for (int i = 0; i < 256; ++i)
    if (seen[i] != 1) { printf("exchange layout is not a bijection\n"); return 1; }
  double worst = 0;
  worst = fmax(worst, run_case<10>(320, 1));
  worst = fmax(worst, run_case<16>(320, 2));
  worst = fmax(worst, run_case<16>(400, 3));
  worst = fmax(worst, run_case<16>(512, 4));
  worst = fmax(worst, run_case<16>(37, 5));
  worst = fmax(worst, run_case<16>(1, 6));
  if (worst > 5e-6) { printf("FAIL worst=%.3e\n", worst); return 1; }
  printf("OK worst=%.3e\n", worst);
  return 0;
}
