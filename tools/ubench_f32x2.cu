// Microbenchmark: issue throughput of scalar FADD/FFMA vs packed add.f32x2 / fma.f32x2 on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench tools/ubench_f32x2.cu && ./ubench
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float *out, int iters) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
  const float c = 1.0001f, d = 0.5f;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = a[i] + c;           // 16 FADD
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], c, d);   // 16 FFMA
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 16; i += 2) {                        // 8 FADD2 = 16 adds
        float2 v = make_float2(a[i], a[i + 1]);
        v = __fadd2_rn(v, make_float2(c, c));
        a[i] = v.x; a[i + 1] = v.y;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; i += 2) {                        // 8 FFMA2 = 16 fmas
        float2 v = make_float2(a[i], a[i + 1]);
        v = __ffma2_rn(v, make_float2(c, c), make_float2(d, d));
        a[i] = v.x; a[i + 1] = v.y;
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name) {
  float *out;
  cudaMalloc(&out, 148 * 8 * 256 * 4);
  const int iters = 20000;
  k<MODE><<<148 * 8, 256>>>(out, 10);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<148 * 8, 256>>>(out, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double ops = 148.0 * 8 * 256 * 16.0 * iters;  // scalar flop-instr equivalents
  printf("%-8s %.3f ms  %.1f G lane-ops/s  (%.1f lane-ops/clk/SM at 1.965 GHz)\n", name, ms,
         ops / ms / 1e6, ops / ms / 1e6 * 1e9 / 148 / 1.965e9);
  cudaFree(out);
}

int main() {
  run<0>("FADD");
  run<1>("FFMA");
  run<2>("FADD2");
  run<3>("FFMA2");
  run<0>("FADD");
  return 0;
}
