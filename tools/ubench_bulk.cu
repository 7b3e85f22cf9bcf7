// Micro-benchmark: per-SM throughput of global -> shared staging on B200, one CTA per SM.
//   mode 0: cp.async.bulk (1-D TMA) of `chunk` bytes, `depth` copies in flight (separate smem slots / mbarriers)
//   mode 1: cp.async (LDGSTS) 16 B per thread, all threads, `depth` commit groups in flight
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_bulk ubench_bulk.cu && ./ubench_bulk
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t su32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(512, 1) k(const char *src, size_t per_cta, int chunk, int depth, int mode, int iters, unsigned long long *sink) {
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ __align__(8) uint64_t bar[8];
  const char *base = src + (size_t)blockIdx.x * per_cta;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(su32(&bar[i])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int nchunks = (int)(per_cta / chunk);
  unsigned long long acc = 0;
  for (int it = 0; it < iters; ++it) {
    if (mode == 0) {
      if (threadIdx.x == 0) {
        uint32_t ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int c = 0; c < nchunks + depth; ++c) {
          const int sl = c % depth;
          if (c >= depth) {
            uint32_t ok = 0;
            while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(su32(&bar[sl])), "r"(ph[sl]) : "memory");
            ph[sl] ^= 1;
          }
          if (c < nchunks) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(su32(&bar[sl])), "r"(chunk) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(su32(sm + (size_t)sl * chunk)), "l"(base + (size_t)c * chunk), "r"(chunk), "r"(su32(&bar[sl])) : "memory");
          }
        }
      }
    } else {
      for (int c = 0; c < nchunks + depth; ++c) {
        const int sl = c % depth;
        if (c < nchunks) {
          for (int off = threadIdx.x * 16; off < chunk; off += blockDim.x * 16)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(su32(sm + (size_t)sl * chunk + off)), "l"(base + (size_t)c * chunk + off) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (depth == 1) asm volatile("cp.async.wait_group 0;" ::: "memory");
        else if (depth == 2) asm volatile("cp.async.wait_group 1;" ::: "memory");
        else if (depth == 4) asm volatile("cp.async.wait_group 3;" ::: "memory");
        else asm volatile("cp.async.wait_group 7;" ::: "memory");
      }
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    acc += sm[threadIdx.x];
  }
  if (acc == 12345) *sink = acc;
}
int main() {
  int sms = 148;
  const size_t per_cta = 48u << 20;  // 48 MiB per CTA: 7 GB in total, far beyond L2
  char *d; unsigned long long *sink;
  cudaMalloc(&d, per_cta * sms); cudaMemset(d, 1, per_cta * sms); cudaMalloc(&sink, 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode)
    for (int chunk : {4096, 16384, 28800, 65536})
      for (int depth : {1, 2, 4, 8}) {
        if ((size_t)chunk * depth > 200 * 1024) continue;
        k<<<sms, 512, 200 * 1024>>>(d, per_cta, chunk, depth, mode, 1, sink);
        cudaEventRecord(e0);
        k<<<sms, 512, 200 * 1024>>>(d, per_cta, chunk, depth, mode, 1, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double bytes = (double)(per_cta / chunk) * chunk * sms;
        printf("mode %d (%s) chunk %6d depth %d: %7.1f GB/s total, %6.2f GB/s per SM  (%s)\n", mode, mode ? "LDGSTS" : "bulk  ", chunk, depth,
               bytes / ms / 1e6, bytes / ms / 1e6 / sms, cudaGetErrorString(cudaGetLastError()));
      }
  return 0;
}
