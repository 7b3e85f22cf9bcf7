// Training feature cubes straight from device-resident log-mel features.
//
// Reference (utils.py:325-397, applied per utterance by the DataLoader):
//   CMVN transform   : if DERIVATIVE: feature = extract_derivative_feature(feature)   (F, M, 3)
//                      if NORMALIZE : every (F, M) channel <- cmvn(channel, variance_normalization=True)
//   FeatureCube[3C]  : idx = np.random.randint(F - frames, size=crops);
//                      cube[ch, k, t, c] = feature[idx[k] + t, c, ch]                  float32
// Here: one stats kernel (per utterance, channel and column: mean and population std of the
// mean-subtracted values, float64 accumulators, two passes like numpy) and one gather kernel
// that derives Delta / Delta-Delta on the fly, normalises and writes the crops.  The random
// crop starts are drawn on the host (the reference uses the global numpy RNG) and passed in.
#include <algorithm>

#include "spb_internal.cuh"

namespace spb {

// The reference's delta operator acts along the FEATURE axis (processing.py:201-236 as written,
// SURVEY App. A.6): d[c] = (x[min(c+1, M-1)] + 2 x[min(c+2, M-1)]) / 10.
__device__ __forceinline__ float delta_at(const float *row, int c, int M) {
  const int c1 = min(c + 1, M - 1), c2 = min(c + 2, M - 1);
  return (row[c1] + 2.f * row[c2]) / 10.f;
}

__device__ __forceinline__ float channel_value(const float *row, int c, int ch, int M) {
  if (ch == 0) return row[c];
  if (ch == 1) return delta_at(row, c, M);
  const int c1 = min(c + 1, M - 1), c2 = min(c + 2, M - 1);
  return (delta_at(row, c1, M) + 2.f * delta_at(row, c2, M)) / 10.f;
}

constexpr int kStatSlices = 4;

// grid = utterances; block = (M * channels) x kStatSlices threads.  stats[u][ch][c] = {mean, std}
__global__ void cube_stats_kernel(const float *__restrict__ lmfe, const int64_t *__restrict__ frame_offsets,
                                  int M, int channels, double2 *__restrict__ stats) {
  extern __shared__ double s_red[];  // [kStatSlices][M * channels]
  const int u = blockIdx.x;
  const int cols = M * channels;
  const int j = threadIdx.x % cols, slice = threadIdx.x / cols;
  const int ch = j / M, c = j % M;
  const long long r0 = frame_offsets[u], F = frame_offsets[u + 1] - r0;
  const float *base = lmfe + r0 * M;
  double sum = 0.0;
  for (long long f = slice; f < F; f += kStatSlices) sum += (double)channel_value(base + f * M, c, ch, M);
  s_red[slice * cols + j] = sum;
  __syncthreads();
  double tot = 0.0;
  for (int s = 0; s < kStatSlices; ++s) tot += s_red[s * cols + j];
  const double mean = F > 0 ? tot / (double)F : 0.0;
  __syncthreads();
  // cmvn: std of the mean-subtracted FLOAT64 values (processing.py:257-266); np.std subtracts
  // their (round-off sized) mean again, which is reproduced by the two accumulators below
  double s1 = 0.0, s2 = 0.0;
  for (long long f = slice; f < F; f += kStatSlices) {
    const double d = (double)channel_value(base + f * M, c, ch, M) - mean;
    s1 += d;
    s2 += d * d;
  }
  s_red[slice * cols + j] = s1;
  __syncthreads();
  double t1 = 0.0;
  for (int s = 0; s < kStatSlices; ++s) t1 += s_red[s * cols + j];
  __syncthreads();
  s_red[slice * cols + j] = s2;
  __syncthreads();
  if (slice == 0) {
    double t2 = 0.0;
    for (int s = 0; s < kStatSlices; ++s) t2 += s_red[s * cols + j];
    double var = 0.0;
    if (F > 0) {
      const double m1 = t1 / (double)F;
      var = t2 / (double)F - m1 * m1;
      if (var < 0.0) var = 0.0;
    }
    stats[((long long)u * channels + ch) * M + c] = make_double2(mean, sqrt(var));
  }
}

// grid = (crops, utterances); out[u][ch][k][t][c]
__global__ void cube_gather_kernel(const float *__restrict__ lmfe, const int64_t *__restrict__ frame_offsets,
                                   const int32_t *__restrict__ crop_starts, const double2 *__restrict__ stats,
                                   int M, int channels, int num_crops, int crop_frames, float *__restrict__ out) {
  const int k = blockIdx.x, u = blockIdx.y;
  const long long r0 = frame_offsets[u] + crop_starts[(long long)u * num_crops + k];
  const int per_ch = crop_frames * M;
  const double eps = 9.313225746154785e-10;  // 2^-30 (processing.py:253)
  for (int i = threadIdx.x; i < per_ch * channels; i += blockDim.x) {
    const int ch = i / per_ch, rem = i % per_ch;
    const int t = rem / M, c = rem % M;
    float v = channel_value(lmfe + (r0 + t) * M, c, ch, M);
    if (stats) {
      const double2 ms = stats[((long long)u * channels + ch) * M + c];
      v = (float)(((double)v - ms.x) / (ms.y + eps));
    }
    out[((((long long)u * channels + ch) * num_crops + k) * crop_frames + t) * M + c] = v;
  }
}

int launch_feature_cubes(spb_context *ctx, const float *d_lmfe, const int64_t *d_frame_offsets, int64_t U, int M,
                         int channels, int normalize, const int32_t *d_crop_starts, int num_crops,
                         int crop_frames, double2 *d_stats, float *d_out) {
  if (U <= 0 || num_crops <= 0) return SPB_OK;
  const int cols = M * channels;
  if (cols * kStatSlices > 1024) return fail(ctx, SPB_ERR_UNSUPPORTED, "feature cubes: num_filters * channels > 256");
  if (U > 65535) return fail(ctx, SPB_ERR_UNSUPPORTED, "feature cubes: more than 65535 utterances per call");
  if (normalize) {
    cube_stats_kernel<<<(unsigned)U, cols * kStatSlices, (size_t)kStatSlices * cols * sizeof(double), ctx->stream>>>(
        d_lmfe, d_frame_offsets, M, channels, d_stats);
    SPB_LAUNCH_CHECK(ctx, "cube_stats_kernel");
  }
  dim3 grid((unsigned)num_crops, (unsigned)U);
  cube_gather_kernel<<<grid, 256, 0, ctx->stream>>>(d_lmfe, d_frame_offsets, d_crop_starts,
                                                     normalize ? d_stats : nullptr, M, channels, num_crops,
                                                     crop_frames, d_out);
  SPB_LAUNCH_CHECK(ctx, "cube_gather_kernel");
  return SPB_OK;
}

}  // namespace spb
