// Training feature cubes straight from device-resident log-mel features.
//
// Reference (utils.py:325-397, applied per utterance by the DataLoader):
//   CMVN transform   : if DERIVATIVE: feature = extract_derivative_feature(feature)   (F, M, 3)
//                      if NORMALIZE : every (F, M) channel <- cmvn(channel, variance_normalization=True)
//   FeatureCube[3C]  : idx = np.random.randint(F - frames, size=crops);
//                      cube[ch, k, t, c] = feature[idx[k] + t, c, ch]                  float32
// Here: one stats kernel (per utterance, channel and column: mean and population std of the
// mean-subtracted values, float64 accumulators) and one gather kernel
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

// grid = utterances; block = kStatSlices x (M * channels) threads.  stats[u][ch][c] = {mean, 1/(std+eps)}.
// One pass: every slice stages its row in shared memory (one coalesced 4*M-byte read per row),
// the M * channels threads of the slice derive their channel value from it and accumulate
// sum(v - p) and sum((v - p)^2) in float64 around the pivot p = value in the utterance's first
// row (keeps the single-pass variance free of cancellation).  mean = p + S1/F,
// var = S2/F - (S1/F)^2 == the population variance of the mean-subtracted column that
// cmvn (processing.py:257-266) takes.
__global__ void cube_stats_kernel(const float *__restrict__ lmfe, const int64_t *__restrict__ frame_offsets,
                                  int M, int channels, double2 *__restrict__ stats) {
  extern __shared__ double s_red[];                         // [2][kStatSlices][cols]
  const int cols = M * channels;
  float *s_row = reinterpret_cast<float *>(s_red + 2 * kStatSlices * cols);  // [kStatSlices + 1][M]
  const int u = blockIdx.x;
  const int j = threadIdx.x % cols, slice = threadIdx.x / cols;
  const int ch = j / M, c = j % M;
  const long long r0 = frame_offsets[u], F = frame_offsets[u + 1] - r0;
  const float *base = lmfe + r0 * M;
  float *row = s_row + slice * M, *row0 = s_row + kStatSlices * M;
  if (threadIdx.x < M) row0[threadIdx.x] = F > 0 ? base[threadIdx.x] : 0.f;
  __syncthreads();
  const double pivot = (double)channel_value(row0, c, ch, M);
  double s1 = 0.0, s2 = 0.0;
  for (long long f0 = 0; f0 < F; f0 += kStatSlices) {
    const long long f = f0 + slice;
    if (j < M && f < F) row[j] = base[f * M + j];
    __syncthreads();
    if (f < F) {
      const double d = (double)channel_value(row, c, ch, M) - pivot;
      s1 += d;
      s2 += d * d;
    }
    __syncthreads();
  }
  s_red[slice * cols + j] = s1;
  s_red[(kStatSlices + slice) * cols + j] = s2;
  __syncthreads();
  if (slice == 0) {
    double t1 = 0.0, t2 = 0.0;
    for (int s = 0; s < kStatSlices; ++s) {
      t1 += s_red[s * cols + j];
      t2 += s_red[(kStatSlices + s) * cols + j];
    }
    double mean = 0.0, var = 0.0;
    if (F > 0) {
      const double m1 = t1 / (double)F;
      mean = pivot + m1;
      var = t2 / (double)F - m1 * m1;
      if (var < 0.0) var = 0.0;
    }
    // {mean, 1 / (std + 2^-30)} (processing.py:253, 266)
    stats[((long long)u * channels + ch) * M + c] = make_double2(mean, 1.0 / (sqrt(var) + 9.313225746154785e-10));
  }
}

// grid = (crops, utterances); out[u][ch][k][t][c].  The crop's rows are staged in shared memory
// once (coalesced), every channel is derived from the staged rows and stored coalesced.
__global__ void cube_gather_kernel(const float *__restrict__ lmfe, const int64_t *__restrict__ frame_offsets,
                                   const int32_t *__restrict__ crop_starts, const double2 *__restrict__ stats,
                                   long long U, int M, int channels, int num_crops, int crop_frames,
                                   float *__restrict__ out) {
  extern __shared__ float s_crop[];  // [crop_frames][M]
  const int k = blockIdx.x;
  const int per_ch = crop_frames * M;
  for (long long u = blockIdx.y; u < U; u += gridDim.y) {  // gridDim.y <= 65535
    const long long r0 = frame_offsets[u] + crop_starts[u * num_crops + k];
    const float *src = lmfe + r0 * M;
    __syncthreads();  // previous utterance's crop fully consumed
    for (int i = threadIdx.x; i < per_ch; i += blockDim.x) s_crop[i] = src[i];
    __syncthreads();
    for (int i = threadIdx.x; i < per_ch; i += blockDim.x) {
      const int t = i / M, c = i - t * M;
      const float *row = s_crop + t * M;
      for (int ch = 0; ch < channels; ++ch) {
        float v = channel_value(row, c, ch, M);
        if (stats) {
          const double2 ms = stats[(u * channels + ch) * M + c];
          v = (v - (float)ms.x) * (float)ms.y;  // float32 like the cube; |error| ~1e-6 of an O(1) value
        }
        out[((u * channels + ch) * num_crops + k) * per_ch + i] = v;
      }
    }
  }
}

int launch_feature_cubes(spb_context *ctx, const float *d_lmfe, const int64_t *d_frame_offsets, int64_t U, int M,
                         int channels, int normalize, const int32_t *d_crop_starts, int num_crops,
                         int crop_frames, double2 *d_stats, float *d_out) {
  if (U <= 0 || num_crops <= 0) return SPB_OK;
  const int cols = M * channels;
  if (cols * kStatSlices > 1024) return fail(ctx, SPB_ERR_UNSUPPORTED, "feature cubes: num_filters * channels > 256");
  if ((size_t)crop_frames * M * sizeof(float) > 48 * 1024)
    return fail(ctx, SPB_ERR_UNSUPPORTED, "feature cubes: crop_frames * num_filters > 12288");
  if (normalize) {
    cube_stats_kernel<<<(unsigned)U, cols * kStatSlices,
                        (size_t)2 * kStatSlices * cols * sizeof(double) + (size_t)(kStatSlices + 1) * M * sizeof(float),
                        ctx->stream>>>(
        d_lmfe, d_frame_offsets, M, channels, d_stats);
    SPB_LAUNCH_CHECK(ctx, "cube_stats_kernel");
  }
  if (num_crops > 65535 * 32) return fail(ctx, SPB_ERR_UNSUPPORTED, "feature cubes: too many crops per utterance");
  dim3 grid((unsigned)num_crops, (unsigned)std::min<int64_t>(U, 65535));
  cube_gather_kernel<<<grid, 256, (size_t)crop_frames * M * sizeof(float), ctx->stream>>>(d_lmfe, d_frame_offsets, d_crop_starts,
                                                     normalize ? d_stats : nullptr, (long long)U, M, channels, num_crops,
                                                     crop_frames, d_out);
  SPB_LAUNCH_CHECK(ctx, "cube_gather_kernel");
  return SPB_OK;
}

}  // namespace spb
