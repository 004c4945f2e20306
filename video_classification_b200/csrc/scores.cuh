// scores.cuh -- what happens to the logits after the network in the reference's evaluation drivers (SURVEY.md 8(f) rows
// f4 and f2), kept on the device:
//   video_scores_kernel : Trainer.run_eval, reference train.py:337-360 -- softmax over classes per clip
//                         (np.exp(x) / np.sum(np.exp(x))), mean over the clips sampled from one video, arg-max;
//                         and SparseFusionTrainer.test, reference train_sparse.py:214-232 (same, without the softmax).
//   sparse_fuse_kernel  : SparseModel.forward, reference train_sparse.py:89-105 -- 249 nn.Linear(P, 1) looped in
//                         Python and concatenated: y[n, c] = sum_p w[c][p] * x[n, p, c] + b[c].
// Tiny, latency-bound kernels (a few hundred KB); the point is that no per-batch .cpu().numpy() sync or pickle round
// trip remains, not bandwidth.
#pragma once
#include "common.cuh"

namespace sfb {

constexpr int kScoreThreads = 256;

__device__ __forceinline__ float block_reduce_max(float v, float* red) {
  for (int o = 16; o >= 1; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = red[0];
  for (int w = 1; w < kScoreThreads / 32; ++w) r = fmaxf(r, red[w]);
  return r;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
  for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  for (int w = 0; w < kScoreThreads / 32; ++w) r += red[w];
  return r;
}

// one block per video; seg[v] .. seg[v+1] are its clips (rows of `scores`).  C <= kScoreThreads * kScoreMaxPer.
constexpr int kScoreMaxPer = 8;
template <bool SOFTMAX>
__global__ void __launch_bounds__(kScoreThreads)
video_scores_kernel(const float* __restrict__ scores, int C, const int* __restrict__ seg, float* __restrict__ clip_probs,
                    float* __restrict__ video_mean, int* __restrict__ top1) {
  __shared__ float red[kScoreThreads / 32];
  __shared__ int redi[kScoreThreads / 32];
  const int v = blockIdx.x, tid = threadIdx.x;
  const int c0 = seg[v], c1 = seg[v + 1];
  float acc[kScoreMaxPer];
#pragma unroll
  for (int j = 0; j < kScoreMaxPer; ++j) acc[j] = 0.f;
  for (int clip = c0; clip < c1; ++clip) {
    const float* row = scores + (size_t)clip * C;
    float x[kScoreMaxPer];
#pragma unroll
    for (int j = 0; j < kScoreMaxPer; ++j) {
      const int c = tid + j * kScoreThreads;
      x[j] = c < C ? row[c] : -INFINITY;
    }
    if (SOFTMAX) {
      float m = -INFINITY;
#pragma unroll
      for (int j = 0; j < kScoreMaxPer; ++j) m = fmaxf(m, x[j]);
      m = block_reduce_max(m, red);
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < kScoreMaxPer; ++j) {
        x[j] = tid + j * kScoreThreads < C ? expf(x[j] - m) : 0.f;
        s += x[j];
      }
      s = block_reduce_sum(s, red);
      const float inv = 1.0f / s;
#pragma unroll
      for (int j = 0; j < kScoreMaxPer; ++j) x[j] *= inv;
    }
#pragma unroll
    for (int j = 0; j < kScoreMaxPer; ++j) {
      const int c = tid + j * kScoreThreads;
      if (c < C) {
        if (clip_probs) clip_probs[(size_t)clip * C + c] = x[j];
        acc[j] += x[j];
      }
    }
  }
  // mean over the clips, arg-max with np.argmax's tie rule (lowest index)
  const float invn = c1 > c0 ? 1.0f / (float)(c1 - c0) : 0.f;
  float best = -INFINITY;
  int besti = 0x7fffffff;
#pragma unroll
  for (int j = 0; j < kScoreMaxPer; ++j) {
    const int c = tid + j * kScoreThreads;
    if (c < C) {
      const float mval = acc[j] * invn;
      if (video_mean) video_mean[(size_t)v * C + c] = mval;
      if (mval > best) { best = mval; besti = c; }   // ascending c: a later equal value does not replace
    }
  }
  for (int o = 16; o >= 1; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
  }
  // the last block_reduce_sum ends with every warp READING red[]: a fast warp must not overwrite it with its arg-max
  // candidate before the slow ones are through
  __syncthreads();
  if ((tid & 31) == 0) { red[tid >> 5] = best; redi[tid >> 5] = besti; }
  __syncthreads();
  if (tid == 0 && top1) {
    for (int w = 1; w < kScoreThreads / 32; ++w)
      if (red[w] > best || (red[w] == best && redi[w] < besti)) { best = red[w]; besti = redi[w]; }
    top1[v] = c1 > c0 ? besti : -1;
  }
}

// x: [N][P][C] part scores, w: [C][P], b: [C] -> y: [N][C]
__global__ void sparse_fuse_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                   int N, int P, int C, float* __restrict__ y) {
  const long long total = (long long)N * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long n = i / C;
    float s = 0.f;
    for (int p = 0; p < P; ++p) s = fmaf(w[(size_t)c * P + p], x[((size_t)n * P + p) * C + c], s);
    y[i] = s + b[c];
  }
}

// backward of sparse_fuse (the reference trains SparseModel with autograd, train_sparse.py:160-190):
// dw[c][p] = sum_n dy[n][c] * x[n][p][c], db[c] = sum_n dy[n][c], dx[n][p][c] = dy[n][c] * w[c][p] (dx may be null).
// One warp per class c: lanes stride over n, shuffle-reduce per part.  P <= 32.
__global__ void sparse_fuse_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ w,
                                       int N, int P, int C, float* __restrict__ dw, float* __restrict__ db,
                                       float* __restrict__ dx) {
  const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  float sb = 0.f;
  for (int n = lane; n < N; n += 32) sb += dy[(size_t)n * C + c];
  for (int o = 16; o >= 1; o >>= 1) sb += __shfl_xor_sync(0xffffffffu, sb, o);
  if (lane == 0) db[c] = sb;
  for (int p = 0; p < P; ++p) {
    const float wcp = w[(size_t)c * P + p];
    float s = 0.f;
    for (int n = lane; n < N; n += 32) {
      const float g = dy[(size_t)n * C + c];
      s = fmaf(g, x[((size_t)n * P + p) * C + c], s);
      if (dx) dx[((size_t)n * P + p) * C + c] = g * wcp;
    }
    for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) dw[(size_t)c * P + p] = s;
  }
}

}  // namespace sfb
