// adam.cuh -- the optimizer step of the training loop as ONE launch over a device-resident job table
// (see include/sfb200.h, sfb_adam_plan / sfb_adam_step; replaces optim.Adam(...).step(), ref train.py:182,231).
#pragma once
#include "common.cuh"

namespace sfb {

// One job = one parameter tensor: 4 fp32 streams of n elements (param, grad, exp_avg, exp_avg_sq).  HBM-bound:
// 16 B read + 12 B written per element, no reuse.  A block owns kAdamChunk consecutive elements of one job and finds
// the job by binary search on the jobs' first-block index (the table is a few KB and stays in L2).
constexpr int kAdamThreads = 256;
constexpr int kAdamVecPerThread = 4;                                  // 4 x float4 in flight per stream per thread
constexpr int kAdamChunk = kAdamThreads * 4 * kAdamVecPerThread;      // 4096 elements per block

struct AdamJobDev {
  float* p;
  const float* g;
  float* m;
  float* v;
  long long n;
  int block0, vec;  // vec: all four pointers 16-byte aligned
};

struct AdamHyper {
  float one_minus_beta1, beta2, one_minus_beta2, eps, weight_decay, neg_step_size, bc2_sqrt, grad_scale;
};

// The update in torch's operation order and roundings (oracle/adam_oracle.py): three fused multiply-adds, every
// other operation rounded on its own (the _rn intrinsics keep ptxas from contracting them differently).
__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, const AdamHyper& h) {
  if (h.grad_scale != 1.0f) g = __fmul_rn(g, h.grad_scale);
  if (h.weight_decay != 0.0f) g = __fmaf_rn(p, h.weight_decay, g);
  m = __fmaf_rn(__fsub_rn(g, m), h.one_minus_beta1, m);
  v = __fmaf_rn(__fmul_rn(h.one_minus_beta2, g), g, __fmul_rn(v, h.beta2));
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), h.bc2_sqrt), h.eps);
  p = __fadd_rn(p, __fdiv_rn(__fmul_rn(h.neg_step_size, m), denom));
}

__global__ void __launch_bounds__(kAdamThreads) adam_step_kernel(const AdamJobDev* __restrict__ jobs, int n_jobs,
                                                                 AdamHyper h) {
  int lo = 0, hi = n_jobs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (jobs[mid].block0 <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const AdamJobDev j = jobs[lo];
  const long long base = (long long)((int)blockIdx.x - j.block0) * kAdamChunk;
  const long long left = j.n - base;
  const int cnt = left < kAdamChunk ? (int)left : kAdamChunk;
  float* __restrict__ p = j.p + base;
  const float* __restrict__ g = j.g + base;
  float* __restrict__ m = j.m + base;
  float* __restrict__ v = j.v + base;
  if (j.vec && cnt == kAdamChunk) {
    // full chunk: all loads of the four streams issued before the first dependent instruction
    float4 P[kAdamVecPerThread], G[kAdamVecPerThread], M[kAdamVecPerThread], V[kAdamVecPerThread];
#pragma unroll
    for (int i = 0; i < kAdamVecPerThread; ++i) {
      const int e = (i * kAdamThreads + threadIdx.x) * 4;
      G[i] = __ldcs(reinterpret_cast<const float4*>(g + e));  // the gradient is dead after this step
      P[i] = __ldcg(reinterpret_cast<const float4*>(p + e));  // ld.global: the table pointers are device memory
      M[i] = __ldcg(reinterpret_cast<const float4*>(m + e));
      V[i] = __ldcg(reinterpret_cast<const float4*>(v + e));
    }
#pragma unroll
    for (int i = 0; i < kAdamVecPerThread; ++i) {
      const int e = (i * kAdamThreads + threadIdx.x) * 4;
      adam_update(P[i].x, G[i].x, M[i].x, V[i].x, h);
      adam_update(P[i].y, G[i].y, M[i].y, V[i].y, h);
      adam_update(P[i].z, G[i].z, M[i].z, V[i].z, h);
      adam_update(P[i].w, G[i].w, M[i].w, V[i].w, h);
      __stcg(reinterpret_cast<float4*>(p + e), P[i]);
      __stcg(reinterpret_cast<float4*>(m + e), M[i]);
      __stcg(reinterpret_cast<float4*>(v + e), V[i]);
    }
  } else {
    for (int e = threadIdx.x; e < cnt; e += kAdamThreads) {
      float pp = __ldcg(p + e), mm = __ldcg(m + e), vv = __ldcg(v + e);
      adam_update(pp, __ldcs(g + e), mm, vv, h);
      __stcg(p + e, pp);
      __stcg(m + e, mm);
      __stcg(v + e, vv);
    }
  }
}

}  // namespace sfb
