// conv_simt.cuh -- CUDA-core implementations of the conv GEMM / wgrad contracts of sfb200.h.
// Same operands, same prologue/epilogue semantics as conv_tc.cuh; used as the on-device cross-check
// of the tcgen05 kernels (SFB_IMPL_SIMT) -- never a CPU path.
#pragma once
#include "conv_tc.cuh"

namespace sfb {

struct ConvSimtParams {
  ConvTcParams c;
  const __nv_bfloat16* wp;
  int k_pad;
};

// one thread per (m, co)
__global__ void conv_gemm_simt_kernel(const ConvSimtParams q) {
  const ConvTcParams& p = q.c;
  const SfbConvGeom& g = p.g;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)p.M * g.Co) return;
  const int co = (int)(idx % g.Co);
  const int m = (int)(idx / g.Co);
  const RowCoord rc = make_row(g, m, p.M);
  const bool unit_stride = (g.sT == 1 && g.sH == 1 && g.sW == 1);
  const __nv_bfloat16* wrow = q.wp + (size_t)co * q.k_pad;
  float acc = 0.f;
  int tap = 0;
  for (int kt = 0; kt < g.kT; ++kt)
    for (int kh = 0; kh < g.kH; ++kh)
      for (int kw = 0; kw < g.kW; ++kw, ++tap) {
        const int pix = gather_pixel(g, rc, kt, kh, kw, unit_stride);
        if (pix < 0) continue;
        const __nv_bfloat16* src = p.G + (size_t)pix * g.g_ld + g.g_coff;
        const __nv_bfloat16* w = wrow + tap * g.Cg;
        for (int c = 0; c < g.Cg; ++c) {
          float x = __bfloat162float(src[c]);
          if (p.pro_scale) {
            x = fmaf(x, p.pro_scale[c], p.pro_shift[c]);
            if (p.pro_relu) x = fmaxf(x, 0.f);
            x = __bfloat162float(__float2bfloat16_rn(x));  // the TC path stages the operand as bf16
          }
          acc = fmaf(x, __bfloat162float(w[c]), acc);
        }
      }
  float y = acc;
  if (p.epi_scale) y = fmaf(y, p.epi_scale[co], p.epi_shift[co]);
  if (p.res) y += __bfloat162float(p.res[(size_t)m * p.res_ld + p.res_coff + co]);
  if (p.epi_relu) y = fmaxf(y, 0.f);
  p.out[(size_t)m * g.o_ld + g.o_coff + co] = __float2bfloat16_rn(y);
  if (p.stats) {
    atomicAdd(p.stats + co, y);
    atomicAdd(p.stats + g.Co + co, y * y);
  }
}

struct WgradSimtParams {
  WgradTcParams w;
  int m_per_split;
};

// one thread per (co, ci, tap); blockIdx.y splits M
__global__ void conv_wgrad_simt_kernel(const WgradSimtParams q) {
  const WgradTcParams& p = q.w;
  const SfbConvGeom& g = p.g;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)g.Co * p.Ci_real * p.taps;
  if (idx >= total) return;
  const int tap = (int)(idx % p.taps);
  const int ci = (int)((idx / p.taps) % p.Ci_real);
  const int co = (int)(idx / ((long long)p.taps * p.Ci_real));
  const int khw = g.kH * g.kW;
  const int kt = tap / khw, kh = (tap % khw) / g.kW, kw = tap % g.kW;
  const int mb = blockIdx.y * q.m_per_split;
  const int me = min(p.M, mb + q.m_per_split);
  float acc = 0.f;
  for (int m = mb; m < me; ++m) {
    const RowCoord rc = make_row(g, m, p.M);
    const int pix = gather_pixel(g, rc, kt, kh, kw, true);
    if (pix < 0) continue;
    float x = __bfloat162float(p.X[(size_t)pix * g.g_ld + g.g_coff + ci]);
    if (p.pro_scale) {
      x = fmaf(x, p.pro_scale[ci], p.pro_shift[ci]);
      if (p.pro_relu) x = fmaxf(x, 0.f);
      x = __bfloat162float(__float2bfloat16_rn(x));
    }
    acc = fmaf(x, __bfloat162float(p.dY[(size_t)m * p.dy_ld + p.dy_coff + co]), acc);
  }
  atomicAdd(p.dW + idx, acc);
}

}  // namespace sfb
