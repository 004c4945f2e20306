// elementwise.cuh -- HBM-bound kernels of the SlowFast hot path: layout conversion, weight packing,
// BatchNorm finalize / fold / backward, residual tail, stem max-pool, head pooling, FC head.
// All activation tensors are bf16 NDHWC views (ptr, ld, coff); every thread moves 16 bytes
// (8 channels) per access; per-channel reductions use registers -> shared -> one atomic per block.
#pragma once
#include "common.cuh"

namespace sfb {

struct View {
  const __nv_bfloat16* p;
  int ld, coff;
};
struct ViewW {
  __nv_bfloat16* p;
  int ld, coff;
};
__device__ __forceinline__ uint4 vload(const View& v, long long m, int c) {
  return *reinterpret_cast<const uint4*>(v.p + (size_t)m * v.ld + v.coff + c);
}
__device__ __forceinline__ void vstore(const ViewW& v, long long m, int c, const uint4& q) {
  *reinterpret_cast<uint4*>(v.p + (size_t)m * v.ld + v.coff + c) = q;
}

// ---- input layout: fp32 strided NCTHW -> bf16 NDHWC (channels zero padded to ld) ----------------
__global__ void layout_input_kernel(const float* __restrict__ x, long long sn, long long sc, long long st,
                                    long long sh, long long sw, int N, int C, int T, int H, int W,
                                    __nv_bfloat16* __restrict__ out, int ld) {
  const long long total = (long long)N * T * H * W;
  for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < total;
       pix += (long long)gridDim.x * blockDim.x) {
    const int w = (int)(pix % W);
    long long q = pix / W;
    const int h = (int)(q % H);
    q /= H;
    const int t = (int)(q % T);
    const int n = (int)(q / T);
    const float* src = x + n * sn + t * st + h * sh + w * sw;
    __nv_bfloat16* dst = out + (size_t)pix * ld;
    for (int c0 = 0; c0 < ld; c0 += 8) {
      float f[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = (c0 + i < C) ? __ldg(src + (long long)(c0 + i) * sc) : 0.f;
      *reinterpret_cast<uint4*>(dst + c0) = pack8(f);
    }
  }
}

// ---- weight packing: fp32 OIDHW -> bf16 [rows_pad][k_pad] ---------------------------------------
// (bodies are __device__ functions over [first, total) with a stride so the batched launch below can run them too)
// IdxT = unsigned for operands below 2^31 elements (every layer of this network): the 64-bit div/mod pair per element
// was most of the kernel's instruction count.
template <typename IdxT>
__device__ __forceinline__ void pack_weight_loop(const float* __restrict__ w, int Co, int Ci, int taps, int Cg,
                                                 int transposed, __nv_bfloat16* __restrict__ wp, IdxT total, IdxT k_pad,
                                                 IdxT first, IdxT step) {
  for (IdxT idx = first; idx < total; idx += step) {
    const int row = (int)(idx / k_pad);
    const int k = (int)(idx - (IdxT)row * k_pad);
    float v = 0.f;
    const int tap = k / Cg, c = k - tap * Cg;
    if (tap < taps) {
      if (!transposed) {
        if (row < Co && c < Ci) v = w[((size_t)row * Ci + c) * taps + tap];
      } else {
        if (row < Ci && c < Co) v = w[((size_t)c * Ci + row) * taps + tap];
      }
    }
    wp[idx] = __float2bfloat16_rn(v);
  }
}
__device__ __forceinline__ void pack_weight_body(const float* __restrict__ w, int Co, int Ci, int taps, int Cg,
                                                 int transposed, __nv_bfloat16* __restrict__ wp, int rows_pad, int k_pad,
                                                 long long first, long long step) {
  const long long total = (long long)rows_pad * k_pad;
  if (total + step < 0x7fffffffll)
    pack_weight_loop<unsigned>(w, Co, Ci, taps, Cg, transposed, wp, (unsigned)total, (unsigned)k_pad, (unsigned)first,
                               (unsigned)step);
  else
    pack_weight_loop<long long>(w, Co, Ci, taps, Cg, transposed, wp, total, (long long)k_pad, first, step);
}
__global__ void pack_weight_kernel(const float* __restrict__ w, int Co, int Ci, int taps, int Cg, int transposed,
                                   __nv_bfloat16* __restrict__ wp, int rows_pad, int k_pad) {
  pack_weight_body(w, Co, Ci, taps, Cg, transposed, wp, rows_pad, k_pad,
                   (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}


// ---- strided dgrad: parity classes -----------------------------------------------------------------------
// dX[o] = sum_k dY[(o + p - k)/s] W[k] over the taps with (o + p - k) % s == 0.  Per dimension the outputs split into
// s classes r = (o + p) % s; class r uses taps k = r + s*j and is a STRIDE-1 transposed conv in a = (o + p - r)/s:
// dX[s*a + r - p] = sum_j dY[a - j] W[r + s*j].  The packed dgrad operand is class-major so each class is one
// contiguous K range (padded to 64) of the same matrix.
struct DgradClass {
  int r[3], nj[3], amin[3], na[3];
  int k0, taps;
};
struct DgradClassTable {
  int n;
  int s[3], k[3];
  DgradClass c[16];
};

// fp32 OIDHW (Co, Ci, kT,kH,kW) -> bf16 [rows_pad (>= Ci)][k_pad], column = class k0 + tap_in_class*Cg + co
template <typename IdxT>
__device__ __forceinline__ void pack_weight_dgrad_loop(const float* __restrict__ w, int Co, int Ci, int Cg,
                                                       const DgradClassTable& tb, __nv_bfloat16* __restrict__ wp,
                                                       IdxT total, IdxT k_pad, IdxT first, IdxT step) {
  const int taps_all = tb.k[0] * tb.k[1] * tb.k[2];
  for (IdxT idx = first; idx < total; idx += step) {
    const int row = (int)(idx / k_pad);
    const int col = (int)(idx - (IdxT)row * k_pad);
    float v = 0.f;
    for (int ci = 0; ci < tb.n; ++ci) {
      const DgradClass& c = tb.c[ci];
      const int rel = col - c.k0;
      if (rel >= 0 && rel < c.taps * Cg) {
        const int tap = rel / Cg, co = rel - tap * Cg;
        const int jw = tap % c.nj[2];
        const int jh = (tap / c.nj[2]) % c.nj[1];
        const int jt = tap / (c.nj[2] * c.nj[1]);
        const int kt = c.r[0] + tb.s[0] * jt, kh = c.r[1] + tb.s[1] * jh, kw = c.r[2] + tb.s[2] * jw;
        const int otap = (kt * tb.k[1] + kh) * tb.k[2] + kw;
        if (row < Ci && co < Co) v = w[((size_t)co * Ci + row) * taps_all + otap];
        break;
      }
    }
    wp[idx] = __float2bfloat16_rn(v);
  }
}
__device__ __forceinline__ void pack_weight_dgrad_body(const float* __restrict__ w, int Co, int Ci, int Cg,
                                                       const DgradClassTable& tb, __nv_bfloat16* __restrict__ wp,
                                                       int rows_pad, int k_pad, long long first, long long step) {
  const long long total = (long long)rows_pad * k_pad;
  if (total + step < 0x7fffffffll)
    pack_weight_dgrad_loop<unsigned>(w, Co, Ci, Cg, tb, wp, (unsigned)total, (unsigned)k_pad, (unsigned)first,
                                     (unsigned)step);
  else
    pack_weight_dgrad_loop<long long>(w, Co, Ci, Cg, tb, wp, total, (long long)k_pad, first, step);
}
__global__ void pack_weight_dgrad_kernel(const float* __restrict__ w, int Co, int Ci, int Cg, DgradClassTable tb,
                                         __nv_bfloat16* __restrict__ wp, int rows_pad, int k_pad) {
  pack_weight_dgrad_body(w, Co, Ci, Cg, tb, wp, rows_pad, k_pad, (long long)blockIdx.x * blockDim.x + threadIdx.x,
                         (long long)gridDim.x * blockDim.x);
}

__global__ void zero_view_kernel(ViewW v, long long M, int C) {
  const int CV = C >> 3;
  const long long total = M * CV;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x)
    vstore(v, idx / CV, (int)(idx % CV) << 3, make_uint4(0, 0, 0, 0));
}


// ---- pixel folding for narrow-channel layers -------------------------------------------------------------
// A stride-1 conv over a dense NDHWC tensor with C < 64 channels moves 16/32-channel (32/64-byte) rows, which
// bounds TMA by its request rate, not by bytes.  Viewing f consecutive W pixels as ONE pixel with f*C channels is the
// same memory; the conv becomes a conv with a block-banded weight
//   Wf[(j,co)][(j',c)][kt][kh][kw'] = W[co][c][kt][kh][kw],  kw = f*(kw' - kW'/2) + j' - j + kW/2   (0 outside [0,kW))
// (kW' = 1 for kW = 1, 3 for kW = 3).  The heavy kernels run unchanged on 128-byte rows; these three small kernels
// translate weights, BatchNorm statistics and weight gradients between the folded and the reference layout.
__device__ __forceinline__ void fold_weight_body(const float* __restrict__ w, int Co, int Ci, int kT, int kH, int kW,
                                                 int f, int ldo, int ldi, float* __restrict__ wf, long long first,
                                                 long long step) {
  const int kWf = kW == 1 ? 1 : 3;
  const int Cof = f * ldo, Cif = f * ldi;
  const long long total = (long long)Cof * Cif * kT * kH * kWf;
  for (long long idx = first; idx < total; idx += step) {
    long long q = idx;
    const int kwf = (int)(q % kWf); q /= kWf;
    const int kh = (int)(q % kH); q /= kH;
    const int kt = (int)(q % kT); q /= kT;
    const int cif = (int)(q % Cif);
    const int cof = (int)(q / Cif);
    const int j = cof / ldo, co = cof - j * ldo;
    const int jp = cif / ldi, c = cif - jp * ldi;
    const int kw = f * (kwf - kWf / 2) + jp - j + kW / 2;
    float v = 0.f;
    if (co < Co && c < Ci && kw >= 0 && kw < kW) v = w[((((size_t)co * Ci + c) * kT + kt) * kH + kh) * kW + kw];
    wf[idx] = v;
  }
}
__global__ void fold_weight_kernel(const float* __restrict__ w, int Co, int Ci, int kT, int kH, int kW, int f, int ldo,
                                   int ldi, float* __restrict__ wf) {
  fold_weight_body(w, Co, Ci, kT, kH, kW, f, ldo, ldi, wf, (long long)blockIdx.x * blockDim.x + threadIdx.x,
                   (long long)gridDim.x * blockDim.x);
}

// dW[co][c][kt][kh][kw] += sum over (j, j', kw') mapping to kw of dWf[(j,co)][(j',c)][kt][kh][kw']
__global__ void unfold_wgrad_kernel(const float* __restrict__ dwf, int Co, int Ci, int kT, int kH, int kW, int f, int ldo,
                                    int ldi, float* __restrict__ dw) {
  const int kWf = kW == 1 ? 1 : 3;
  const int Cif = f * ldi;
  const long long total = (long long)Co * Ci * kT * kH * kW;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    long long q = idx;
    const int kw = (int)(q % kW); q /= kW;
    const int kh = (int)(q % kH); q /= kH;
    const int kt = (int)(q % kT); q /= kT;
    const int c = (int)(q % Ci);
    const int co = (int)(q / Ci);
    float acc = 0.f;
    for (int j = 0; j < f; ++j)
      for (int kwf = 0; kwf < kWf; ++kwf) {
        const int jp = kw - kW / 2 + j - f * (kwf - kWf / 2);
        if (jp < 0 || jp >= f) continue;
        acc += dwf[((((size_t)(j * ldo + co) * Cif + (jp * ldi + c)) * kT + kt) * kH + kh) * kWf + kwf];
      }
    dw[idx] += acc;
  }
}

// stats[c] += sum_j sf[j*ld + c] ; stats[C + c] += sum_j sf[f*ld + j*ld + c]
__global__ void unfold_stats_kernel(const float* __restrict__ sf, int f, int ld, int C, float* __restrict__ stats) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float a = 0.f, b = 0.f;
  for (int j = 0; j < f; ++j) { a += sf[j * ld + c]; b += sf[f * ld + j * ld + c]; }
  stats[c] += a;
  stats[C + c] += b;
}

// one channel: (sum, sum of squares) -> mean, biased variance (the cancelling difference in double), invstd
__device__ __forceinline__ void bn_moments(float sum, float sq, double inv_count, float eps, float* mean, float* var,
                                           float* invstd) {
  const double m = (double)sum * inv_count;
  double v = (double)sq * inv_count - m * m;
  if (v < 0.0) v = 0.0;
  *mean = (float)m;
  *var = (float)v;
  *invstd = 1.0f / sqrtf((float)v + eps);
}
// ---- BatchNorm finalize / fold ---------------------------------------------------------------------
__global__ void bn_finalize_kernel(const float* __restrict__ stats, int C, double inv_count, double unbias,
                                   const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                   float momentum, float* running_mean, float* running_var, float* scale,
                                   float* shift, float* save_mean, float* save_invstd) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float mean, var, invstd;
  bn_moments(stats[c], stats[C + c], inv_count, eps, &mean, &var, &invstd);
  const float sc = gamma[c] * invstd;
  scale[c] = sc;
  shift[c] = beta[c] - mean * sc;
  save_mean[c] = mean;
  save_invstd[c] = invstd;
  if (running_mean) {
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (var * (float)unbias);
  }
}

__global__ void bn_fold_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                               const float* __restrict__ rm, const float* __restrict__ rv, float eps, int C,
                               float* scale, float* shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float sc = gamma[c] * rsqrtf(rv[c] + eps);
  scale[c] = sc;
  shift[c] = beta[c] - rm[c] * sc;
}

// ---- residual tail: out = relu(y*s+b + (r*rs+rb | r)) -------------------------------------------------
// block = (CV channel-vectors) x (RL row lanes): a thread keeps its 8 channels' constants in registers and has
// kTailUnroll rows of every operand in flight (one 16-byte load per row and tensor was latency-bound at 75 % of HBM).
// mask_bits (optional, [M][C/8] bytes, bit i = channel 8*cv+i): the ReLU mask of `out`, so that the backward pass
// reads one byte instead of 16 to know where the gradient passes.
// RAW: the BatchNorm coefficients are derived here from the raw sums the producing conv accumulated (what
// bn_finalize_kernel does in a launch of its own: ~150 launches of ~7 us per training step); every thread needs only
// its 8 channels, the first block's first row lane also stores scale / shift / mean / invstd for the backward pass
// and updates the running statistics.  fold_f > 1: the sums are in the pixel-folded layout (f column groups).
struct BnRawDev {
  const float* stats;
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  float* scale;
  float* shift;
  float* save_mean;
  float* save_invstd;
  double inv_count, unbias;
  float eps, momentum;
  int fold_f, fold_ld;
};
__device__ __forceinline__ void bn_raw_channel(const BnRawDev& q, int C, int ch, bool writer, float* s, float* b) {
  float sum, sq;
  if (q.fold_f > 1) {
    sum = 0.f; sq = 0.f;
    for (int j = 0; j < q.fold_f; ++j) {
      sum += q.stats[j * q.fold_ld + ch];
      sq += q.stats[q.fold_f * q.fold_ld + j * q.fold_ld + ch];
    }
  } else {
    sum = q.stats[ch];
    sq = q.stats[C + ch];
  }
  float mean, var, invstd;
  bn_moments(sum, sq, q.inv_count, q.eps, &mean, &var, &invstd);
  const float sc = q.gamma[ch] * invstd;
  *s = sc;
  *b = q.beta[ch] - mean * sc;
  if (writer) {
    q.scale[ch] = sc;
    q.shift[ch] = *b;
    q.save_mean[ch] = mean;
    q.save_invstd[ch] = invstd;
    if (q.running_mean) {
      q.running_mean[ch] = (1.f - q.momentum) * q.running_mean[ch] + q.momentum * mean;
      q.running_var[ch] = (1.f - q.momentum) * q.running_var[ch] + q.momentum * (var * (float)q.unbias);
    }
  }
}

constexpr int kTailThreads = 256;
template <bool HAS_R, bool RAW>
__global__ void __launch_bounds__(kTailThreads, 2)
bn_add_relu_kernel(View y, const float* __restrict__ scale, const float* __restrict__ shift, View r,
                   const float* __restrict__ r_scale, const float* __restrict__ r_shift, ViewW out,
                   uint8_t* __restrict__ mask_bits, long long M, int C, int relu, BnRawDev yraw, BnRawDev rraw) {
  const int CV = C >> 3;
  const int RL = kTailThreads / CV;
  const int tx = threadIdx.x % CV;
  const int ty = threadIdx.x / CV;
  const int c = tx << 3;
  float s[8], b[8], rs[8], rb[8];
  constexpr bool has_r = HAS_R;
  constexpr int kTailUnroll = HAS_R ? 4 : 8;  // rows in flight per thread: 8 x 16 B loads either way
  const bool r_bn = RAW ? rraw.stats != nullptr : r_scale != nullptr;
  if constexpr (RAW) {
    // each channel is finalised once per block and shared (the double-precision moments are too slow to repeat in
    // every row lane), block 0 stores the coefficients / updates the running statistics
    __shared__ float coef[4][2048];
    const bool writer = blockIdx.x == 0;
    for (int ch = threadIdx.x; ch < C; ch += kTailThreads) {
      bn_raw_channel(yraw, C, ch, writer, &coef[0][ch], &coef[1][ch]);
      if (has_r && r_bn) bn_raw_channel(rraw, C, ch, writer, &coef[2][ch], &coef[3][ch]);
    }
    __syncthreads();
    if (ty >= RL) return;
    load8f(&coef[0][c], s);
    load8f(&coef[1][c], b);
    if (has_r && r_bn) { load8f(&coef[2][c], rs); load8f(&coef[3][c], rb); }
  } else {
    if (ty >= RL) return;
    load8f(scale + c, s);
    load8f(shift + c, b);
    if (r_bn) { load8f(r_scale + c, rs); load8f(r_shift + c, rb); }
  }
  const long long stride = (long long)gridDim.x * RL;
  for (long long m0 = (long long)blockIdx.x * RL + ty; m0 < M; m0 += stride * kTailUnroll) {
    uint4 qy[kTailUnroll], qr[kTailUnroll];
#pragma unroll
    for (int u = 0; u < kTailUnroll; ++u) {
      const long long m = m0 + u * stride;
      if (m < M) {
        qy[u] = vload(y, m, c);
        if (has_r) qr[u] = vload(r, m, c);
      }
    }
#pragma unroll
    for (int u = 0; u < kTailUnroll; ++u) {
      const long long m = m0 + u * stride;
      if (m < M) {
        float f[8];
        unpack8(qy[u], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = fmaf(f[i], s[i], b[i]);
        if (has_r) {
          float rf[8];
          unpack8(qr[u], rf);
          if (r_bn) {
#pragma unroll
            for (int i = 0; i < 8; ++i) rf[i] = fmaf(rf[i], rs[i], rb[i]);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] += rf[i];
        }
        if (relu) {
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = fmaxf(f[i], 0.f);
        }
        const uint4 q = pack8(f);
        vstore(out, m, c, q);
        if (mask_bits) {
          // from the bf16-rounded output, i.e. exactly (out > 0) as the backward would have recomputed it
          float o[8];
          unpack8(q, o);
          uint32_t bits = 0;
#pragma unroll
          for (int i = 0; i < 8; ++i) bits |= (o[i] > 0.f ? 1u : 0u) << i;
          mask_bits[m * CV + tx] = (uint8_t)bits;
        }
      }
    }
  }
}

// ---- BatchNorm(+ReLU) backward ----------------------------------------------------------------------
constexpr int kRedThreads = 256;
constexpr int kBnUnroll = 4;  // independent 16-byte loads in flight per thread and tensor (HBM latency hiding)

// raw operands of kBnUnroll rows (loads issued back to back), then the masked gradient / xhat per row
struct BnBwdRaw {
  uint4 dz[kBnUnroll], y[kBnUnroll], mk[kBnUnroll];
};
// ReLU mask source: 0 = recomputed from y (scale*y+shift > 0), 1 = sign of a bf16 tensor, 2 = packed bits [M][C/8]
__device__ __forceinline__ void bn_bwd_fetch(const View& dz, const View& y, const View& msk, const uint8_t* bits, int CV,
                                             int mode, const long long* m, const bool* ok, int c, BnBwdRaw& r) {
#pragma unroll
  for (int u = 0; u < kBnUnroll; ++u) {
    if (ok[u]) {
      r.dz[u] = vload(dz, m[u], c);
      r.y[u] = vload(y, m[u], c);
      if (mode == 1) r.mk[u] = vload(msk, m[u], c);
      else if (mode == 2) r.mk[u].x = bits[m[u] * CV + (c >> 3)];
    }
  }
}
__device__ __forceinline__ void bn_apply_mask(const BnBwdRaw& r, int u, int mode, const float* yv, const float* s,
                                              const float* b, float* g) {
  if (mode == 1) {
    float mv[8];
    unpack8(r.mk[u], mv);
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = mv[i] > 0.f ? g[i] : 0.f;
  } else if (mode == 2) {
    const uint32_t bits = r.mk[u].x;
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = (bits >> i) & 1u ? g[i] : 0.f;
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = fmaf(yv[i], s[i], b[i]) > 0.f ? g[i] : 0.f;
  }
}
__device__ __forceinline__ void bn_bwd_math(const BnBwdRaw& r, int u, int mode, int relu, const float* s, const float* b,
                                            const float* mu, const float* is, float* g, float* xhat) {
  float yv[8];
  unpack8(r.dz[u], g);
  unpack8(r.y[u], yv);
#pragma unroll
  for (int i = 0; i < 8; ++i) xhat[i] = (yv[i] - mu[i]) * is[i];
  if (relu) bn_apply_mask(r, u, mode, yv, s, b, g);
}

// block = (CV channel-vectors) x (RL row lanes); grid.x strides over row groups.
// MODE = ReLU mask source (bn_bwd_fetch; 3 = no ReLU).  A read-only pass lives on bytes in flight: with 121
// registers (2 blocks / SM) and 4 rows per tensor it reached 5.3 TB/s (ncu r1g: 24 % warps active, 40 % long
// scoreboard), so the mask source is a template parameter (dead operands cost no registers), the centred sum
// sum g*(y - mean) replaces sum g*xhat (invstd is applied once per block) and kRedUnroll rows are in flight.
constexpr int kRedUnroll = 8;
template <int MODE>
__global__ void __launch_bounds__(kRedThreads, 2)
bn_bwd_reduce_kernel(View dz, View y, View msk, const uint8_t* __restrict__ mask_bits, const float* __restrict__ scale,
                     const float* __restrict__ shift, const float* __restrict__ mean, const float* __restrict__ invstd,
                     long long M, int C, float* dgamma, float* dbeta) {
  __shared__ float red[kRedThreads * 16];
  const int CV = C >> 3;
  const int RL = kRedThreads / CV;  // >= 1 (C <= 2048)
  const int tx = threadIdx.x % CV;
  const int ty = threadIdx.x / CV;
  const int c = tx << 3;
  float s1[8], s2[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { s1[i] = 0.f; s2[i] = 0.f; }
  if (ty < RL) {
    float sc[8], sh[8], mu[8];
    if (MODE == 0) { load8f(scale + c, sc); load8f(shift + c, sh); }
    load8f(mean + c, mu);
    const long long stride = (long long)gridDim.x * RL;
    for (long long m0 = (long long)blockIdx.x * RL + ty; m0 < M; m0 += stride * kRedUnroll) {
      uint4 qd[kRedUnroll], qy[kRedUnroll], qm[kRedUnroll];
#pragma unroll
      for (int u = 0; u < kRedUnroll; ++u) {
        const long long m = m0 + u * stride;
        if (m < M) {
          qd[u] = vload(dz, m, c);
          qy[u] = vload(y, m, c);
          if (MODE == 1) qm[u] = vload(msk, m, c);
          else if (MODE == 2) qm[u].x = mask_bits[m * CV + tx];
        }
      }
#pragma unroll
      for (int u = 0; u < kRedUnroll; ++u) {
        if (m0 + u * stride < M) {
          float g[8], yv[8];
          unpack8(qd[u], g);
          unpack8(qy[u], yv);
          if (MODE == 1) {
            float mv[8];
            unpack8(qm[u], mv);
#pragma unroll
            for (int i = 0; i < 8; ++i) g[i] = mv[i] > 0.f ? g[i] : 0.f;
          } else if (MODE == 2) {
            const uint32_t bits = qm[u].x;
#pragma unroll
            for (int i = 0; i < 8; ++i) g[i] = (bits >> i) & 1u ? g[i] : 0.f;
          } else if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) g[i] = fmaf(yv[i], sc[i], sh[i]) > 0.f ? g[i] : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) { s1[i] += g[i]; s2[i] = fmaf(g[i], yv[i] - mu[i], s2[i]); }
        }
      }
    }
    float is[8];
    load8f(invstd + c, is);
#pragma unroll
    for (int i = 0; i < 8; ++i) s2[i] *= is[i];
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    red[threadIdx.x * 16 + i] = s1[i];
    red[threadIdx.x * 16 + 8 + i] = s2[i];
  }
  __syncthreads();
  // thread t < CV*16 sums over row lanes
  for (int t = threadIdx.x; t < CV * 16; t += kRedThreads) {
    const int cv = t / 16, e = t % 16;
    float acc = 0.f;
    for (int r = 0; r < RL; ++r) acc += red[(r * CV + cv) * 16 + e];
    if (e < 8) atomicAdd(dbeta + cv * 8 + e, acc);
    else atomicAdd(dgamma + cv * 8 + (e - 8), acc);
  }
}

// block = (CV channel-vectors) x (RL row lanes) like the reduction, so the per-channel constants are loaded once
template <int MASK, bool GOUT>  // MASK: ReLU mask source (see bn_bwd_fetch)
__global__ void __launch_bounds__(kRedThreads, 2)
bn_bwd_apply_kernel(View dz, View y, View msk, const uint8_t* __restrict__ mask_bits, const float* __restrict__ scale,
                    const float* __restrict__ shift, const float* __restrict__ mean,
                    const float* __restrict__ invstd, const float* __restrict__ gamma,
                    const float* __restrict__ dgamma, const float* __restrict__ dbeta, int relu,
                    long long M, int C, float invM, ViewW dy, ViewW gout, int g_accumulate) {
  const int CV = C >> 3;
  const int RL = kRedThreads / CV;
  const int tx = threadIdx.x % CV;
  const int ty = threadIdx.x / CV;
  if (ty >= RL) return;
  const int c = tx << 3;
  // dy = gamma*invstd*(g - dbeta/M - xhat*dgamma/M) = k0*g - ka*y + kb   (xhat = (y-mean)*invstd folded in)
  float sc[8], sh[8], k0[8], ka[8], kb[8];
  constexpr int mode = MASK;
  if (relu && mode == 0) { load8f(scale + c, sc); load8f(shift + c, sh); }
  {
    float ga[8], dg[8], db[8], mu[8], is[8];
    load8f(mean + c, mu);
    load8f(invstd + c, is);
    load8f(gamma + c, ga);
    load8f(dgamma + c, dg);
    load8f(dbeta + c, db);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      k0[i] = ga[i] * is[i];
      ka[i] = k0[i] * dg[i] * invM * is[i];
      kb[i] = ka[i] * mu[i] - k0[i] * db[i] * invM;
    }
  }
  const View gv{gout.p, gout.ld, gout.coff};
  const long long stride = (long long)gridDim.x * RL;
  for (long long m0 = (long long)blockIdx.x * RL + ty; m0 < M; m0 += stride * kBnUnroll) {
    long long m[kBnUnroll];
    bool ok[kBnUnroll];
#pragma unroll
    for (int u = 0; u < kBnUnroll; ++u) { m[u] = m0 + u * stride; ok[u] = m[u] < M; }
    BnBwdRaw r;
    bn_bwd_fetch(dz, y, msk, mask_bits, CV, relu ? mode : 3, m, ok, c, r);
    uint4 old[kBnUnroll];
    if (GOUT && g_accumulate) {
#pragma unroll
      for (int u = 0; u < kBnUnroll; ++u)
        if (ok[u]) old[u] = vload(gv, m[u], c);
    }
#pragma unroll
    for (int u = 0; u < kBnUnroll; ++u) {
      if (ok[u]) {
        float g[8], yv[8], o[8];
        unpack8(r.dz[u], g);
        unpack8(r.y[u], yv);
        if (relu) bn_apply_mask(r, u, mode, yv, sc, sh, g);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = fmaf(k0[i], g[i], fmaf(-ka[i], yv[i], kb[i]));
        vstore(dy, m[u], c, pack8(o));
        if (GOUT) {
          if (g_accumulate) {
            float of[8];
            unpack8(old[u], of);
#pragma unroll
            for (int i = 0; i < 8; ++i) g[i] += of[i];
          }
          vstore(gout, m[u], c, pack8(g));
        }
      }
    }
  }
}

// ---- block tail with a projection shortcut: both BatchNorms in one reduce pass and one apply pass ----------------------------
// O = ReLU(BN_c(y_c) + BN_s(y_s)) (first block of every stage): the two BatchNorm backward passes share the masked
// gradient g = dO * mask(O).  Run one after the other they read dO and the mask four times; here one reduction
// (dO, bits, y_c, y_s -> both layers' sums) and one apply (-> dy_c and dy_s) read them twice.
struct BnTail2 {
  View yc, ys;
  const float* mean_c; const float* invstd_c; const float* gamma_c; float* dgamma_c; float* dbeta_c;
  const float* mean_s; const float* invstd_s; const float* gamma_s; float* dgamma_s; float* dbeta_s;
};
constexpr int kTail2Unroll = 4;
__global__ void __launch_bounds__(kRedThreads, 2)
bn_bwd_tail2_reduce_kernel(View dz, const uint8_t* __restrict__ mask_bits, BnTail2 t, long long M, int C) {
  __shared__ float red[kRedThreads * 24];
  const int CV = C >> 3;
  const int RL = kRedThreads / CV;
  const int tx = threadIdx.x % CV;
  const int ty = threadIdx.x / CV;
  const int c = tx << 3;
  float s1[8], sc_[8], ss_[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { s1[i] = 0.f; sc_[i] = 0.f; ss_[i] = 0.f; }
  if (ty < RL) {
    float mc[8], ms[8];
    load8f(t.mean_c + c, mc);
    load8f(t.mean_s + c, ms);
    const long long stride = (long long)gridDim.x * RL;
    for (long long m0 = (long long)blockIdx.x * RL + ty; m0 < M; m0 += stride * kTail2Unroll) {
      uint4 qd[kTail2Unroll], qc[kTail2Unroll], qs[kTail2Unroll];
      uint32_t qb[kTail2Unroll];
#pragma unroll
      for (int u = 0; u < kTail2Unroll; ++u) {
        const long long m = m0 + u * stride;
        if (m < M) {
          qd[u] = vload(dz, m, c);
          qc[u] = vload(t.yc, m, c);
          qs[u] = vload(t.ys, m, c);
          qb[u] = mask_bits[m * CV + tx];
        }
      }
#pragma unroll
      for (int u = 0; u < kTail2Unroll; ++u) {
        if (m0 + u * stride < M) {
          float g[8], a[8], b[8];
          unpack8(qd[u], g);
          unpack8(qc[u], a);
          unpack8(qs[u], b);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            g[i] = (qb[u] >> i) & 1u ? g[i] : 0.f;
            s1[i] += g[i];
            sc_[i] = fmaf(g[i], a[i] - mc[i], sc_[i]);
            ss_[i] = fmaf(g[i], b[i] - ms[i], ss_[i]);
          }
        }
      }
    }
    float ic[8], is[8];
    load8f(t.invstd_c + c, ic);
    load8f(t.invstd_s + c, is);
#pragma unroll
    for (int i = 0; i < 8; ++i) { sc_[i] *= ic[i]; ss_[i] *= is[i]; }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    red[threadIdx.x * 24 + i] = s1[i];
    red[threadIdx.x * 24 + 8 + i] = sc_[i];
    red[threadIdx.x * 24 + 16 + i] = ss_[i];
  }
  __syncthreads();
  for (int q = threadIdx.x; q < CV * 24; q += kRedThreads) {
    const int cv = q / 24, e = q % 24;
    float acc = 0.f;
    for (int r = 0; r < RL; ++r) acc += red[(r * CV + cv) * 24 + e];
    if (e < 8) {
      atomicAdd(t.dbeta_c + cv * 8 + e, acc);
      atomicAdd(t.dbeta_s + cv * 8 + e, acc);
    } else if (e < 16) {
      atomicAdd(t.dgamma_c + cv * 8 + (e - 8), acc);
    } else {
      atomicAdd(t.dgamma_s + cv * 8 + (e - 16), acc);
    }
  }
}

__global__ void __launch_bounds__(kRedThreads, 2)
bn_bwd_tail2_apply_kernel(View dz, const uint8_t* __restrict__ mask_bits, BnTail2 t, long long M, int C, float invM,
                          ViewW dyc, ViewW dys) {
  const int CV = C >> 3;
  const int RL = kRedThreads / CV;
  const int tx = threadIdx.x % CV;
  const int ty = threadIdx.x / CV;
  if (ty >= RL) return;
  const int c = tx << 3;
  // dy = gamma*invstd*(g - dbeta/M - xhat*dgamma/M) = k0*g - ka*y + kb per layer
  float k0c[8], kac[8], kbc[8], k0s[8], kas[8], kbs[8];
  {
    float ga[8], dg[8], db[8], mu[8], is[8];
    load8f(t.mean_c + c, mu); load8f(t.invstd_c + c, is); load8f(t.gamma_c + c, ga);
    load8f(t.dgamma_c + c, dg); load8f(t.dbeta_c + c, db);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      k0c[i] = ga[i] * is[i];
      kac[i] = k0c[i] * dg[i] * invM * is[i];
      kbc[i] = kac[i] * mu[i] - k0c[i] * db[i] * invM;
    }
    load8f(t.mean_s + c, mu); load8f(t.invstd_s + c, is); load8f(t.gamma_s + c, ga);
    load8f(t.dgamma_s + c, dg); load8f(t.dbeta_s + c, db);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      k0s[i] = ga[i] * is[i];
      kas[i] = k0s[i] * dg[i] * invM * is[i];
      kbs[i] = kas[i] * mu[i] - k0s[i] * db[i] * invM;
    }
  }
  constexpr int U = 2;
  const long long stride = (long long)gridDim.x * RL;
  for (long long m0 = (long long)blockIdx.x * RL + ty; m0 < M; m0 += stride * U) {
    uint4 qd[U], qc[U], qs[U];
    uint32_t qb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long m = m0 + u * stride;
      if (m < M) {
        qd[u] = vload(dz, m, c);
        qc[u] = vload(t.yc, m, c);
        qs[u] = vload(t.ys, m, c);
        qb[u] = mask_bits[m * CV + tx];
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long m = m0 + u * stride;
      if (m < M) {
        float g[8], a[8], b[8], oc[8], os[8];
        unpack8(qd[u], g);
        unpack8(qc[u], a);
        unpack8(qs[u], b);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          g[i] = (qb[u] >> i) & 1u ? g[i] : 0.f;
          oc[i] = fmaf(k0c[i], g[i], fmaf(-kac[i], a[i], kbc[i]));
          os[i] = fmaf(k0s[i], g[i], fmaf(-kas[i], b[i], kbs[i]));
        }
        vstore(dyc, m, c, pack8(oc));
        vstore(dys, m, c, pack8(os));
      }
    }
  }
}

// ---- stem max-pool (1,3,3)/s(1,2,2)/p(0,1,1) with fused BN+ReLU on load ---------------------------------
// general path of the pooling (some BatchNorm scale <= 0: the window maximum must be taken on the transformed values);
// not inlined, so that its registers do not count against the kernel's common path
__device__ __noinline__ void maxpool_fwd_general(const View& y, const float* scale, const float* shift, long long q,
                                                 int ho, int wo, int c, int H, int W, int C, long long opix,
                                                 const ViewW& out, uint8_t* idx_out) {
  float s[8], b[8], best[8];
  int bi[8];
  load8f(scale + c, s);
  load8f(shift + c, b);
#pragma unroll
  for (int i = 0; i < 8; ++i) { best[i] = -INFINITY; bi[i] = 0; }
#pragma unroll
  for (int kh = 0; kh < 3; ++kh) {
    const int h = ho * 2 - 1 + kh;
    if ((unsigned)h >= (unsigned)H) continue;
#pragma unroll
    for (int kw = 0; kw < 3; ++kw) {
      const int w = wo * 2 - 1 + kw;
      if ((unsigned)w >= (unsigned)W) continue;
      float f[8];
      unpack8(vload(y, (q * H + h) * W + w, c), f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float v = fmaxf(fmaf(f[i], s[i], b[i]), 0.f);
        if (v > best[i]) { best[i] = v; bi[i] = kh * 3 + kw; }
      }
    }
  }
  vstore(out, opix, c, pack8(best));
  if (idx_out) {
    uint2 pk;
    pk.x = (uint32_t)bi[0] | ((uint32_t)bi[1] << 8) | ((uint32_t)bi[2] << 16) | ((uint32_t)bi[3] << 24);
    pk.y = (uint32_t)bi[4] | ((uint32_t)bi[5] << 8) | ((uint32_t)bi[6] << 16) | ((uint32_t)bi[7] << 24);
    *reinterpret_cast<uint2*>(idx_out + (size_t)opix * C + c) = pk;
  }
}

// 42 registers at most (6 blocks of 256 threads per SM): at 60 registers / 4 blocks the kernel waited on its nine tap
// loads with half the warps an SM can hold (ncu r3u: 42 % warps active, 12 long-scoreboard stalls per issue, 2.7 TB/s)
__global__ void __launch_bounds__(256, 6)
maxpool_fwd_kernel(View y, const float* __restrict__ scale, const float* __restrict__ shift, int N,
                   int T, int H, int W, int C, ViewW out, uint8_t* __restrict__ idx_out) {
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  const int CV = C >> 3;
  // grid.y = frame (n*T + t), grid.x strides over the frame's (pixel, channel vector) pairs: 32-bit index math only
  // (the 64-bit div/mod chain of a flat index cost more instructions than the pooling itself)
  const unsigned per_frame = (unsigned)(Ho * Wo * CV);
  const long long q = blockIdx.y;
  for (unsigned idx = blockIdx.x * blockDim.x + threadIdx.x; idx < per_frame; idx += gridDim.x * blockDim.x) {
    const int c = (int)(idx % (unsigned)CV) << 3;
    const unsigned pf = idx / (unsigned)CV;
    const int wo = (int)(pf % (unsigned)Wo);
    const int ho = (int)(pf / (unsigned)Wo);
    const long long opix = q * (Ho * Wo) + pf;
    float s[8], b[8], best[8];
    bool monotone = true;
    if (scale) {
      load8f(scale + c, s);
#pragma unroll
      for (int i = 0; i < 8; ++i) monotone = monotone && s[i] > 0.f;
    }
    // the coefficients are fetched again where they are used (through a pointer the compiler cannot match with the
    // loads above), so that they do not occupy 16 registers across the tap loop
    const float* const scale2 = keep_ptr(scale);
    if (monotone) {
      // BatchNorm scale > 0 (or no affine): ReLU(scale*y + shift) is non-decreasing in y, so the window maximum is taken
      // on the RAW bf16 pairs (one compare-mask + two bitwise selects per pair and tap, against ~5 fp32 instructions
      // per element and tap on the general path below -- the kernel was instruction-bound: ncu r3l, 2.9 TB/s) and the
      // affine + ReLU is applied once to the winner.  Where the transformed values tie only because ReLU clamps them
      // to 0 the recorded tap may differ from the general path's first-of-ties; the gradient routed there is masked by
      // the BatchNorm backward's ReLU mask either way.
      uint32_t bq[4], tq[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) { bq[j] = 0xFF80FF80u; tq[j] = 0u; }  // -inf pairs
#pragma unroll
      for (int kh = 0; kh < 3; ++kh) {
        const int h = ho * 2 - 1 + kh;
        if ((unsigned)h >= (unsigned)H) continue;
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const int w = wo * 2 - 1 + kw;
          if ((unsigned)w >= (unsigned)W) continue;
          const uint4 qv = vload(y, (q * H + h) * W + w, c);
          const uint32_t in[4] = {qv.x, qv.y, qv.z, qv.w};
          const uint32_t tap2 = (uint32_t)(kh * 3 + kw) * 0x00010001u;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (idx_out) {
              const uint32_t m = __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&in[j]),
                                             *reinterpret_cast<const __nv_bfloat162*>(&bq[j]));
              bq[j] = (in[j] & m) | (bq[j] & ~m);
              tq[j] = (tap2 & m) | (tq[j] & ~m);
            } else {
              const __nv_bfloat162 mx = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&in[j]),
                                                *reinterpret_cast<const __nv_bfloat162*>(&bq[j]));
              bq[j] = *reinterpret_cast<const uint32_t*>(&mx);
            }
          }
        }
      }
      uint4 o = make_uint4(bq[0], bq[1], bq[2], bq[3]);
      if (scale2) {
        load8f(scale2 + c, s);
        load8f(shift + c, b);
        unpack8(o, best);
#pragma unroll
        for (int i = 0; i < 8; ++i) best[i] = fmaxf(fmaf(best[i], s[i], b[i]), 0.f);
        o = pack8(best);
      }
      vstore(out, opix, c, o);
      if (idx_out) {
        uint2 pk;
        pk.x = __byte_perm(tq[0], tq[1], 0x6420);
        pk.y = __byte_perm(tq[2], tq[3], 0x6420);
        *reinterpret_cast<uint2*>(idx_out + (size_t)opix * C + c) = pk;
      }
      continue;
    }
    maxpool_fwd_general(y, scale2, shift, q, ho, wo, c, H, W, C, opix, out, idx_out);
  }
}

// Backward of the stem max-pool.  One thread owns a 2x2 block of input pixels (rows 2i, 2i+1; columns 2j, 2j+1) and 8
// channels: the only windows that cover those pixels are (i+u, j+v), u, v in {0, 1} -- row 2i lies in window row i alone
// (tap 1), row 2i+1 in window rows i (tap 2) and i+1 (tap 0), the same along w -- so four (arg-max bytes, gradient)
// pairs are loaded once and serve four pixels; per input pixel the kernel loaded up to four of them (2.25 on average)
// and was bound by its instruction count (ncu r3u: 96 % warps active, math-pipe throttle the top stall, 2.4 TB/s).
// The arg-max bytes are compared four at a time and turned into bf16-pair masks; a pixel covered by ONE window takes
// the masked gradient as it is, the others sum in fp32 (the same values as the per-pixel kernel produced).
__device__ __forceinline__ uint4 maxpool_mask_tap(const uint4& d, const uint2& k, int tap) {
  const uint32_t tap4 = (uint32_t)tap * 0x01010101u;
  const uint32_t m0 = __vcmpeq4(k.x, tap4), m1 = __vcmpeq4(k.y, tap4);  // 0xFF per matching channel
  uint4 r;
  r.x = d.x & __byte_perm(m0, 0, 0x1100);
  r.y = d.y & __byte_perm(m0, 0, 0x3322);
  r.z = d.z & __byte_perm(m1, 0, 0x1100);
  r.w = d.w & __byte_perm(m1, 0, 0x3322);
  return r;
}

__global__ void __launch_bounds__(256, 4)
maxpool_bwd_kernel(View dout, const uint8_t* __restrict__ idxs, int N, int T, int H, int W, int C, ViewW dz) {
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;  // = ceil(H/2), ceil(W/2): one 2x2 block per window origin
  const int CV = C >> 3;
  const unsigned per_frame = (unsigned)(Ho * Wo * CV);
  const long long q = blockIdx.y;  // frame n*T + t
  for (unsigned idx = blockIdx.x * blockDim.x + threadIdx.x; idx < per_frame; idx += gridDim.x * blockDim.x) {
    const int c = (int)(idx % (unsigned)CV) << 3;
    const unsigned pf = idx / (unsigned)CV;
    const int j = (int)(pf % (unsigned)Wo);
    const int i = (int)(pf / (unsigned)Wo);
    uint4 d[2][2];
    uint2 k[2][2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
#pragma unroll
      for (int v = 0; v < 2; ++v) {
        d[u][v] = make_uint4(0, 0, 0, 0);
        k[u][v] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);  // matches no tap
        if (i + u < Ho && j + v < Wo) {
          const long long opix = (q * Ho + (i + u)) * Wo + (j + v);
          k[u][v] = *reinterpret_cast<const uint2*>(idxs + (size_t)opix * C + c);
          d[u][v] = vload(dout, opix, c);
        }
      }
    }
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const int h = 2 * i + a;
      if (h >= H) continue;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int w = 2 * j + b;
        if (w >= W) continue;
        const long long ipix = (q * H + h) * W + w;
        // window rows / columns that cover this pixel and the tap it is in each of them
        const int nu = a == 0 ? 1 : 2, nv = b == 0 ? 1 : 2;
        if (nu == 1 && nv == 1) {
          vstore(dz, ipix, c, maxpool_mask_tap(d[0][0], k[0][0], 1 * 3 + 1));
          continue;
        }
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (u >= nu) continue;
          const int kh = a == 0 ? 1 : (u == 0 ? 2 : 0);
#pragma unroll
          for (int v = 0; v < 2; ++v) {
            if (v >= nv) continue;
            const int kw = b == 0 ? 1 : (v == 0 ? 2 : 0);
            float f[8];
            unpack8(maxpool_mask_tap(d[u][v], k[u][v], kh * 3 + kw), f);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += f[e];
          }
        }
        vstore(dz, ipix, c, pack8(acc));
      }
    }
  }
}

// ---- head: AvgPool3d(k, stride 1) -> dropout -> mean over positions -----------------------------------
__device__ __forceinline__ float dropout_keep_scale(uint64_t seed, int n, int c, int pos, float p) {
  // counter-based hash (splitmix64 finaliser) -> uniform in [0,1)
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * ((uint64_t)n * 0x100000001B3ull + (uint64_t)c * 0x1000193ull + (uint64_t)pos + 1ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  const float u = (float)(z >> 40) * (1.0f / 16777216.0f);
  return u >= p ? 1.f / (1.f - p) : 0.f;
}

// One block per (clip, window, group of <= 32 channel vectors); its 256 threads are (channel vector) x (row lanes
// splitting the window's positions), reduced through shared memory.  The first version walked a whole window per
// thread (392 / 1568 dependent loads with N*C/8 threads in flight: 0.72 ms for 130 MB).
// p_drop == 0: the windows collapse into ONE weighted sum over the map (weight = windows covering a position).
// feat must be zeroed beforehand when there is more than one window (atomic accumulation over windows).
constexpr int kHeadThreads = 256;
__global__ void __launch_bounds__(kHeadThreads)
head_pool_fwd_kernel(View x, int N, int T, int H, int W, int C, int kT, int kH, int kW, float p_drop,
                     uint64_t seed, const uint64_t* __restrict__ seed_dev, float* __restrict__ feat, int f_ld, int f_off,
                     int cvb, int npos) {
  __shared__ float red[kHeadThreads * 8];
  if (seed_dev) seed = *seed_dev;
  const int CV = C >> 3;
  const int RL = kHeadThreads / cvb;
  const int tx = threadIdx.x % cvb, ty = threadIdx.x / cvb;
  const int cv = blockIdx.y * cvb + tx;
  const int n = blockIdx.x / npos, pos = blockIdx.x - n * npos;
  const int Pt = T - kT + 1, Ph = H - kH + 1, Pw = W - kW + 1;
  const bool drop = p_drop > 0.f;
  // window origin / extent: the whole map when the windows are collapsed
  int pt = 0, ph = 0, pw = 0, eT = T, eH = H, eW = W;
  if (drop) {
    pw = pos % Pw; ph = (pos / Pw) % Ph; pt = pos / (Pw * Ph);
    eT = kT; eH = kH; eW = kW;
  }
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  if (ty < RL && cv < CV) {
    const int ne = eT * eH * eW;
    for (int e = ty; e < ne; e += RL) {
      const int w = pw + e % eW, h = ph + (e / eW) % eH, t = pt + e / (eW * eH);
      float wgt = 1.f;
      if (!drop) {
        const int wt = min(t, Pt - 1) - max(0, t - kT + 1) + 1;
        const int wh = min(h, Ph - 1) - max(0, h - kH + 1) + 1;
        const int ww = min(w, Pw - 1) - max(0, w - kW + 1) + 1;
        wgt = (float)(wt * wh * ww);
      }
      float f[8];
      unpack8(vload(x, (((long long)n * T + t) * H + h) * W + w, cv << 3), f);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = fmaf(f[i], wgt, acc[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[threadIdx.x * 8 + i] = acc[i];
  __syncthreads();
  if (ty == 0 && cv < CV) {
    const float inv = 1.f / ((float)(Pt * Ph * Pw) * (float)(kT * kH * kW));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float s = 0.f;
      for (int r = 0; r < RL; ++r) s += red[(r * cvb + tx) * 8 + i];
      const int c = (cv << 3) + i;
      if (drop) s *= dropout_keep_scale(seed, n, f_off + c, pos, p_drop);
      float* dst = feat + (size_t)n * f_ld + f_off + c;
      if (npos == 1) *dst = s * inv;
      else atomicAdd(dst, s * inv);
    }
  }
}

__global__ void head_pool_bwd_kernel(const float* __restrict__ dfeat, int f_ld, int f_off, int N, int T, int H, int W,
                                     int C, int kT, int kH, int kW, float p_drop, uint64_t seed,
                                     const uint64_t* __restrict__ seed_dev, ViewW dx) {
  if (seed_dev) seed = *seed_dev;
  const int CV = C >> 3;
  const long long total = (long long)N * T * H * W * CV;
  const int Pt = T - kT + 1, Ph = H - kH + 1, Pw = W - kW + 1;
  const float inv = 1.f / ((float)(Pt * Ph * Pw) * (float)(kT * kH * kW));
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % CV) << 3;
    long long q = idx / CV;
    const long long pix = q;
    const int w = (int)(q % W);
    q /= W;
    const int h = (int)(q % H);
    q /= H;
    const int t = (int)(q % T);
    const int n = (int)(q / T);
    const int t_lo = max(0, t - kT + 1), t_hi = min(t, Pt - 1);
    const int h_lo = max(0, h - kH + 1), h_hi = min(h, Ph - 1);
    const int w_lo = max(0, w - kW + 1), w_hi = min(w, Pw - 1);
    float wsum[8];
    if (p_drop <= 0.f) {
      const float cnt = (float)((t_hi - t_lo + 1) * (h_hi - h_lo + 1) * (w_hi - w_lo + 1));
#pragma unroll
      for (int i = 0; i < 8; ++i) wsum[i] = cnt;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) wsum[i] = 0.f;
      for (int pt = t_lo; pt <= t_hi; ++pt)
        for (int ph = h_lo; ph <= h_hi; ++ph)
          for (int pw = w_lo; pw <= w_hi; ++pw) {
            const int pos = (pt * Ph + ph) * Pw + pw;
#pragma unroll
            for (int i = 0; i < 8; ++i) wsum[i] += dropout_keep_scale(seed, n, f_off + c + i, pos, p_drop);
          }
    }
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = dfeat[(size_t)n * f_ld + f_off + c + i] * wsum[i] * inv;
    vstore(dx, pix, c, pack8(o));
  }
}

// ---- FC head (fp32) ---------------------------------------------------------------------------------------
// one warp per (n, k)
__global__ void fc_fwd_kernel(const float* __restrict__ feat, const float* __restrict__ w, const float* __restrict__ b,
                              int N, int K, int C, float* __restrict__ logits) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (gw >= N * C) return;
  const int k = gw % C, n = gw / C;
  const float* f = feat + (size_t)n * K;
  const float* wr = w + (size_t)k * K;
  float acc = 0.f;
  for (int i = lane; i < K; i += 32) acc = fmaf(f[i], wr[i], acc);
#pragma unroll
  for (int off = 16; off; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (lane == 0) logits[(size_t)n * C + k] = acc + (b ? b[k] : 0.f);
}
// dfeat[n][i] = sum_k dl[n][k] w[k][i]
__global__ void fc_bwd_dfeat_kernel(const float* __restrict__ dl, const float* __restrict__ w, int N, int K, int C,
                                    float* __restrict__ dfeat) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * K) return;
  const int i = idx % K, n = idx / K;
  float acc = 0.f;
  for (int k = 0; k < C; ++k) acc = fmaf(dl[(size_t)n * C + k], w[(size_t)k * K + i], acc);
  dfeat[idx] = acc;
}
// dw[k][i] += sum_n dl[n][k] feat[n][i];  db[k] += sum_n dl[n][k]
__global__ void fc_bwd_dw_kernel(const float* __restrict__ dl, const float* __restrict__ feat, int N, int K, int C,
                                 float* __restrict__ dw, float* __restrict__ db) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < C * K) {
    const int i = idx % K, k = idx / K;
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc = fmaf(dl[(size_t)n * C + k], feat[(size_t)n * K + i], acc);
    dw[idx] += acc;
  }
  if (idx < C && db) {
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc += dl[(size_t)n * C + idx];
    db[idx] += acc;
  }
}

}  // namespace sfb
