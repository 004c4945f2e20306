// conv_tc.cuh -- implicit-GEMM Conv3d on tcgen05 tensor cores (sm_100a).
//
//   conv_gemm_tc  : fprop and dgrad.  D[128 x BN] (TMEM, fp32) = A[128 x K] * B[BN x K]^T
//                   A = im2col rows of the NDHWC bf16 activation, gathered by 8 producer warps
//                       (16-byte granules, optional BatchNorm+ReLU prologue in fp32) into
//                       SWIZZLE_128B K-major shared-memory tiles;
//                   B = packed bf16 weights, TMA 2D tiled loads (SWIZZLE_128B);
//                   one thread issues tcgen05.mma (M=128, N=BN, K=16); the accumulator lives in TMEM;
//                   epilogue warps read it back with tcgen05.ld and fuse affine / residual / ReLU /
//                   per-channel sum & sum-of-squares (training BatchNorm statistics).
//   conv_wgrad_tc : dW[co][(tap,ci)] = sum_m dY[m][co] * im2col(X)[m][(tap,ci)].
//                   Both operands are staged position-major ([64 positions][64 channels] blocks, i.e.
//                   MN-major for the MMA), D[128 co x <=256 cols] in TMEM, split over M across CTAs,
//                   fp32 atomics into the OIDHW gradient.
//
// Replaces cuDNN's cudnn_convolution fprop / bwd-data / bwd-filter as used by nn.Conv3d in the
// reference graph (reference model/my_slowfast.py:94-125 -> pytorchvideo create_slowfast).
#pragma once
#include "common.cuh"
#include "../../include/sfb200.h"
#include <cuda.h>

namespace sfb {

constexpr int kProducerThreads = 256;
constexpr int kTileM = 128;
constexpr int kChunkK = 64;                 // bf16 elements per K chunk = one 128-byte swizzle row
constexpr int kAStageBytes = kTileM * 128;  // 16 KB

struct ConvTcParams {
  SfbConvGeom g;
  const __nv_bfloat16* G;
  const float* pro_scale;
  const float* pro_shift;
  int pro_relu;
  __nv_bfloat16* out;
  const float* epi_scale;
  const float* epi_shift;
  const __nv_bfloat16* res;
  int res_ld, res_coff;
  int epi_relu;
  float* stats;
  int M;     // N*To*Ho*Wo
  int Ktot;  // taps*Cg
  int nk;    // k_pad / 64
};

// decomposed coordinates of one GEMM row (an output pixel), ready for tap offsets
struct RowCoord {
  int nT;  // n * Ti
  int t0, h0, w0;
};

__device__ __forceinline__ RowCoord make_row(const SfbConvGeom& g, int m, int M) {
  RowCoord r;
  if (m < M) {
    int wo = m % g.Wo;
    int q = m / g.Wo;
    int ho = q % g.Ho;
    q /= g.Ho;
    int to = q % g.To;
    int n = q / g.To;
    r.nT = n * g.Ti;
    if (!g.transposed) {
      r.t0 = to * g.sT - g.pT; r.h0 = ho * g.sH - g.pH; r.w0 = wo * g.sW - g.pW;
    } else {
      r.t0 = to + g.pT; r.h0 = ho + g.pH; r.w0 = wo + g.pW;
    }
  } else {
    r.nT = 0; r.t0 = -(1 << 28); r.h0 = 0; r.w0 = 0;
  }
  return r;
}

// pixel index (n,t,h,w flattened) of the gathered tensor for this row and tap, or -1 (padding / hole)
__device__ __forceinline__ int gather_pixel(const SfbConvGeom& g, const RowCoord& r, int kt, int kh, int kw,
                                            bool unit_stride) {
  int t, h, w;
  if (!g.transposed) {
    t = r.t0 + kt; h = r.h0 + kh; w = r.w0 + kw;
  } else {
    t = r.t0 - kt; h = r.h0 - kh; w = r.w0 - kw;
    if (!unit_stride) {
      if (t < 0 || h < 0 || w < 0) return -1;
      int tq = t / g.sT, hq = h / g.sH, wq = w / g.sW;
      if (tq * g.sT != t || hq * g.sH != h || wq * g.sW != w) return -1;
      t = tq; h = hq; w = wq;
    }
  }
  if ((unsigned)t >= (unsigned)g.Ti || (unsigned)h >= (unsigned)g.Hi || (unsigned)w >= (unsigned)g.Wi) return -1;
  return ((r.nT + t) * g.Hi + h) * g.Wi + w;
}

__device__ __forceinline__ uint4 apply_prologue(const uint4& raw, const float* sc, const float* sh, int relu) {
  float f[8];
  unpack8(raw, f);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    f[i] = fmaf(f[i], sc[i], sh[i]);
    if (relu) f[i] = fmaxf(f[i], 0.f);
  }
  return pack8(f);
}

// column sums over the 32 lanes of a warp: on return lane L holds sum over lanes of v[L]
__device__ __forceinline__ float warp_colsum32(float* v, int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      float keep = up ? v[i + off] : v[i];
      float send = up ? v[i] : v[i + off];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

// -------------------------------------------------------------------------------------------------
// fprop / dgrad
// -------------------------------------------------------------------------------------------------
template <int BN, int STAGES>
struct ConvTcSmem {
  static constexpr int kBStageBytes = BN * 128;
  static constexpr int kABytes = STAGES * kAStageBytes;
  static constexpr int kBBytes = STAGES * kBStageBytes;
  static constexpr int kBarOff = kABytes + kBBytes;             // full[STAGES], empty[STAGES], tmem_full
  static constexpr int kHolderOff = kBarOff + (2 * STAGES + 1) * 8;
  static constexpr int kStatsOff = kHolderOff + 8;               // float[2*BN]
  static constexpr int kProOff = kStatsOff + 2 * BN * 4;         // float scale[Cg], shift[Cg]
  static size_t bytes(int Cg, bool pro) { return 1024 + kProOff + (pro ? (size_t)Cg * 8 : 0); }
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kProducerThreads + 64, 1)
conv_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_w, const ConvTcParams p) {
  using L = ConvTcSmem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + L::kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(smem + L::kHolderOff);
  float* s_stats = reinterpret_cast<float*>(smem + L::kStatsOff);
  float* s_pro = reinterpret_cast<float*>(smem + L::kProOff);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int m0 = blockIdx.x * kTileM;
  const int n0 = blockIdx.y * BN;
  const SfbConvGeom& g = p.g;
  const bool has_pro = p.pro_scale != nullptr;
  constexpr uint32_t kTmemCols = BN < 32 ? 32 : BN;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], kProducerThreads + 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_fence_init();
  }
  if (warp == 8 && lane == 0) tma_prefetch_desc(&tmap_w);
  if (warp == 9) {
    tmem_alloc(tmem_holder, kTmemCols);
    tmem_relinquish();
  }
  for (int i = tid; i < 2 * BN; i += blockDim.x) s_stats[i] = 0.f;
  if (has_pro) {
    for (int i = tid; i < g.Cg; i += blockDim.x) {
      s_pro[i] = p.pro_scale[i];
      s_pro[g.Cg + i] = p.pro_shift[i];
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (tid < kProducerThreads) {
    // ===================== A-operand producers (im2col gather) =====================
    const int j = tid & 7;        // 16-byte granule within the 128-byte K row
    const int rbase = tid >> 3;   // rows rbase + 32*i
    const uint32_t sw_off = (uint32_t)((j ^ (rbase & 7)) << 4);
    const bool unit_stride = (g.sT == 1 && g.sH == 1 && g.sW == 1);
    const int khw = g.kH * g.kW;
    RowCoord rows[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) rows[i] = make_row(g, m0 + rbase + 32 * i, p.M);
    const __nv_bfloat16* gbase = p.G + g.g_coff;

    auto issue = [&](int kc, uint4* v, uint32_t& okmask, int& cch) {
      const int k = kc * kChunkK + j * 8;
      okmask = 0;
      cch = 0;
      if (k < p.Ktot) {
        const int tap = k / g.Cg;
        const int c = k - tap * g.Cg;
        const int kt = tap / khw;
        const int r2 = tap - kt * khw;
        const int kh = r2 / g.kW;
        const int kw = r2 - kh * g.kW;
        cch = c;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int pix = gather_pixel(g, rows[i], kt, kh, kw, unit_stride);
          if (pix >= 0) {
            v[i] = ldg_nc_v4(gbase + (size_t)pix * g.g_ld + c);
            okmask |= 1u << i;
          }
        }
      }
    };
    auto commit = [&](int kc, uint4* v, uint32_t okmask, int cch) {
      const int s = kc % STAGES;
      const uint32_t par = ((kc / STAGES) & 1) ^ 1;
      mbar_wait(&empty_bar[s], par, 1);
      const uint32_t abase = smem_u32(sA + s * kAStageBytes) + sw_off;
      float sc[8], sh[8];
      if (has_pro && okmask) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { sc[i] = s_pro[cch + i]; sh[i] = s_pro[g.Cg + cch + i]; }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 q = make_uint4(0, 0, 0, 0);
        if (okmask & (1u << i)) q = has_pro ? apply_prologue(v[i], sc, sh, p.pro_relu) : v[i];
        sts_v4(abase + (uint32_t)((rbase + 32 * i) * 128), q);
      }
      fence_proxy_async_smem();
      mbar_arrive(&full_bar[s]);
    };

    uint4 va[4], vb[4];
    uint32_t oka = 0, okb = 0;
    int ca = 0, cb = 0;
    issue(0, va, oka, ca);
    for (int kc = 0; kc < p.nk; kc += 2) {
      if (kc + 1 < p.nk) issue(kc + 1, vb, okb, cb);
      commit(kc, va, oka, ca);
      if (kc + 1 < p.nk) {
        if (kc + 2 < p.nk) issue(kc + 2, va, oka, ca);
        commit(kc + 1, vb, okb, cb);
      }
    }

    // ===================== epilogue (warps 0..3 own TMEM lanes 32*warp ..) =====================
    if (warp < 4) {
      mbar_wait(tmem_full, 0, 2);
      tc_fence_after();
      const int m = m0 + warp * 32 + lane;
      const bool mv = m < p.M;
      __nv_bfloat16* orow = p.out + (size_t)(mv ? m : 0) * g.o_ld + g.o_coff;
      const __nv_bfloat16* rrow = p.res ? p.res + (size_t)(mv ? m : 0) * p.res_ld + p.res_coff : nullptr;
#pragma unroll 1
      for (int cc = 0; cc < BN / 32; ++cc) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(cc * 32), v);
        tmem_ld_wait();
        const int nb = n0 + cc * 32;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int n = nb + q * 8;
          if (n < g.Co) {
            float* f = v + q * 8;
            if (p.epi_scale) {
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = fmaf(f[i], __ldg(p.epi_scale + n + i), __ldg(p.epi_shift + n + i));
            }
            if (rrow && mv) {
              float rf[8];
              unpack8(*reinterpret_cast<const uint4*>(rrow + n), rf);
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] += rf[i];
            }
            if (p.epi_relu) {
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = fmaxf(f[i], 0.f);
            }
            if (mv) *reinterpret_cast<uint4*>(orow + n) = pack8(f);
          }
        }
        if (p.stats) {
          float sq[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            v[i] = mv ? v[i] : 0.f;
            sq[i] = v[i] * v[i];
          }
          const float s1 = warp_colsum32(v, lane);
          const float s2 = warp_colsum32(sq, lane);
          atomicAdd(&s_stats[cc * 32 + lane], s1);
          atomicAdd(&s_stats[BN + cc * 32 + lane], s2);
        }
      }
      if (p.stats) {
        named_bar_sync(1, 128);
        for (int i = tid; i < BN; i += 128) {
          const int n = n0 + i;
          if (n < g.Co) {
            atomicAdd(p.stats + n, s_stats[i]);
            atomicAdd(p.stats + g.Co + n, s_stats[BN + i]);
          }
        }
      }
    }
  } else if (warp == 8) {
    // ===================== B-operand producer (TMA) =====================
    if (lane == 0) {
      for (int kc = 0; kc < p.nk; ++kc) {
        const int s = kc % STAGES;
        const uint32_t par = ((kc / STAGES) & 1) ^ 1;
        mbar_wait(&empty_bar[s], par, 3);
        mbar_arrive_expect_tx(&full_bar[s], (uint32_t)L::kBStageBytes);
        tma_load_2d(sB + s * L::kBStageBytes, &tmap_w, &full_bar[s], kc * kChunkK, n0);
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kTileM, BN, 0, 0);
      for (int kc = 0; kc < p.nk; ++kc) {
        const int s = kc % STAGES;
        const uint32_t par = (kc / STAGES) & 1;
        mbar_wait(&full_bar[s], par, 4);
        tc_fence_after();
        const uint64_t ad = umma_smem_desc_sw128(smem_u32(sA + s * kAStageBytes), 16, 1024);
        const uint64_t bd = umma_smem_desc_sw128(smem_u32(sB + s * L::kBStageBytes), 16, 1024);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          // advance 16 bf16 = 32 bytes inside the 128-byte swizzle row: +2 in the (addr>>4) field
          umma_bf16(tmem_base, ad + (uint64_t)(kk * 2), bd + (uint64_t)(kk * 2), idesc, (kc | kk) != 0 ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);
      }
      umma_commit(tmem_full);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, kTmemCols);
}

// -------------------------------------------------------------------------------------------------
// wgrad
// -------------------------------------------------------------------------------------------------
struct WgradTcParams {
  SfbConvGeom g;   // same geometry as the fprop of this layer (transposed = 0)
  const __nv_bfloat16* X;
  const float* pro_scale;
  const float* pro_shift;
  int pro_relu;
  const __nv_bfloat16* dY;
  int dy_ld, dy_coff;
  float* dW;  // [Co][Ci_real][taps]
  int Ci_real;
  int taps;
  int M;
  int Ktot;          // taps*Cg
  int chunks_total;  // ceil(M/64)
  int chunks_per_split;
};

constexpr int kWgBlockBytes = 64 * 128;  // [64 positions][64 channels] bf16
constexpr int kWgStages = 4;
constexpr int kWgMaxNB = 4;              // up to 256 (tap,ci) columns per CTA
constexpr int kWgStageBytes = (2 + kWgMaxNB) * kWgBlockBytes;  // 48 KB
constexpr int kWgBarOff = kWgStages * kWgStageBytes;
constexpr int kWgHolderOff = kWgBarOff + (2 * kWgStages + 1) * 8;
constexpr int kWgProOff = kWgHolderOff + 8;
inline size_t wgrad_smem_bytes(int Cg, bool pro) { return 1024 + kWgProOff + (pro ? (size_t)Cg * 8 : 0); }

__global__ void __launch_bounds__(kProducerThreads + 32, 1)
conv_wgrad_tc_kernel(const WgradTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kWgBarOff);
  uint64_t* empty_bar = full_bar + kWgStages;
  uint64_t* tmem_full = empty_bar + kWgStages;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(smem + kWgHolderOff);
  float* s_pro = reinterpret_cast<float*>(smem + kWgProOff);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const SfbConvGeom& g = p.g;
  const int kg0 = blockIdx.x * (kWgMaxNB * 64);
  const int co0 = blockIdx.y * 128;
  const int chunk_begin = blockIdx.z * p.chunks_per_split;
  const int chunk_end = min(p.chunks_total, chunk_begin + p.chunks_per_split);
  const int nchunks = chunk_end - chunk_begin;
  const int kpad = (p.Ktot + 63) & ~63;
  const int nb = min(kWgMaxNB, (kpad - kg0) / 64);  // live 64-column blocks of this CTA
  const bool has_pro = p.pro_scale != nullptr;
  constexpr uint32_t kTmemCols = 256;

  if (tid == 0) {
    for (int s = 0; s < kWgStages; ++s) {
      mbar_init(&full_bar[s], kProducerThreads);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_fence_init();
  }
  if (warp == 8) {
    tmem_alloc(tmem_holder, kTmemCols);
    tmem_relinquish();
  }
  if (has_pro) {
    for (int i = tid; i < g.Cg; i += blockDim.x) {
      s_pro[i] = p.pro_scale[i];
      s_pro[g.Cg + i] = p.pro_shift[i];
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (nchunks > 0) {
    if (tid < kProducerThreads) {
      const int j = tid & 7;
      const int rbase = tid >> 3;  // rows rbase, rbase+32 of each 64-row block
      const uint32_t sw_off = (uint32_t)((j ^ (rbase & 7)) << 4);
      const bool unit_stride = true;  // fprop geometry: never the transposed mapping
      const int khw = g.kH * g.kW;
      // per-thread column descriptors of the B blocks (fixed for the whole CTA)
      int bc[kWgMaxNB], bkt[kWgMaxNB], bkh[kWgMaxNB], bkw[kWgMaxNB];
#pragma unroll
      for (int b = 0; b < kWgMaxNB; ++b) {
        const int kg = kg0 + b * 64 + j * 8;
        if (b < nb && kg < p.Ktot) {
          const int tap = kg / g.Cg;
          bc[b] = kg - tap * g.Cg;
          bkt[b] = tap / khw;
          const int r2 = tap - bkt[b] * khw;
          bkh[b] = r2 / g.kW;
          bkw[b] = r2 - bkh[b] * g.kW;
        } else {
          bc[b] = -1; bkt[b] = 0; bkh[b] = 0; bkw[b] = 0;
        }
      }
      const __nv_bfloat16* xbase = p.X + g.g_coff;
      const __nv_bfloat16* dybase = p.dY + p.dy_coff;

      for (int ci = 0; ci < nchunks; ++ci) {
        const int mchunk = (chunk_begin + ci) * 64;
        const int s = ci % kWgStages;
        const uint32_t par = ((ci / kWgStages) & 1) ^ 1;
        uint4 va[2][2];
        uint4 vb[kWgMaxNB][2];
        uint32_t okb = 0;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int m = mchunk + rbase + 32 * i;
          // A blocks: dY[m][co0 + b*64 + j*8 ..]
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            const int co = co0 + b * 64 + j * 8;
            va[b][i] = make_uint4(0, 0, 0, 0);
            if (m < p.M && co < g.Co) va[b][i] = ldg_nc_v4(dybase + (size_t)m * p.dy_ld + co);
          }
          const RowCoord rc = make_row(g, m, p.M);
#pragma unroll
          for (int b = 0; b < kWgMaxNB; ++b) {
            if (bc[b] >= 0) {
              const int pix = gather_pixel(g, rc, bkt[b], bkh[b], bkw[b], unit_stride);
              if (pix >= 0) {
                vb[b][i] = ldg_nc_v4(xbase + (size_t)pix * g.g_ld + bc[b]);
                okb |= 1u << (b * 2 + i);
              }
            }
          }
        }
        mbar_wait(&empty_bar[s], par, 5);
        const uint32_t sbase = smem_u32(smem + s * kWgStageBytes) + sw_off;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const uint32_t roff = (uint32_t)((rbase + 32 * i) * 128);
#pragma unroll
          for (int b = 0; b < 2; ++b) sts_v4(sbase + b * kWgBlockBytes + roff, va[b][i]);
        }
#pragma unroll
        for (int b = 0; b < kWgMaxNB; ++b) {
          if (b < nb) {
            float sc[8], sh[8];
            if (has_pro && bc[b] >= 0) {
#pragma unroll
              for (int e = 0; e < 8; ++e) { sc[e] = s_pro[bc[b] + e]; sh[e] = s_pro[g.Cg + bc[b] + e]; }
            }
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              uint4 q = make_uint4(0, 0, 0, 0);
              if (okb & (1u << (b * 2 + i))) q = has_pro ? apply_prologue(vb[b][i], sc, sh, p.pro_relu) : vb[b][i];
              sts_v4(sbase + (2 + b) * kWgBlockBytes + (uint32_t)((rbase + 32 * i) * 128), q);
            }
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(&full_bar[s]);
      }

      // ---------------- epilogue: TMEM lane = co, columns = (tap,ci) ----------------
      if (warp < 4) {
        mbar_wait(tmem_full, 0, 6);
        tc_fence_after();
        const int co = co0 + warp * 32 + lane;
        const bool cov = co < g.Co;
        float* dwrow = p.dW + (size_t)(cov ? co : 0) * p.Ci_real * p.taps;
#pragma unroll 1
        for (int cc = 0; cc < nb * 2; ++cc) {
          float v[32];
          tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(cc * 32), v);
          tmem_ld_wait();
          const int kgb = kg0 + cc * 32;
          int tap = kgb / g.Cg;
          int c = kgb - tap * g.Cg;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (cov && tap < p.taps && c < p.Ci_real) atomicAdd(dwrow + (size_t)c * p.taps + tap, v[i]);
            if (++c == g.Cg) { c = 0; ++tap; }
          }
        }
      }
    } else {
      // ---------------- MMA issuer (warp 8) ----------------
      if (lane == 0) {
        const uint32_t idesc = umma_idesc_bf16(128, nb * 64, 1, 1);
        for (int ci = 0; ci < nchunks; ++ci) {
          const int s = ci % kWgStages;
          const uint32_t par = (ci / kWgStages) & 1;
          mbar_wait(&full_bar[s], par, 7);
          tc_fence_after();
          const uint32_t sbase = smem_u32(smem + s * kWgStageBytes);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            // 16 positions = 16 rows of 128 bytes = 2 swizzle atoms of 1024 bytes
            const uint64_t ad = umma_smem_desc_sw128(sbase + kk * 2048, kWgBlockBytes, 1024);
            const uint64_t bd = umma_smem_desc_sw128(sbase + 2 * kWgBlockBytes + kk * 2048, kWgBlockBytes, 1024);
            umma_bf16(tmem_base, ad, bd, idesc, (ci | kk) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);
        }
        umma_commit(tmem_full);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace sfb
