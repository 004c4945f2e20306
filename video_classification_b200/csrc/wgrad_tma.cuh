// wgrad_tma.cuh -- Conv3d weight gradient on tcgen05, operands fetched by TMA (sm_100a).
//
//   dW[co][(tap, c)] = sum over output pixels m of  dY[m][co] * X[in(m, tap)][c]
//
//   The reduction runs over pixels, so both MMA operands are "MN-major": a shared-memory row is one pixel, its 128 /
//   64 / 32 bytes are consecutive channels -- exactly what a TMA box of an NDHWC tensor produces.
//     A = dY patch  : [R pixels][128 co]  two SWIZZLE_128B blocks of 64 channels (TMA zero-fills co >= Co)
//     B = X  patch  : per unit u = (tap, 64-channel chunk) one shifted box [R pixels][box_k channels]
//                     (the same shifted-box im2col as conv_tma.cuh); UPC units side by side = N of the MMA
//   D[128 co][UPC*box_k] lives in TMEM for the whole kernel; a CTA walks its share of the pixel patches
//   (split over blockIdx.z), then adds its partial sums into the fp32 OIDHW gradient with atomics.
//   R <= 64 pixels per stage; stages are zeroed once so the rows a box never writes (R not a multiple of 16) add 0.
//
// Replaces cuDNN bwd-filter behind nn.Conv3d (reference model/my_slowfast.py:94-125 via autograd).
#pragma once
#include "conv_tma.cuh"

namespace sfb {

struct WgradTmaParams {
  int N, To, Ho, Wo;
  int Co, Cg, Ci_real, taps;
  int kT, kH, kW, sT, sH, sW, pT, pH, pW;
  int box_k, cchunks;          // channels per X box (64/32/16); 64-channel chunks per tap
  int n_units;                 // taps * cchunks
  int upc;                     // units per CTA (upc * box_k <= 256)
  int Wt, Ht, Tt, tilesW, tilesH, tilesT, total_patches;
  int patches_per_split;
  int mh;                      // 128-channel halves of dY per CTA: 2 when Co > 128 (two TMEM accumulators share the X boxes)
  float* dW;
  // pixel-folded conv (elementwise.cuh, fold_weight_kernel): the GEMM runs on the folded operands, the epilogue adds
  // each element straight into the REFERENCE-layout gradient (Co, Ci, kT, kH, kW) through the band map
  // kw = f*(kw' - kW'/2) + j' - j + kW/2 -- no folded gradient buffer, no unfold launch.  fold_f = 0: plain conv.
  int fold_f, fold_ldo, fold_ldi, fold_kW, fold_Co, fold_Ci;
};

constexpr int kWtStages = 4;
constexpr int kWtRows = 64;
constexpr int kWtABytes = 2 * kWtRows * 128;          // 16 KB: two [64 px][64 co] blocks
constexpr int kWtBBytes = 4 * kWtRows * 128;          // 32 KB: up to 256 columns
constexpr int kWtStage = kWtABytes + kWtBBytes;       // 48 KB
constexpr int kWtBarOff = kWtStages * kWtStage;
constexpr int kWtHolderOff = kWtBarOff + (2 * kWtStages + 1) * 8;
constexpr int kWtSmem = 1024 + kWtHolderOff + 16;

// MN-major operand descriptor: rows (K) are `span` bytes, 8-row groups SBO = 8*span apart, the next block of
// span/2 MN-elements is `lbo` bytes away.
__device__ __forceinline__ uint64_t umma_smem_desc_mnmajor(uint32_t smem_addr, uint32_t span, uint32_t lbo) {
  const uint64_t layout = span == 128 ? 2ull : (span == 64 ? 4ull : 6ull);
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(((8u * span) >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= layout << 61;
  return d;
}

template <int MH>
__global__ void __launch_bounds__(192, 1)
wgrad_tma_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_dy,
                 const WgradTmaParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kWtBarOff);
  uint64_t* empty_bar = full_bar + kWtStages;
  uint64_t* done_bar = empty_bar + kWtStages;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(smem + kWtHolderOff);

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx();
  const int lane = tid & 31;
  const int u0 = blockIdx.x * p.upc;                       // first unit of this CTA
  const int nu = min(p.upc, p.n_units - u0);               // live units
  // With one 128 x 256 accumulator a stage moves 48 KB for 4 MMAs (94 B/clk at the tensor floor, twice what TMA
  // sustains per SM); a second accumulator (co0+128..) reuses the 32 KB of X boxes: 64 KB for 8 MMAs.
  constexpr int mh = MH;
  constexpr int n_stages = MH == 2 ? 3 : kWtStages;
  constexpr uint32_t stage_bytes = MH == 2 ? 4 * kWtRows * 128 + kWtBBytes : kWtStage;   // 64 KB | 48 KB
  constexpr uint32_t a_bytes = (uint32_t)MH * kWtABytes;
  const int co0 = blockIdx.y * 128 * mh;
  const int patch_begin = blockIdx.z * p.patches_per_split;
  const int patch_end = min(p.total_patches, patch_begin + p.patches_per_split);
  const int rows = p.Wt * p.Ht * p.Tt;
  const int ksteps = (rows + 15) >> 4;
  const uint32_t span = (uint32_t)(p.box_k * 2);
  const uint32_t bblock = (uint32_t)kWtRows * span;        // bytes of one unit block
  const int nblk_a = min(2 * mh, (p.Co - co0 + 63) / 64);
  const int halves = MH == 2 ? (nblk_a + 1) >> 1 : 1;
  constexpr uint32_t kTmemCols = 256 * MH;

  if (tid == 0) {
    for (int s = 0; s < kWtStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(done_bar, 1);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_dy);
  }
  if (warp == 2) {
    tmem_alloc(tmem_holder, kTmemCols);
    tmem_relinquish();
  }
  // zero the operand stages once: rows a box never writes must contribute 0 (and never NaN)
  for (int i = tid; i < kWtStages * kWtStage / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = uniform_u32(*tmem_holder);

  if (patch_end > patch_begin) {
    if (warp == 0) {
      // TMA producer: warp-uniform loop (operands in uniform registers), one elected lane issues; taps and patches
      // advance with counters -- no division and no table lookup per box.
      const bool leader = elect_one();
      const uint32_t tx = (uint32_t)(nblk_a * rows * 128 + nu * rows * (int)span);
      const uint32_t smem0 = smem_u32(smem);
      const int khw = p.kH * p.kW;
      const int tap0 = u0 / p.cchunks, cc0 = u0 - tap0 * p.cchunks;
      const int kt0 = tap0 / khw, kh0 = (tap0 - kt0 * khw) / p.kW, kw0 = tap0 - kt0 * khw - kh0 * p.kW;
      int s = 0;
      uint32_t ph = 1;
      int q = patch_begin;
      int tw = q % p.tilesW; q /= p.tilesW;
      int th = q % p.tilesH; q /= p.tilesH;
      int tt = q % p.tilesT;
      int n = q / p.tilesT;
      for (int patch = patch_begin; patch < patch_end; ++patch) {
        const int w0 = tw * p.Wt, h0 = th * p.Ht, t0 = tt * p.Tt;
        mbar_wait(&empty_bar[s], ph, 21);
        const uint32_t sa = smem0 + (uint32_t)s * stage_bytes;
        if (leader) {
          mbar_arrive_expect_tx(&full_bar[s], tx);
          for (int b = 0; b < nblk_a; ++b)
            tma_load_5d(sa + b * (kWtRows * 128), &tmap_dy, &full_bar[s], co0 + b * 64, w0, h0, t0, n);
        }
        const uint32_t sb = sa + a_bytes;
        const int xw = w0 * p.sW - p.pW, xh = h0 * p.sH - p.pH, xt = t0 * p.sT - p.pT;
        int cc = cc0, kw = kw0, kh = kh0, kt = kt0;
        for (int u = 0; u < nu; ++u) {
          if (leader) tma_load_5d(sb + u * bblock, &tmap_x, &full_bar[s], cc * p.box_k, xw + kw, xh + kh, xt + kt, n);
          if (++cc == p.cchunks) { cc = 0; if (++kw == p.kW) { kw = 0; if (++kh == p.kH) { kh = 0; ++kt; } } }
        }
        __syncwarp();
        if (++s == n_stages) { s = 0; ph ^= 1; }
        if (++tw == p.tilesW) { tw = 0; if (++th == p.tilesH) { th = 0; if (++tt == p.tilesT) { tt = 0; ++n; } } }
      }
    } else if (warp == 1) {
      // MMA issuer: warp-uniform loop, one elected lane issues
      const bool leader = elect_one();
      const int ncols = ((nu * p.box_k + 15) >> 4) << 4;
      const uint32_t idesc = umma_idesc_bf16(128, ncols, 1, 1);
      const uint64_t adesc0 = umma_smem_desc_mnmajor(smem_u32(smem), 128, kWtRows * 128);
      const uint64_t bdesc0 = umma_smem_desc_mnmajor(smem_u32(smem) + a_bytes, span, bblock);
      const uint32_t bstep = (16 * span) >> 4;
      int s = 0;
      uint32_t ph = 0, first = 1;
      for (int patch = patch_begin; patch < patch_end; ++patch) {
        mbar_wait(&full_bar[s], ph, 22);
        tc_fence_after();
        uint64_t ad = adesc0 + (uint64_t)(((uint32_t)s * stage_bytes) >> 4);
        uint64_t bd = bdesc0 + (uint64_t)(((uint32_t)s * stage_bytes) >> 4);
        if (leader) {
          for (int kk = 0; kk < ksteps; ++kk) {
            umma_bf16(tmem_base, ad, bd, idesc, first ? 0u : 1u);
            if (MH == 2 && halves == 2) umma_bf16(tmem_base + 256, ad + (uint64_t)(kWtABytes >> 4), bd, idesc, first ? 0u : 1u);
            first = 0;
            ad += (16 * 128) >> 4;
            bd += bstep;
          }
          umma_commit(&empty_bar[s]);
        }
        first = 0;
        __syncwarp();
        if (++s == n_stages) { s = 0; ph ^= 1; }
      }
      if (leader) umma_commit(done_bar);
      __syncwarp();
    } else {
      // ---------------- epilogue: TMEM lane = co, columns = (unit, channel) ----------------
      // A thread reads one co row of the accumulator, but in the OIDHW gradient consecutive addresses run along
      // (channel, tap) of ONE co: adding straight from the TMEM layout makes every lane of an atomic hit its own
      // line (32 sectors per instruction; the tail of 65 k scattered atomics per CTA was longer than the MMA phase
      // of the small-map layers).  Each warp therefore transposes its 32 x 32 block through shared memory (the
      // operand stages are idle by now) and adds with lane = channel: `taps` floats apart, 1 (pointwise) to 9 lines.
      const int q4 = warp & 3;
      mbar_wait(done_bar, 0, 23);
      tc_fence_after();
      const int ncols = nu * p.box_k;
      float* scratch = reinterpret_cast<float*>(smem) + (warp - 2) * (32 * 33);
      const size_t co_stride = (size_t)p.Ci_real * p.taps;
#pragma unroll 1
      for (int hc = 0; hc * 32 < halves * 256; ++hc) {
        const int h = hc >> 3, cc = hc & 7;
        if (cc * 32 >= ncols) continue;
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(h * 256 + cc * 32), v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) scratch[lane * 33 + i] = v[i];
        __syncwarp();
        // this lane's column of the block -> (tap, channel) of the gradient
        const int col = cc * 32 + lane;
        const int u = col / p.box_k;
        const int unit = u0 + u;
        const int tap = unit / p.cchunks;
        const int c = (unit - tap * p.cchunks) * p.box_k + (col - u * p.box_k);
        const bool cv = col < ncols && c < p.Ci_real;
        const int co_b = co0 + h * 128 + q4 * 32;
        float* dst = p.dW + (size_t)co_b * co_stride + (size_t)(cv ? c : 0) * p.taps + tap;
        const int nrows = min(32, p.Co - co_b);
        if (p.fold_f > 1) {
          // folded column (j', c) at folded tap (kt, kh, kw'); rows are (j, co)
          const int kWf = p.fold_kW == 1 ? 1 : 3;
          const int jp = c / p.fold_ldi, cc = c - jp * p.fold_ldi;
          const int kwf = tap % kWf, tkh = tap / kWf;              // tkh = kt*kH + kh
          const int kw_base = p.fold_f * (kwf - kWf / 2) + jp + p.fold_kW / 2;
          const int tkh_n = p.taps / kWf;
          if (cv && cc < p.fold_Ci) {
            int j = co_b / p.fold_ldo, co = co_b - j * p.fold_ldo;
            float* base = p.dW + ((size_t)cc * tkh_n + tkh) * p.fold_kW;
            const size_t cstride = (size_t)p.fold_Ci * tkh_n * p.fold_kW;
            for (int rr = 0; rr < nrows; ++rr) {
              const int kw = kw_base - j;
              if (co < p.fold_Co && kw >= 0 && kw < p.fold_kW) atomicAdd(base + (size_t)co * cstride + kw, scratch[rr * 33 + lane]);
              if (++co == p.fold_ldo) { co = 0; ++j; }
            }
          }
        } else if (cv) {
#pragma unroll 4
          for (int rr = 0; rr < nrows; ++rr) atomicAdd(dst + (size_t)rr * co_stride, scratch[rr * 33 + lane]);
        }
        __syncwarp();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace sfb
