// stem_win.cuh -- the (1,7,7)/s(1,2,2)/p(0,3,3) stem convolution of ResNetBasicStem (reference
// model/my_slowfast.py:63-68 -> pytorchvideo create_res_basic_stem) as a "row window" implicit GEMM on sm_100a.
//
// Why a dedicated kernel: Cin = 5 / 15 and 49 taps.  Fetching every tap as its own TMA box reads the input 49x from L2
// in 32-byte rows and is bound by the TMA request rate (ncu: 3 % DRAM, 24 % L2, profiles/r1_stem_before.txt), 27x off
// the HBM roofline.  Here the input is stored space-to-depth ("s2d": each 2x2 pixel block becomes 4*Cin channels,
// padded to Cs = 32 or 64), which turns the conv into a STRIDE-1 4x4 conv over a (H/2, W/2) image with 64/128-byte
// pixels.  One M tile is one output row (<= 125 pixels).  The s2d input rows live in a shared-memory ring, one
// [128 pixels][Cs] TMA box per row (pixels -2 .. 125; out-of-range pixels and rows are TMA zero fill = the conv's
// padding), and each row is loaded ONCE and used by the four output rows (dh) it contributes to.  Tap (dh, dw) of
// the tile is the ring slot of input row ho-2+dh with the MMA descriptor's start address advanced by dw pixels:
// no im2col copy exists anywhere.
//
//   fprop : D[128 pixels][BN co]  += A(slot dh, shifted dw)[128][Cs] (K-major)  x  Wp[BN][(dh,dw,cs)] (K-major)
//   wgrad : D[(dw,cs) 128][BN co] += A(slot dh, shifted)[pixels][(dw,cs)] (MN-major; the M blocks of the descriptor
//           are one pixel apart = consecutive dw taps)  x  dY row [pixels][BN co] (MN-major), accumulated in TMEM
//           over every row the CTA owns, then added to the fp32 OIDHW gradient with the s2d -> (kh,kw,ci) mapping.
#pragma once
#include "conv_tma.cuh"
#include "wgrad_tma.cuh"

namespace sfb {

__device__ __forceinline__ void tma_load_4d(uint32_t smem_dst, const void* tmap, uint64_t* bar, int32_t c0, int32_t c1,
                                            int32_t c2, int32_t c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// descriptor "matrix base offset" (bits 49..51): phase of the swizzle pattern when the start address is not aligned
// to the pattern's repeat (8 rows); a window shifted by dw pixels starts dw rows into the pattern.
__device__ __forceinline__ uint64_t desc_base_offset(uint32_t smem_addr) { return (uint64_t)((smem_addr >> 7) & 7u) << 49; }

struct StemParams {
  int F, Ho, Wo;        // frames (N*T), s2d image = output image size
  int Cs;               // s2d channels: 32 or 64
  int Co, o_ld, o_coff; // real output channels; output / dY view
  int Ci;               // real input channels (wgrad mapping)
  int tiles_per_cta, total_tiles;
  int ksteps;           // wgrad: ceil(Wo/16)
  int base_off_mode;    // 1: set the descriptor base offset for shifted windows
  __nv_bfloat16* out;   // fprop
  float* stats;         // fprop: [2*Co] or null
  float* dW;            // wgrad: fp32 (Co, Ci, 7, 7)
};

constexpr int kStemSlots = 8;

// butterfly reduce-scatter of v[NC] over the 32 lanes of a warp: afterwards v[0..nf) hold full-warp sums of columns
// base .. base+nf-1 (nf = max(1, NC/32)); with NC < 32 groups of 32/NC adjacent lanes hold the same column.
template <int NC>
__device__ __forceinline__ int warp_reduce_columns(float* v, int lane) {
  int n = NC, base = 0;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    if (n > 1) {
      n >>= 1;
      const bool up = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < NC / 2; ++i) {
        if (i < n) {
          const float send = up ? v[i] : v[i + n];
          const float keep = up ? v[i + n] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
      }
      if (up) base += n;
    } else {
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
    }
  }
  return base;
}

template <int BN>
struct StemFpropCfg {
  static size_t smem_bytes(int Cs) {
    return 1024 + (size_t)kStemSlots * 128 * Cs * 2 + 1024 /*tail pad*/ + (size_t)BN * 16 * Cs * 2 + 256;
  }
};

// ---------------------------------------------------------------------------------------------------------------
// fprop
// ---------------------------------------------------------------------------------------------------------------
template <int BN, int CS>
__global__ void __launch_bounds__(192, 1)
stem_fprop_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                  const StemParams p) {
  constexpr int NS = kStemSlots;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t span = (uint32_t)CS * 2;
  constexpr uint32_t slot_bytes = 128u * span;
  constexpr uint32_t ring_bytes = NS * slot_bytes;
  uint8_t* bsm = smem + ring_bytes + 1024;
  constexpr int Ktot = 16 * CS;
  constexpr int nk = Ktot / 64;
  uint64_t* bars = reinterpret_cast<uint64_t*>(bsm + (size_t)BN * Ktot * 2);
  uint64_t* full_bar = bars;            // [NS]
  uint64_t* empty_bar = bars + NS;      // [NS]
  uint64_t* tfull_bar = bars + 2 * NS;  // [2]
  uint64_t* tempty_bar = tfull_bar + 2; // [2]
  uint64_t* b_bar = tempty_bar + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(b_bar + 1);

  const int tid = threadIdx.x, warp = uniform_warp_idx(), lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 4); }
    mbar_init(b_bar, 1);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmap_x); tma_prefetch_desc(&tmap_w); }
  constexpr uint32_t kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;
  if (warp == 2) { tmem_alloc(tmem_holder, kTmemCols); tmem_relinquish(); }
  // zero the ring and its tail pad once: shifted windows read up to 3 pixels past a slot
  for (uint32_t i = tid; i < (ring_bytes + 1024) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = uniform_u32(*tmem_holder);
  const int g0 = blockIdx.x * p.tiles_per_cta;
  const int g1 = min(p.total_tiles, g0 + p.tiles_per_cta);
  const uint32_t ring0 = smem_u32(smem);

  if (warp == 0) {
    // TMA producer: the loop runs warp-uniform, one elected lane issues (see common.cuh)
    if (g0 < g1) {
      const bool leader = elect_one();
      if (leader) {
        mbar_arrive_expect_tx(b_bar, (uint32_t)(BN * Ktot * 2));
        for (int kc = 0; kc < nk; ++kc) tma_load_2d_u32(smem_u32(bsm) + kc * (BN * 128), &tmap_w, b_bar, kc * 64, 0);
      }
      uint32_t L = 0;
      int frame = g0 / p.Ho, ho = g0 - frame * p.Ho;
      bool fresh = true;  // first tile of a frame (or of the CTA): the whole 4-row window is loaded
      for (int g = g0; g < g1; ++g) {
        for (int r = fresh ? ho - 2 : ho + 1; r <= ho + 1; ++r, ++L) {
          const int s = L % NS;
          mbar_wait(&empty_bar[s], ((L / NS) & 1) ^ 1, 31);
          if (leader) {
            mbar_arrive_expect_tx(&full_bar[s], slot_bytes);
            tma_load_4d(ring0 + s * slot_bytes, &tmap_x, &full_bar[s], 0, -2, r, frame);
          }
          __syncwarp();
        }
        fresh = false;
        if (++ho == p.Ho) { ho = 0; ++frame; fresh = true; }
      }
    }
  } else if (warp == 1) {
    // MMA issuer: warp-uniform loop, one elected lane issues
    if (g0 < g1) {
      const bool leader = elect_one();
      const uint32_t idesc = umma_idesc_bf16(128, BN, 0, 0);
      constexpr int KS = CS / 16;
      // 16*KS MMAs per output row: descriptors are a constant plus a compile-time offset in the 14-bit start-address
      // field (everything lives below 256 KB, no carry), i.e. one integer add per operand.
      const uint64_t adesc0 = umma_smem_desc_kmajor(ring0, span);
      const uint64_t bdesc0 = umma_smem_desc_kmajor(smem_u32(bsm), 128);
      mbar_wait(b_bar, 0, 32);
      uint32_t L = 0;
      int ho = g0 % p.Ho, it = 0;
      bool fresh = true;
      for (int g = g0; g < g1; ++g, ++it) {
        L += fresh ? 4 : 1;
        const int buf = it & 1;
        mbar_wait(&tempty_bar[buf], (uint32_t)(((it >> 1) & 1) ^ 1), 33);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BN);
#pragma unroll
        for (int dh = 0; dh < 4; ++dh) {
          const uint32_t l = L - 4 + dh;
          const int s = l % NS;
          mbar_wait(&full_bar[s], (l / NS) & 1, 34);
          tc_fence_after();
          const uint64_t ad_slot = adesc0 + (uint64_t)((s * slot_bytes) >> 4);
          if (leader) {
#pragma unroll
            for (int dw = 0; dw < 4; ++dw) {
#pragma unroll
              for (int ks = 0; ks < KS; ++ks) {
                const int kidx = (dh * 4 + dw) * CS + ks * 16;
                const uint32_t b_off = (uint32_t)((kidx >> 6) * (BN * 128) + (kidx & 63) * 2);
                const uint64_t ad = ad_slot + (uint64_t)((dw * span + ks * 32) >> 4);
                const uint64_t bd = bdesc0 + (uint64_t)(b_off >> 4);
                umma_bf16(d_tmem, ad, bd, idesc, (dh | dw | ks) != 0 ? 1u : 0u);
              }
            }
          }
          __syncwarp();
        }
        fresh = false;
        if (++ho == p.Ho) { ho = 0; fresh = true; }
        const bool last_of_run = (g + 1 == g1) || fresh;
        if (leader) {
          if (last_of_run) {
            for (int dh = 0; dh < 4; ++dh) umma_commit(&empty_bar[(L - 4 + dh) % NS]);
          } else {
            umma_commit(&empty_bar[(L - 4) % NS]);
          }
          umma_commit(&tfull_bar[buf]);
        }
        __syncwarp();
      }
    }
  } else {
    const int q4 = warp & 3;
    const int r = q4 * 32 + lane;
    const bool valid = r < p.Wo;
    float acc_s[BN >= 32 ? BN / 32 : 1], acc_q[BN >= 32 ? BN / 32 : 1];
#pragma unroll
    for (int i = 0; i < (BN >= 32 ? BN / 32 : 1); ++i) acc_s[i] = acc_q[i] = 0.f;
    int col_base = 0, it = 0;
    for (int g = g0; g < g1; ++g, ++it) {
      const int buf = it & 1;
      mbar_wait(&tfull_bar[buf], (uint32_t)((it >> 1) & 1), 35);
      tc_fence_after();
      float v[BN];
      const uint32_t ta = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(buf * BN);
      if constexpr (BN == 16) {
        tmem_ld16(ta, v);
      } else {
#pragma unroll
        for (int c = 0; c < BN / 32; ++c) tmem_ld32(ta + c * 32, v + c * 32);
      }
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[buf]);
      __nv_bfloat16* orow = p.out + ((size_t)g * p.Wo + (valid ? r : 0)) * p.o_ld + p.o_coff;
      float sq[BN];
#pragma unroll
      for (int j = 0; j < BN / 8; ++j) {
        uint4 pk = pack8(v + j * 8);
        if (valid && j * 8 < p.Co) *reinterpret_cast<uint4*>(orow + j * 8) = pk;
        // BatchNorm statistics are taken from the bf16-rounded values the consumer will read
        float f[8];
        unpack8(pk, f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float x = valid ? f[i] : 0.f;
          v[j * 8 + i] = x;
          sq[j * 8 + i] = x * x;
        }
      }
      if (p.stats) {
        col_base = warp_reduce_columns<BN>(v, lane);
        warp_reduce_columns<BN>(sq, lane);
#pragma unroll
        for (int i = 0; i < (BN >= 32 ? BN / 32 : 1); ++i) { acc_s[i] += v[i]; acc_q[i] += sq[i]; }
      }
    }
    if (p.stats && g0 < g1) {
      constexpr int NF = BN >= 32 ? BN / 32 : 1;
      constexpr int REP = BN >= 32 ? 1 : 32 / BN;
      if ((lane & (REP - 1)) == 0) {
#pragma unroll
        for (int i = 0; i < NF; ++i) {
          const int col = col_base + i;
          if (col < p.Co) {
            atomicAdd(p.stats + col, acc_s[i]);
            atomicAdd(p.stats + p.Co + col, acc_q[i]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------------------------------------------------------
// fprop, four output rows per tile (narrow Co)
// ---------------------------------------------------------------------------------------------------------------
// With Co = 8 the one-row kernel issues N = 16 MMAs (half channel padding) that each re-read a 4 KB A window: it is
// bound by that shared-memory operand traffic.  Here a tile is FOUR output rows h0..h0+3 whose channels sit side by
// side in N (N = 4*BR, no padding for Co = 8): s2d input row h0-2+j, j = 0..6, is multiplied once per (dw, k-step)
// against the weights of all output rows it feeds.  Output row h0+p takes W[dh = j-p] from input row j; with the
// column blocks in reverse row order (block p' <-> row h0+3-p') that is the window of blocks j .. j+3 of the sequence
// [0,0,0,W[dh=0],W[1],W[2],W[3],0,0,0] -- one resident operand, the descriptor start moves by j blocks (whole 8-row
// swizzle atoms).  28 MMAs per output row instead of 64.
constexpr int kStemRowsF = 4;
constexpr int kStemRowsSlots = 11;  // 7 rows of the current tile + the 4 new rows of the next

template <int BR>  // channels per row block: 8 or 16
struct StemFpropRowsCfg {
  static size_t smem_bytes(int Cs) {
    return 1024 + (size_t)kStemRowsSlots * 128 * Cs * 2 + 1024 + (size_t)(4 * Cs / 64) * 10 * BR * 128 + 256;
  }
};

template <int BR, int CS>
__global__ void __launch_bounds__(192, 1)
stem_fprop_rows_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                       const StemParams p) {
  constexpr int NS = kStemRowsSlots;
  constexpr int F = kStemRowsF;
  constexpr int NW = F * BR;                       // accumulator width
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t span = (uint32_t)CS * 2;
  constexpr uint32_t slot_bytes = 128u * span;
  constexpr uint32_t ring_bytes = NS * slot_bytes;
  constexpr int NKC = 4 * CS / 64;                 // 64-wide K chunks of one input row (K = 4 dw * CS)
  constexpr uint32_t chunk_bytes = 10u * BR * 128u;  // [10 blocks * BR rows][128 B]
  uint8_t* bsm = smem + ring_bytes + 1024;
  uint64_t* bars = reinterpret_cast<uint64_t*>(bsm + NKC * chunk_bytes);
  uint64_t* full_bar = bars;            // [NS]
  uint64_t* empty_bar = bars + NS;      // [NS]
  uint64_t* tfull_bar = bars + 2 * NS;  // [2]
  uint64_t* tempty_bar = tfull_bar + 2; // [2]
  uint64_t* b_bar = tempty_bar + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(b_bar + 1);

  const int tid = threadIdx.x, warp = uniform_warp_idx(), lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 4); }
    mbar_init(b_bar, 1);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmap_x); tma_prefetch_desc(&tmap_w); }
  constexpr uint32_t kTmemCols = 2 * NW < 32 ? 32 : (2 * NW <= 64 ? 64 : 128);
  if (warp == 2) { tmem_alloc(tmem_holder, kTmemCols); tmem_relinquish(); }
  // zero the ring, its tail pad and the weight sequence (its first and last three blocks stay zero)
  for (uint32_t i = tid; i < (ring_bytes + 1024 + NKC * chunk_bytes) / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = uniform_u32(*tmem_holder);
  const int tpf = (p.Ho + F - 1) / F;              // tiles per frame
  const int g0 = blockIdx.x * p.tiles_per_cta;
  const int g1 = min(p.total_tiles, g0 + p.tiles_per_cta);
  const uint32_t ring0 = smem_u32(smem);

  if (warp == 0) {
    if (g0 < g1) {
      const bool leader = elect_one();
      if (leader) {
        // W[dh] of the packed operand [rows][(dh,dw,cs)] -> block dh+3 of every K chunk
        mbar_arrive_expect_tx(b_bar, (uint32_t)(4 * NKC * BR * 128));
        for (int dh = 0; dh < 4; ++dh)
          for (int kc = 0; kc < NKC; ++kc)
            tma_load_2d_u32(smem_u32(bsm) + kc * chunk_bytes + (dh + 3) * (BR * 128), &tmap_w, b_bar,
                            dh * 4 * CS + kc * 64, 0);
      }
      uint32_t L = 0;
      int frame = g0 / tpf, t = g0 - frame * tpf;
      bool fresh = true;
      for (int g = g0; g < g1; ++g) {
        const int h0 = t * F;
        for (int r = fresh ? h0 - 2 : h0 + 1; r <= h0 + F; ++r, ++L) {
          const int s = L % NS;
          mbar_wait(&empty_bar[s], ((L / NS) & 1) ^ 1, 31);
          if (leader) {
            mbar_arrive_expect_tx(&full_bar[s], slot_bytes);
            tma_load_4d(ring0 + s * slot_bytes, &tmap_x, &full_bar[s], 0, -2, r, frame);
          }
          __syncwarp();
        }
        fresh = false;
        if (++t == tpf) { t = 0; ++frame; fresh = true; }
      }
    }
  } else if (warp == 1) {
    if (g0 < g1) {
      const bool leader = elect_one();
      const uint32_t idesc = umma_idesc_bf16(128, NW, 0, 0);
      constexpr int KS = CS / 16;
      const uint64_t adesc0 = umma_smem_desc_kmajor(ring0, span);
      const uint64_t bdesc0 = umma_smem_desc_kmajor(smem_u32(bsm), 128);
      mbar_wait(b_bar, 0, 32);
      uint32_t L = 0;
      int t = g0 % tpf, it = 0;
      bool fresh = true;
      for (int g = g0; g < g1; ++g, ++it) {
        L += fresh ? F + 3 : F;
        const int buf = it & 1;
        mbar_wait(&tempty_bar[buf], (uint32_t)(((it >> 1) & 1) ^ 1), 33);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(buf * NW);
#pragma unroll
        for (int j = 0; j < F + 3; ++j) {
          const uint32_t l = L - (F + 3) + j;
          const int s = l % NS;
          mbar_wait(&full_bar[s], (l / NS) & 1, 34);
          tc_fence_after();
          const uint64_t ad_slot = adesc0 + (uint64_t)((s * slot_bytes) >> 4);
          if (leader) {
#pragma unroll
            for (int dw = 0; dw < 4; ++dw) {
#pragma unroll
              for (int ks = 0; ks < KS; ++ks) {
                const int kidx = dw * CS + ks * 16;
                const uint32_t b_off = (uint32_t)((kidx >> 6) * chunk_bytes + j * (BR * 128) + (kidx & 63) * 2);
                const uint64_t ad = ad_slot + (uint64_t)((dw * span + ks * 32) >> 4);
                const uint64_t bd = bdesc0 + (uint64_t)(b_off >> 4);
                umma_bf16(d_tmem, ad, bd, idesc, (j | dw | ks) != 0 ? 1u : 0u);
              }
            }
          }
          __syncwarp();
        }
        fresh = false;
        if (++t == tpf) { t = 0; fresh = true; }
        const bool last_of_run = (g + 1 == g1) || fresh;
        if (leader) {
          const int nfree = last_of_run ? F + 3 : F;  // the last three rows also serve the next tile of the frame
          for (int j = 0; j < nfree; ++j) umma_commit(&empty_bar[(L - (F + 3) + j) % NS]);
          umma_commit(&tfull_bar[buf]);
        }
        __syncwarp();
      }
    }
  } else {
    const int q4 = warp & 3;
    const int r = q4 * 32 + lane;        // pixel of the row
    const bool pv = r < p.Wo;
    float acc_s = 0.f, acc_q = 0.f;
    int col_base = 0, it = 0;
    int frame = g0 / tpf, t = g0 - frame * tpf;
    for (int g = g0; g < g1; ++g, ++it) {
      const int h0 = t * F;
      const int buf = it & 1;
      mbar_wait(&tfull_bar[buf], (uint32_t)((it >> 1) & 1), 35);
      tc_fence_after();
      float v[NW];
      const uint32_t ta = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(buf * NW);
      if constexpr (NW == 32) {
        tmem_ld32(ta, v);
      } else {
        tmem_ld32(ta, v);
        tmem_ld32(ta + 32, v + 32);
      }
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[buf]);
      float sv[BR], sq[BR];
#pragma unroll
      for (int i = 0; i < BR; ++i) { sv[i] = 0.f; sq[i] = 0.f; }
#pragma unroll
      for (int pb = 0; pb < F; ++pb) {           // column block pb <-> output row h0 + 3 - pb
        const int row = h0 + (F - 1) - pb;
        const bool valid = pv && row < p.Ho;
        __nv_bfloat16* orow = p.out + (((size_t)frame * p.Ho + (valid ? row : 0)) * p.Wo + (valid ? r : 0)) * p.o_ld + p.o_coff;
#pragma unroll
        for (int jv = 0; jv < BR / 8; ++jv) {
          uint4 pk = pack8(v + pb * BR + jv * 8);
          if (valid && jv * 8 < p.Co) *reinterpret_cast<uint4*>(orow + jv * 8) = pk;
          float f[8];
          unpack8(pk, f);  // statistics from the bf16-rounded values the consumer will read
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float x = valid ? f[i] : 0.f;
            sv[jv * 8 + i] += x;
            sq[jv * 8 + i] = fmaf(x, x, sq[jv * 8 + i]);
          }
        }
      }
      if (p.stats) {
        col_base = warp_reduce_columns<BR>(sv, lane);
        warp_reduce_columns<BR>(sq, lane);
        acc_s += sv[0];
        acc_q += sq[0];
      }
      if (++t == tpf) { t = 0; ++frame; }
    }
    if (p.stats && g0 < g1) {
      constexpr int REP = 32 / BR;
      if ((lane & (REP - 1)) == 0 && col_base < p.Co) {
        atomicAdd(p.stats + col_base, acc_s);
        atomicAdd(p.stats + p.Co + col_base, acc_q);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------------------------------------------------------
// wgrad
// ---------------------------------------------------------------------------------------------------------------
// wgrad.  The loop runs over INPUT rows: s2d row r contributes to the four output rows ho = r-1 .. r+2 (dh = r+2-ho), so
// its [pixels][(dw,cs)] window is multiplied ONCE against the four dY rows placed side by side in the MMA's N
// dimension -- N = 4*BN, accumulator column block i <-> dY row r-1+i <-> dh = 3-i -- instead of four N = BN MMAs
// that each re-read the same 4 KB A window from shared memory (the kernel was bound by exactly that traffic, and half
// of every N = 16 tile was channel padding).  The dY rows live in a ring whose first three slots are mirrored behind
// its end, so any four consecutive rows are contiguous for the descriptor (next N block = next slot).  dY rows that
// belong to another CTA's range (or lie outside the frame) are fetched with an out-of-range row coordinate, i.e. as
// TMA zero fill.
template <int BN>
struct StemDyRing {
  static constexpr int NB = BN == 64 ? 5 : 8;   // ring slots (+3 mirror slots)
  static constexpr int kSlots = NB + 3;
};

template <int BN, int CS>  // BN = padded Co: 16, 32 or 64
__global__ void __launch_bounds__(192, 1)
stem_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_dy,
                  const StemParams p) {
  constexpr int NS = kStemSlots;
  constexpr int NB = StemDyRing<BN>::NB;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t span = (uint32_t)CS * 2;
  constexpr uint32_t slot_bytes = 128u * span;
  constexpr uint32_t ring_bytes = NS * slot_bytes;
  constexpr uint32_t span_b = BN * 2;
  constexpr uint32_t bslot_bytes = 128u * span_b;
  constexpr uint32_t bring_bytes = StemDyRing<BN>::kSlots * bslot_bytes;
  uint8_t* bsm = smem + ring_bytes + 1024;
  uint64_t* bars = reinterpret_cast<uint64_t*>(bsm + bring_bytes);
  uint64_t* full_bar = bars;                 // [NS]
  uint64_t* empty_bar = bars + NS;           // [NS]
  uint64_t* bfull_bar = bars + 2 * NS;       // [NB]
  uint64_t* bempty_bar = bfull_bar + NB;     // [NB]
  uint64_t* done_bar = bempty_bar + NB;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(done_bar + 1);

  const int tid = threadIdx.x, warp = uniform_warp_idx(), lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < NB; ++s) { mbar_init(&bfull_bar[s], 1); mbar_init(&bempty_bar[s], 1); }
    mbar_init(done_bar, 1);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmap_x); tma_prefetch_desc(&tmap_dy); }
  constexpr int G = CS / 32;                 // M groups: 4 dw taps * Cs / 128
  constexpr int bpg = 4 / G;                 // dw taps (= M blocks of Cs channels) per group
  constexpr int NW = 4 * BN;                 // accumulator width: four dY rows
  constexpr uint32_t kTmemCols = G * NW <= 64 ? 64 : (G * NW <= 128 ? 128 : 256);
  if (warp == 2) { tmem_alloc(tmem_holder, kTmemCols); tmem_relinquish(); }
  for (uint32_t i = tid; i < (ring_bytes + 1024 + bring_bytes) / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = uniform_u32(*tmem_holder);
  const int g0 = blockIdx.x * p.tiles_per_cta;
  const int g1 = min(p.total_tiles, g0 + p.tiles_per_cta);
  const uint32_t ring0 = smem_u32(smem);
  const uint32_t bring0 = smem_u32(bsm);

  if (g0 < g1) {
    if (warp == 0) {
      // TMA producer (warp-uniform loop, one elected lane issues).  A frame segment [ha, hb) of this CTA's output
      // rows needs X rows ha-2+j, j = 0 .. n+2, and dY positions q = 0 .. n+5 (row ha-3+q; only ha <= row < hb is real)
      const bool leader = elect_one();
      uint32_t L = 0, Q = 0;  // running X-row / dY-position counters (ring slot = counter % slots)
      for (int g = g0; g < g1;) {
        const int frame = g / p.Ho, ha = g - frame * p.Ho;
        const int hb = min(p.Ho, ha + (g1 - g));
        const int n = hb - ha;
        int q = 0;
        for (int j = 0; j < n + 3; ++j, ++L) {
          for (; q <= j + 3; ++q, ++Q) {
            const int sb = Q % NB;
            mbar_wait(&bempty_bar[sb], ((Q / NB) & 1) ^ 1, 42);
            if (leader) {
              const int row = ha - 3 + q;
              const int rc = (row >= ha && row < hb) ? row : -8;  // foreign rows arrive as zeros
              mbar_arrive_expect_tx(&bfull_bar[sb], sb < 3 ? 2 * bslot_bytes : bslot_bytes);
              tma_load_4d(bring0 + sb * bslot_bytes, &tmap_dy, &bfull_bar[sb], 0, 0, rc, frame);
              if (sb < 3) tma_load_4d(bring0 + (sb + NB) * bslot_bytes, &tmap_dy, &bfull_bar[sb], 0, 0, rc, frame);
            }
            __syncwarp();
          }
          const int s = L % NS;
          mbar_wait(&empty_bar[s], ((L / NS) & 1) ^ 1, 41);
          if (leader) {
            mbar_arrive_expect_tx(&full_bar[s], slot_bytes);
            tma_load_4d(ring0 + s * slot_bytes, &tmap_x, &full_bar[s], 0, -2, ha - 2 + j, frame);
          }
          __syncwarp();
        }
        g += n;
      }
    } else if (warp == 1) {
      // MMA issuer: per X row, G * ksteps MMAs of N = 4*BN
      const bool leader = elect_one();
      const uint32_t idesc = umma_idesc_bf16(128, NW, 1, 1);
      const uint64_t adesc0 = umma_smem_desc_mnmajor(ring0, span, span);       // next M block = next pixel (dw tap)
      const uint64_t bdesc0 = umma_smem_desc_mnmajor(bring0, span_b, bslot_bytes);  // next N block = next dY row
      uint32_t L = 0, Q = 0, Qw = 0;  // Qw: dY positions already waited for
      for (int g = g0; g < g1;) {
        const int frame = g / p.Ho, ha = g - frame * p.Ho;
        const int hb = min(p.Ho, ha + (g1 - g));
        const int n = hb - ha;
        for (int j = 0; j < n + 3; ++j, ++L) {
          const uint32_t qj = Q + (uint32_t)j;  // first dY position of this row's window
          for (; Qw <= qj + 3; ++Qw) mbar_wait(&bfull_bar[Qw % NB], (Qw / NB) & 1, 43);
          const int s = L % NS;
          mbar_wait(&full_bar[s], (L / NS) & 1, 44);
          tc_fence_after();
          const uint64_t ad_slot = adesc0 + (uint64_t)((s * slot_bytes) >> 4);
          const uint64_t bd_slot = bdesc0 + (uint64_t)(((qj % NB) * bslot_bytes) >> 4);
          if (leader) {
#pragma unroll
            for (int gi = 0; gi < G; ++gi) {
              const uint32_t d_tmem = tmem_base + (uint32_t)(gi * NW);
              uint64_t ad = ad_slot + (uint64_t)((gi * bpg * span) >> 4);
              uint64_t bd = bd_slot;
              for (int k = 0; k < p.ksteps; ++k) {
                umma_bf16(d_tmem, ad, bd, idesc, (L | (uint32_t)k) != 0 ? 1u : 0u);
                ad += (uint64_t)((16 * span) >> 4);
                bd += (uint64_t)((16 * span_b) >> 4);
              }
            }
            umma_commit(&empty_bar[s]);
            umma_commit(&bempty_bar[qj % NB]);  // dY position qj is not used by later rows
          }
          __syncwarp();
        }
        // positions n+3 .. n+5 of this segment were only needed as window tails: free them as well
        if (leader) {
          for (uint32_t t = 0; t < 3; ++t) umma_commit(&bempty_bar[(Q + (uint32_t)n + 3 + t) % NB]);
        }
        __syncwarp();
        Q += (uint32_t)n + 6;
        g += n;
      }
      if (leader) umma_commit(done_bar);
      __syncwarp();
    } else {
      // epilogue: TMEM lane m = (dw block j, s2d channel cs); column block i = dY row r-1+i <-> dh = 3-i
      const int q4 = warp & 3;
      mbar_wait(done_bar, 0, 45);
      tc_fence_after();
      const int m = q4 * 32 + lane;
      const int j = m / CS, cs = m - j * CS;
      const int quad = cs / p.Ci, ci = cs - quad * p.Ci;   // s2d channel -> (sub-pixel, input channel)
      const int sh = quad >> 1, sw = quad & 1;
      for (int i = 0; i < 4; ++i) {
        const int dh = 3 - i;
        for (int gi = 0; gi < G; ++gi) {
          float v[BN];
          const uint32_t ta = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(gi * NW + i * BN);
          if constexpr (BN == 16) {
            tmem_ld16(ta, v);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 32; ++c) tmem_ld32(ta + c * 32, v + c * 32);
          }
          tmem_ld_wait();
          const int dw = gi * bpg + j;
          const int kh = 2 * dh + sh - 1, kw = 2 * dw + sw - 1;
          if (quad < 4 && kh >= 0 && kw >= 0 && kh < 7 && kw < 7) {
#pragma unroll
            for (int co = 0; co < BN; ++co)
              if (co < p.Co) atomicAdd(p.dW + (((size_t)co * p.Ci + ci) * 7 + kh) * 7 + kw, v[co]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// ---- s2d input layout: fp32 strided NCTHW -> bf16 [N][T][H/2][W/2][Cs], channel = (sh*2+sw)*C + c ----------------
__global__ void layout_input_s2d_kernel(const float* __restrict__ x, long long sn, long long sc, long long st,
                                        long long sh, long long sw, int N, int C, int T, int H2, int W2,
                                        __nv_bfloat16* __restrict__ out, int Cs) {
  // one thread per s2d pixel: for a fixed (channel, sub-pixel) the lanes of a warp read consecutive W positions of one
  // fp32 plane (coalesced through L1; the other sub-pixel of the pair hits the same lines), and write 16 B each
  const int CV = Cs >> 3;
  const long long total = (long long)N * T * H2 * W2;
  for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < total;
       pix += (long long)gridDim.x * blockDim.x) {
    const int w = (int)(pix % W2);
    long long q = pix / W2;
    const int h = (int)(q % H2);
    q /= H2;
    const int t = (int)(q % T);
    const int n = (int)(q / T);
    const float* src = x + n * sn + t * st + (long long)(2 * h) * sh + (long long)(2 * w) * sw;
    __nv_bfloat16* dst = out + (size_t)pix * Cs;
    int quad = 0, c = 0;
    for (int cv = 0; cv < CV; ++cv) {
      float f[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        f[i] = quad < 4 ? __ldg(src + (long long)c * sc + (long long)(quad >> 1) * sh + (long long)(quad & 1) * sw) : 0.f;
        if (++c == C) { c = 0; ++quad; }
      }
      *reinterpret_cast<uint4*>(dst + cv * 8) = pack8(f);
    }
  }
}

// Same layout for W-contiguous input (stride 1 along W, 8-byte aligned rows) and a compile-time channel count: the
// two sub-pixels of a row come in one 8-byte load, all 2*C loads of a pixel are independent and issued back to back,
// and the channel shuffle is resolved at compile time (registers only).
template <int C, int CS>
__global__ void __launch_bounds__(128)
layout_input_s2d_vec_kernel(const float* __restrict__ x, long long sn, long long sc, long long st, long long sh,
                            int N, int T, int H2, int W2, __nv_bfloat16* __restrict__ out) {
  __shared__ uint4 tile[4][32 * (CS / 8)];
  const long long total = (long long)N * T * H2 * W2;
  // whole warps iterate together (the tail warp keeps its out-of-range lanes for the shared-memory exchange)
  for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix - (threadIdx.x & 31) < total;
       pix += (long long)gridDim.x * blockDim.x) {
    const long long pc = pix < total ? pix : total - 1;
    const int w = (int)(pc % W2);
    long long q = pc / W2;
    const int h = (int)(q % H2);
    q /= H2;
    const int t = (int)(q % T);
    const int n = (int)(q / T);
    const float* src = x + n * sn + t * st + (long long)(2 * h) * sh + 2 * w;
    float f[CS];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const float2 v = __ldg(reinterpret_cast<const float2*>(src + (long long)c * sc + (long long)r * sh));
        f[(r * 2 + 0) * C + c] = v.x;
        f[(r * 2 + 1) * C + c] = v.y;
      }
    }
#pragma unroll
    for (int i = 4 * C; i < CS; ++i) f[i] = 0.f;
    // A thread owns one pixel = KV 16-byte chunks, CS*2 bytes apart from its neighbour's: stored directly, every
    // instruction writes half sectors 128 B apart and DRAM sees 5.5x the bytes (ncu r1o: 4.5 GB written for 0.82).
    // The warp's 32 pixels are contiguous in memory, so they go through shared memory (chunk index XOR-swizzled
    // against bank conflicts) and leave as KV fully coalesced 512-byte rows.
    constexpr int KV = CS / 8;
    uint4* tw = tile[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < KV; ++k) tw[lane * KV + (k ^ (lane & (KV - 1)))] = pack8(f + k * 8);
    __syncwarp();
    const long long pix0 = pix - lane;  // first pixel of the warp
    uint4* dst = reinterpret_cast<uint4*>(out + (size_t)pix0 * CS);
#pragma unroll
    for (int j = 0; j < KV; ++j) {
      const int sidx = j * 32 + lane;
      const int pl = sidx / KV, kk = sidx % KV;
      if (pix0 + pl < total) dst[pl * KV + (kk ^ (pl & (KV - 1)))] = tw[sidx];
    }
    __syncwarp();
  }
}

// ---- stem weight packing: fp32 (Co, Ci, 1, 7, 7) -> bf16 [rows_pad][16*Cs], k = (dh*4+dw)*Cs + (sh*2+sw)*Ci + c ---
__device__ __forceinline__ void pack_weight_stem_body(const float* __restrict__ w, int Co, int Ci, int Cs,
                                                      __nv_bfloat16* __restrict__ wp, int rows_pad, long long first,
                                                      long long step) {
  const int K = 16 * Cs;
  const long long total = (long long)rows_pad * K;
  for (long long idx = first; idx < total; idx += step) {
    const int k = (int)(idx % K);
    const int row = (int)(idx / K);
    const int tap = k / Cs, cs = k - tap * Cs;
    const int dh = tap >> 2, dw = tap & 3;
    const int quad = cs / Ci, c = cs - quad * Ci;
    const int kh = 2 * dh + (quad >> 1) - 1, kw = 2 * dw + (quad & 1) - 1;
    float v = 0.f;
    if (row < Co && quad < 4 && kh >= 0 && kw >= 0 && kh < 7 && kw < 7) v = w[(((size_t)row * Ci + c) * 7 + kh) * 7 + kw];
    wp[idx] = __float2bfloat16_rn(v);
  }
}
__global__ void pack_weight_stem_kernel(const float* __restrict__ w, int Co, int Ci, int Cs,
                                        __nv_bfloat16* __restrict__ wp, int rows_pad) {
  pack_weight_stem_body(w, Co, Ci, Cs, wp, rows_pad, (long long)blockIdx.x * blockDim.x + threadIdx.x,
                        (long long)gridDim.x * blockDim.x);
}

}  // namespace sfb
