// conv_tma.cuh -- persistent, warp-specialised implicit-GEMM Conv3d for sm_100a (fprop, and dgrad of stride-1 convs).
//
//   A operand : NDHWC bf16 activation, fetched by TMA as shifted 5-D boxes.  An M tile is a rectangular patch of
//               Wt x Ht x Tt output pixels (<= 128) of one clip; filter tap (kt,kh,kw) of that tile is ONE box
//               [box_k channels][Wt*sW][Ht*sH][Tt*sT][1] whose origin is the patch origin shifted by the tap
//               (traversal strides = conv strides), so zero padding and partial edge tiles are TMA out-of-bounds
//               zero fill -- no index arithmetic, no register staging.  box_k = 64 / 32 / 16 channels lands in the
//               SWIZZLE_128B / 64B / 32B K-major layouts tcgen05.mma reads directly.
//   B operand : packed bf16 weights [rows_pad][k_pad] (k = tap*Cg + c), 2-D TMA, SWIZZLE_128B, 64-wide K chunks.
//   D         : fp32 in TMEM, two accumulator buffers so the epilogue of tile i overlaps the MMAs of tile i+1.
//   roles     : warp 0 = TMA producer, warp 1 = MMA issuer, warps 2..5 = epilogue (tcgen05.ld -> affine / residual /
//               ReLU -> bf16 stores; per-channel sum and sum-of-squares for training BatchNorm kept in shared memory
//               across all tiles of the CTA and flushed with one atomic per channel per CTA).
//   schedule  : persistent, grid = #SMs, static round-robin over (pixel patch, N tile) with the N tile fastest.
//
// Replaces cuDNN fprop / bwd-data behind nn.Conv3d in the reference graph (model/my_slowfast.py:94-125).
#pragma once
#include "common.cuh"
#include "../../include/sfb200.h"
#include <cuda.h>

namespace sfb {

__device__ __forceinline__ void tma_load_5d(uint32_t smem_dst, const void* tmap, uint64_t* bar, int32_t c0, int32_t c1,
                                            int32_t c2, int32_t c3, int32_t c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3),
      "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_u32(uint32_t smem_dst, const void* tmap, uint64_t* bar, int32_t x, int32_t y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(x), "r"(y)
      : "memory");
}

// CTA-pair loads: the data lands in THIS CTA's shared memory, the transaction bytes are signalled on a barrier of the
// pair's leader CTA (a shared::cluster address from mapa) -- what .cta_group::2 allows
__device__ __forceinline__ void tma_load_5d_pair(uint32_t smem_dst, const void* tmap, uint32_t bar_cluster, int32_t c0,
                                                 int32_t c1, int32_t c2, int32_t c3, int32_t c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t smem_dst, const void* tmap, uint32_t bar_cluster, int32_t x,
                                                 int32_t y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar_cluster), "r"(x), "r"(y)
      : "memory");
}

// K-major shared-memory operand descriptor for a swizzle span of `span` bytes (128 / 64 / 32):
// rows are `span` bytes apart, 8-row atoms are 8*span bytes apart (SBO); layout type 2 / 4 / 6.
__device__ __forceinline__ uint64_t umma_smem_desc_kmajor(uint32_t smem_addr, uint32_t span) {
  const uint64_t layout = span == 128 ? 2ull : (span == 64 ? 4ull : 6ull);
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;                                   // LBO (unused for swizzled K-major): 16 bytes
  d |= (uint64_t)(((8u * span) >> 4) & 0x3FFF) << 32;       // SBO
  d |= (uint64_t)1 << 46;                                   // descriptor version (sm_100)
  d |= layout << 61;
  return d;
}

struct ConvTmaParams {
  int N, To, Ho, Wo;           // output tensor (rows of the GEMM)
  int Co, o_ld, o_coff;
  int kT, kH, kW, sT, sH, sW, pT, pH, pW;
  int transposed;              // dgrad of a stride-1 conv: tap shifts are mirrored
  int Cg;                      // channels of the gathered view
  int box_k;                   // channels per A box: 64, 32 or 16
  int cchunks;                 // 64-channel chunks per tap (Cg > 64), else 1
  int taps, Ktot, nk;          // Ktot = taps*Cg ; nk = number of 64-wide K chunks of the packed weights
  int wk0;                     // first K column of this launch inside the packed weights (strided-dgrad classes)
  int Wt, Ht, Tt;              // output patch of one M tile; Wt*Ht*Tt <= 128
  int tilesW, tilesH, tilesT;
  int n_tiles;                 // ceil(Co / BN); a CTA works on ONE N tile: blockIdx.x % n_tiles
  int total_patches;           // N*tilesT*tilesH*tilesW
  int ocols;                   // channels per output store box: min(64, Co) rounded to 8/16/32/64
  __nv_bfloat16* out;
  const float* epi_scale;
  const float* epi_shift;
  const __nv_bfloat16* res;
  int res_ld, res_coff;
  int epi_relu;
  float* stats;                // [2*Co] or null
  int reduce_add;              // 1: the TMA store is a reduce-add into the output (in-place gradient accumulation)
  uint32_t stg2_off;           // byte offset of a second staging tile (0: single buffer): the TMA store of chunk k drains
                               // while chunk k+1 is converted and staged
  int nstages;                 // pipeline stages in use (<= ConvTmaCfg::kStages; one fewer when the y tile of the
                               // BatchNorm-backward epilogue takes the last stage's shared memory)
  // ---- fused BatchNorm(+ReLU) backward reduction (sfb_conv_dgrad_bn): this launch produces dz of a conv -> BN -> ReLU
  //      layer; the epilogue fetches the matching tile of the layer's raw output y by TMA (tmap_y, same boxes as the
  //      output store), masks the gradient where the forward ReLU was off, stores the MASKED gradient g and
  //      accumulates dbeta += sum g, dgamma += invstd * sum g*(y - mean) -- the separate bn_bwd_reduce pass (two
  //      tensor reads) disappears.
  int bn_mode;                 // 0: off, 1: mask recomputed from y*scale+shift > 0
  uint32_t ybuf_off;           // byte offset of the y tile (one staging tile)
  uint32_t ybuf2_off;          // second side tile (0: none): side tiles are then requested TWO chunks ahead.  With one
                               // tile in flight a chunk could not take less than the HBM latency of its side tile
                               // (request -> arrival -> math -> store -> next request): the residual convs of the eval
                               // forward ran at 3 us per 128x128 chunk, 800 issue cycles of work in 5800
  int res_tile;                // kEpiAffine: the residual tile arrives by TMA through tmap_y into the same buffer, one
                               // chunk ahead (per-thread global loads left the epilogue waiting on HBM once per chunk:
                               // the block-tail convs of the eval forward ran at half their roofline)
  // ---- halo mode ((1,3,3) stride-1 convs on 64-channel pixels, weights resident): ONE box per tile brings the patch
  //      with its one-pixel border, [Ht+2][Wt+2] pixels of 128 bytes; tap (kh, kw) is the same shared memory read
  //      through a descriptor whose start is kh*(Wt+2) + kw pixel rows further on.  MMA row r stands for pixel
  //      (r / pitch, r % pitch): the two rows per line that wrap into the next line are computed and thrown away.
  //      Nine 16 KB boxes per tile became one of <= 21 KB: these layers were bound by the shared-memory bandwidth that
  //      the TMA fills and the MMA operand reads share (ncu r3x: 36 % tensor, 36 % L2, 64 % SM throughput).
  int pair;                    // 1: launch the CTA-pair instantiation (cta_group::2) when the weights are streamed
  int halo;                    // 0 / 1
  int pitch;                   // Wt + 2
  uint32_t halo_stride;        // bytes between pipeline stages in halo mode (one stage = one tile's box)
  uint32_t halo_bytes;         // bytes of one box: pitch * (Ht + 2) * 128
  const float* bn_scale;
  const float* bn_shift;
  const float* bn_mean;
  const float* bn_invstd;
  float* bn_dgamma;
  float* bn_dbeta;
  int bn_C, bn_ld;             // real channels; channel of output column j = j % bn_ld (pixel-folded launches)
};

__device__ __forceinline__ void tma_store_5d(const void* tmap, uint32_t smem_src, int32_t c0, int32_t c1, int32_t c2,
                                             int32_t c3, int32_t c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];"
      ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_src), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tma_reduce_add_5d(const void* tmap, uint32_t smem_src, int32_t c0, int32_t c1, int32_t c2,
                                                  int32_t c3, int32_t c4) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.5d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];"
      ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_src), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ uint4 lds_v4(uint32_t saddr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
  return v;
}

// Walk of a persistent CTA over its pixel patches (patch0, patch0 + step, ...) as (tw, th, tt, n) counters: the
// mixed-radix decomposition is done once, every further tile is a few adds and compares instead of four divisions.
struct PatchWalk {
  int tw, th, tt, n, dw, dh, dt, dn;
  __device__ __forceinline__ void init(int patch0, int step, const ConvTmaParams& p) {
    int q = patch0;
    tw = q % p.tilesW; q /= p.tilesW;
    th = q % p.tilesH; q /= p.tilesH;
    tt = q % p.tilesT; n = q / p.tilesT;
    q = step;
    dw = q % p.tilesW; q /= p.tilesW;
    dh = q % p.tilesH; q /= p.tilesH;
    dt = q % p.tilesT; dn = q / p.tilesT;
  }
  __device__ __forceinline__ void next(const ConvTmaParams& p) {
    tw += dw;
    if (tw >= p.tilesW) { tw -= p.tilesW; ++th; }
    th += dh;
    if (th >= p.tilesH) { th -= p.tilesH; ++tt; }
    tt += dt;
    if (tt >= p.tilesT) { tt -= p.tilesT; ++n; }
    n += dn;
  }
};

// BRES: the CTA's whole weight slice [BN][Ktot] stays resident in shared memory (loaded once; a persistent CTA works
// on one N tile for its whole life), so a pipeline stage carries the activation boxes only.  Small-K layers were bound
// by L2->SM traffic, of which the per-tile weight re-fetch was 1/3 (64->64 (1,3,3)) to 2/3 (64->256 pointwise).
// PAIR: two CTAs of a cluster run one 256-row MMA (cta_group::2) on a 256-wide N tile; each CTA stages its own 128 pixel
// rows and HALF of the weight tile (128 of the 256 rows): a stage is 32 KB instead of 48 and an MMA reads 8 KB of
// operands per SM instead of 12 -- the streamed-weight layers were bound by that shared-memory traffic at ~68 % of the
// MMA floor (fill 48 KB + operand reads 48 KB per 512 MMA cycles against a 128 B/clk port).
template <int BN, bool BRES = false, bool PAIR = false>
struct ConvTmaCfg {
  static_assert(!PAIR || (BN == 256 && !BRES), "the CTA-pair kernel is the 256-wide streamed-weight one");
  static constexpr int kStages = PAIR ? 6 : (BRES ? (BN >= 128 ? 6 : 8) : (BN >= 256 ? 4 : (BN >= 128 ? 6 : 8)));
  static constexpr int kAStage = 128 * 128;  // 16 KB: 128 rows x (64 | 2x32 | 4x16) channels
  static constexpr int kBStage = BRES ? 0 : (PAIR ? 128 * 128 : BN * 128);
  static constexpr int kStage = kAStage + kBStage;
  static constexpr int kChunk = BN < 128 ? BN : 128;                  // output columns staged per pass
  static constexpr int kChunks = BN / kChunk;
  static constexpr int kStagingOff = kStages * kStage;
  static constexpr int kStagingBytes = 128 * kChunk * 2;
  static constexpr int kBarOff = kStagingOff + kStagingBytes;          // full[S], empty[S], tfull[2], tempty[2]
  static constexpr int kHolderOff = kBarOff + (2 * kStages + 7) * 8;   // + bres, ybar[2]
  static constexpr int kEnd = kHolderOff + 16;
  static constexpr int kBResOff = (kEnd + 1023) / 1024 * 1024;        // resident weights: nk blocks of [BN][128 B]
  static constexpr uint32_t kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;
  // epilogue warps: one per (TMEM lane quadrant, 32-column slice of a staged chunk).  With a single warp per SM
  // sub-partition the epilogue was dependency-stall bound (ncu r1b: 38 % "wait", 21 % issuing, 1 warp / scheduler).
  static constexpr int kEpiWarps = kChunk / 8;        // 16 (BN >= 128), 8 (BN = 64), 4 (BN = 32)
  static constexpr int kEpiThreads = 32 * kEpiWarps;
  static constexpr int kThreads = 64 + kEpiThreads;
  static size_t smem_bytes(int nk = 0) { return 1024 + (BRES ? (size_t)kBResOff + (size_t)nk * BN * 128 : (size_t)kEnd); }
};

// MODE selects the epilogue, as a template parameter so that no variant pays registers for another one's state (the
// kernel sits at the 96-register cap of its 576-thread block):
//   kEpiPlain  training fprop / dgrad: convert + stage (+ BatchNorm statistics of the forward, + reduce-add stores)
//   kEpiAffine inference: folded-BatchNorm affine, residual, ReLU
//   kEpiBnBwd  dgrad with the fused BatchNorm-backward reduction (ConvTmaParams::bn_*)
enum { kEpiPlain = 0, kEpiAffine = 1, kEpiBnBwd = 2 };
template <int BN, bool BRES = false, int MODE = kEpiPlain, bool PAIR = false>
__global__ void __launch_bounds__(ConvTmaCfg<BN, BRES, PAIR>::kThreads, 1)
conv_tma_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                const __grid_constant__ CUtensorMap tmap_o, const __grid_constant__ CUtensorMap tmap_y,
                const ConvTmaParams p) {
  using L = ConvTmaCfg<BN, BRES, PAIR>;
  // CTA pair: launched as clusters of two; rank 0 (the leader) issues the MMAs for both
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  constexpr int SMAX = L::kStages;
  const int S = p.nstages;
  const bool halo = BRES && BN == 64 && p.halo != 0;  // only the 64-wide resident-weight instantiations carry the code
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* empty_bar = full_bar + SMAX;
  uint64_t* tfull_bar = empty_bar + SMAX;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* bres_bar = tempty_bar + 2;
  uint64_t* y_bar = bres_bar + 1;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(smem + L::kHolderOff);

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx();
  const int lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < SMAX; ++s) {
      mbar_init(&full_bar[s], 1);  // pair: the leader arrives and expects the bytes of BOTH CTAs' loads
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull_bar[b], 1);
      mbar_init(&tempty_bar[b], PAIR ? 2 * L::kEpiWarps : L::kEpiWarps);  // pair: the epilogues of both CTAs
    }
    mbar_init(bres_bar, 1);
    mbar_init(&y_bar[0], 1);
    mbar_init(&y_bar[1], 1);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_o);
    if (p.bn_mode) tma_prefetch_desc(&tmap_y);
  }
  if (halo) {
    // staging rows past the patch's Wt*Ht pixels are never written in halo mode and must read as zeros (statistics)
    for (int i = tid; i < L::kStagingBytes / 16; i += L::kThreads) {
      reinterpret_cast<uint4*>(smem + L::kStagingOff)[i] = make_uint4(0, 0, 0, 0);
      if (p.stg2_off) reinterpret_cast<uint4*>(smem + p.stg2_off)[i] = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async_smem();
  }
  if (p.bn_mode) {
    // rows of the y tile a box never writes (patches of fewer than 128 pixels) must read as finite numbers
    for (int i = tid; i < L::kStagingBytes / 16; i += L::kThreads) {
      reinterpret_cast<uint4*>(smem + p.ybuf_off)[i] = make_uint4(0, 0, 0, 0);
      if (p.ybuf2_off) reinterpret_cast<uint4*>(smem + p.ybuf2_off)[i] = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async_smem();
  }
  if (warp == 2) {
    tmem_alloc(tmem_holder, L::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anything is signalled on them
  tc_fence_after();
  const uint32_t tmem_base = uniform_u32(*tmem_holder);
  const int rows = p.Wt * p.Ht * p.Tt;
  const uint32_t box_bytes = (uint32_t)(rows * p.box_k * 2);
  const int bpc = p.Cg >= 64 ? 1 : 64 / p.box_k;  // A boxes per 64-wide K chunk
  const uint32_t sub_stride = (uint32_t)(128 * p.box_k * 2);
  const int khw = p.kH * p.kW;
  // this CTA's N tile and its walk over the pixel patches
  // (pair: the two CTAs take neighbouring patches 2q + rank of the same N tile; the last pair of an odd patch count
  //  gives rank 1 a patch past the end -- its boxes are out of range: zero operands, clipped stores -- so that both
  //  CTAs run the same number of tiles)
  const int unit = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int n_units = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int nt = unit % p.n_tiles;
  const int patch0 = PAIR ? 2 * (unit / p.n_tiles) + (int)rank : unit / p.n_tiles;
  const int patch_step = PAIR ? 2 * (n_units / p.n_tiles) : n_units / p.n_tiles;
  const int patch_lim = PAIR ? p.total_patches + (int)rank : p.total_patches;
  const int n0 = nt * BN;

  if (warp == 0) {
    // ======================= TMA producer =======================
    // One instruction stream feeds the whole pipeline, so its per-chunk instruction count is on the critical path of
    // every small-K / small-N layer: the loop is warp-uniform (operands in uniform registers, one elected lane
    // issues), taps and stages advance with counters, no division inside the chunk loop.
    const bool leader = elect_one();
    if (BRES && patch0 < p.total_patches && leader) {
      mbar_arrive_expect_tx(bres_bar, (uint32_t)(p.nk * BN * 128));
      for (int kc = 0; kc < p.nk; ++kc)
        tma_load_2d_u32(smem_u32(smem + L::kBResOff) + kc * (BN * 128), &tmap_w, bres_bar, p.wk0 + kc * 64, n0);
    }
    const int sgn = p.transposed ? -1 : 1;
    const uint32_t smem0 = smem_u32(smem);
    int s = 0;
    uint32_t ph = 1;  // parity to wait on for "stage s is free"
    PatchWalk pw;
    pw.init(patch0, patch_step, p);
    for (int patch = patch0; patch < patch_lim; patch += patch_step, pw.next(p)) {
      const int tw = pw.tw, th = pw.th, tt = pw.tt, n = pw.n;
      if (halo) {
        mbar_wait(&empty_bar[s], ph, 11);
        if (leader) {
          mbar_arrive_expect_tx(&full_bar[s], p.halo_bytes);
          tma_load_5d(smem0 + (uint32_t)s * p.halo_stride, &tmap_a, &full_bar[s], 0, tw * p.Wt - 1, th * p.Ht - 1, tt, n);
        }
        __syncwarp();
        if (++s == S) { s = 0; ph ^= 1; }
        continue;
      }
      const int bw = p.transposed ? tw * p.Wt + p.pW : tw * p.Wt * p.sW - p.pW;
      const int bh = p.transposed ? th * p.Ht + p.pH : th * p.Ht * p.sH - p.pH;
      const int bt = p.transposed ? tt * p.Tt + p.pT : tt * p.Tt * p.sT - p.pT;
      int kt = 0, kh = 0, kw = 0, cc = 0, tap = 0;
      for (int kc = 0; kc < p.nk; ++kc) {
        mbar_wait(&empty_bar[s], ph, 11);
        const uint32_t sa = smem0 + (uint32_t)s * L::kStage;
        if (PAIR) {
          // both CTAs' loads count their bytes on the leader's barrier (its own 16 KB of pixels and its half of the
          // weight tile each); only the leader arrives.  The peer cannot get a phase ahead: it refills a stage only after
          // the leader's MMAs on it were committed, i.e. after the leader saw this phase complete.
          if (leader) {
            const uint32_t fb = mapa_u32(smem_u32(&full_bar[s]), 0);
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[s], 2u * (box_bytes + (uint32_t)L::kBStage));
            tma_load_5d_pair(sa, &tmap_a, fb, cc * 64, bw + sgn * kw, bh + sgn * kh, bt + sgn * kt, n);
            tma_load_2d_pair(sa + L::kAStage, &tmap_w, fb, p.wk0 + kc * 64, n0 + (int)rank * 128);
          }
          if (++cc == p.cchunks) {
            cc = 0;
            if (++kw == p.kW) { kw = 0; if (++kh == p.kH) { kh = 0; ++kt; } }
          }
        } else if (p.Cg >= 64) {
          if (leader) {
            mbar_arrive_expect_tx(&full_bar[s], box_bytes + (uint32_t)L::kBStage);
            tma_load_5d(sa, &tmap_a, &full_bar[s], cc * 64, bw + sgn * kw, bh + sgn * kh, bt + sgn * kt, n);
          }
          if (++cc == p.cchunks) {
            cc = 0;
            if (++kw == p.kW) { kw = 0; if (++kh == p.kH) { kh = 0; ++kt; } }
          }
        } else {
          const int nbox = min(bpc, p.taps - tap);
          if (leader) mbar_arrive_expect_tx(&full_bar[s], (uint32_t)nbox * box_bytes + (uint32_t)L::kBStage);
          for (int b = 0; b < nbox; ++b) {
            if (leader)
              tma_load_5d(sa + b * sub_stride, &tmap_a, &full_bar[s], 0, bw + sgn * kw, bh + sgn * kh, bt + sgn * kt, n);
            if (++kw == p.kW) { kw = 0; if (++kh == p.kH) { kh = 0; ++kt; } }
          }
          tap += nbox;
        }
        if (!BRES && !PAIR && leader) tma_load_2d_u32(sa + L::kAStage, &tmap_w, &full_bar[s], p.wk0 + kc * 64, n0);
        __syncwarp();
        if (++s == S) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ======================= MMA issuer (warp-uniform loop, one elected lane issues; pair: the leader CTA only) =======
    const bool leader = elect_one();
    const uint32_t idesc = umma_idesc_bf16(PAIR ? 256 : 128, BN, 0, 0);
    const uint32_t span = (uint32_t)(p.box_k * 2);
    // descriptors = constant + offset in the 14-bit start-address field: one add per operand per MMA
    const uint64_t adesc0 = umma_smem_desc_kmajor(smem_u32(smem), span);
    const uint64_t bdesc0 = umma_smem_desc_kmajor(BRES ? smem_u32(smem + L::kBResOff) : smem_u32(smem) + L::kAStage, 128);
    uint32_t aoff[4];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const int koff = kk * 16;
      const int sub = koff / p.box_k;
      aoff[kk] = (uint32_t)((sub * sub_stride + (koff - sub * p.box_k) * 2) >> 4);
    }
    const int n_last = (p.Ktot - (p.nk - 1) * 64 + 15) >> 4;  // 16-wide K steps of the last chunk
    int s = 0;
    uint32_t ph = 0;
    int it = 0;
    if (BRES && patch0 < p.total_patches) mbar_wait(bres_bar, 0, 15);
    for (int patch = patch0; patch < patch_lim; patch += patch_step, ++it) {
      const int buf = it & 1;
      mbar_wait(&tempty_bar[buf], (uint32_t)(((it >> 1) & 1) ^ 1), 12);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BN);
      if (halo) {
        // one stage = the tile's box; tap (kh, kw) starts kh*pitch + kw pixel rows into it (mirrored for dgrad)
        mbar_wait(&full_bar[s], ph, 13);
        tc_fence_after();
        const uint64_t ad_s = adesc0 + (uint64_t)(((uint32_t)s * p.halo_stride) >> 4);
        if (leader) {
          int kh = 0, kw = 0;
          for (int tap = 0; tap < 9; ++tap) {
            const int roff = p.transposed ? (2 - kh) * p.pitch + (2 - kw) : kh * p.pitch + kw;
            const uint64_t ad0 = ad_s + (uint64_t)((uint32_t)(roff * 128) >> 4);
            const uint64_t bd0 = bdesc0 + (uint64_t)(((uint32_t)tap * (BN * 128)) >> 4);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16(d_tmem, ad0 + (uint64_t)(kk * 2), bd0 + (uint64_t)(kk * 2), idesc, (tap | kk) != 0 ? 1u : 0u);
            if (++kw == 3) { kw = 0; ++kh; }
          }
          umma_commit(&empty_bar[s]);
          umma_commit(&tfull_bar[buf]);
        }
        __syncwarp();
        if (++s == S) { s = 0; ph ^= 1; }
        continue;
      }
      for (int kc = 0; kc < p.nk; ++kc) {
        mbar_wait(&full_bar[s], ph, 13);
        tc_fence_after();
        const uint64_t ad0 = adesc0 + (uint64_t)(((uint32_t)s * L::kStage) >> 4);
        const uint64_t bd0 = bdesc0 + (uint64_t)((BRES ? (uint32_t)kc * (BN * 128) : (uint32_t)s * L::kStage) >> 4);
        const int nkk = kc == p.nk - 1 ? n_last : 4;
        if (leader) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            if (kk < nkk) {
              if (PAIR) umma_bf16_pair(d_tmem, ad0 + aoff[kk], bd0 + (uint64_t)(kk * 2), idesc, (kc | kk) != 0 ? 1u : 0u);
              else umma_bf16(d_tmem, ad0 + aoff[kk], bd0 + (uint64_t)(kk * 2), idesc, (kc | kk) != 0 ? 1u : 0u);
            }
          }
          if (PAIR) umma_commit_pair(&empty_bar[s]); else umma_commit(&empty_bar[s]);
        }
        __syncwarp();
        if (++s == S) { s = 0; ph ^= 1; }
      }
      if (leader) {
        if (PAIR) umma_commit_pair(&tfull_bar[buf]); else umma_commit(&tfull_bar[buf]);
      }
      __syncwarp();
    }
  } else if (warp >= 2) {
    // ======================= epilogue (warps 2..5; TMEM lanes 32*(warp&3)..) =======================
    // TMEM -> registers -> (affine, residual, ReLU) -> bf16 -> swizzled staging tile in shared memory -> TMA store.
    // BatchNorm statistics are column sums of the staged (bf16-rounded) tile, kept in registers across all tiles.
    constexpr int CH = L::kChunk;
    constexpr int NCH = L::kChunks;
    constexpr int ET = L::kEpiThreads;
    const int q4 = warp & 3;             // TMEM lane quadrant this warp may read
    const int te = tid - 64;
    const int cslice = (warp - 2) >> 2;  // 32-column slice of the chunk
    const int r_mma = q4 * 32 + lane;    // accumulator row (TMEM lane) of this thread
    // halo mode: row r_mma is pixel (r_mma / pitch, r_mma % pitch) of the patch; its staging row is the compact index
    const bool row_live = !halo || (r_mma % p.pitch < p.Wt && r_mma / p.pitch < p.Ht);
    const int r = halo ? (row_live ? (r_mma / p.pitch) * p.Wt + r_mma % p.pitch : 0) : r_mma;
    const int ww = r % p.Wt;
    const int hh = (r / p.Wt) % p.Ht;
    const int tq = r / (p.Wt * p.Ht);
    const int ocols = p.ocols;                       // 8, 16, 32 or 64
    const int olog = 31 - __clz(ocols);
    const uint32_t rb = (uint32_t)(ocols * 2);
    const uint32_t swz = rb == 128 ? 7u : (rb == 64 ? 3u : (rb == 32 ? 1u : 0u));
    const uint32_t smem_e = keep_u32(smem_u32(smem));
    const uint32_t stg_a = keep_u32(smem_e + L::kStagingOff);
    const uint32_t stg_b = keep_u32(p.stg2_off ? smem_e + p.stg2_off : stg_a);
    uint32_t nstaged = 0;
    const uint32_t sub_bytes = 128u * rb;
    const int last_ch = min(NCH, (p.Co - n0 + CH - 1) / CH) - 1;
    float ssum[NCH][8], ssq[NCH][8];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { ssum[c][i] = 0.f; ssq[c][i] = 0.f; }
    }
    // The epilogue is instruction-bound on the memory-bound layers (ncu r1o, 64->256 pointwise: 15 k warp
    // instructions per 128x256 tile, 58 % issue utilisation, of which ~6 k were per-group branches, parameter
    // reloads and address arithmetic): everything that does not change from tile to tile is computed once here.
    constexpr bool plain = MODE != kEpiAffine;
    uint32_t soff[4], toff[4];  // swizzled staging offsets: this thread's four 16-byte stores / four stats loads
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int cl = cslice * 32 + j * 8;
      uint32_t o = (uint32_t)(cl >> olog) * sub_bytes + (uint32_t)r * rb + (uint32_t)((cl & (ocols - 1)) * 2);
      soff[j] = keep_u32(o ^ (((o >> 7) & swz) << 4));
      const int cs = (warp - 2) * 8;
      o = (uint32_t)(cs >> olog) * sub_bytes + (uint32_t)(lane + 32 * j) * rb + (uint32_t)((cs & (ocols - 1)) * 2);
      toff[j] = o ^ (((o >> 7) & swz) << 4);
      if (MODE != kEpiAffine) toff[j] = keep_u32(toff[j]);
    }
    const bool want_stats = MODE == kEpiPlain && p.stats != nullptr;
    constexpr bool bnred = MODE == kEpiBnBwd;
    // a side tile (y of the fused BatchNorm backward / the residual of the inference epilogue) travels by TMA
    const bool tile = bnred || (MODE == kEpiAffine && p.res_tile);
    // TMA stores alternate between two staging tiles when the second one exists and is not lent to the side tile
    const bool dbl = p.stg2_off != 0 && (!tile || p.ybuf_off != p.stg2_off);
    const uint32_t ybuf_a = keep_u32(smem_e + p.ybuf_off);
    const uint32_t ybuf_b = keep_u32(p.ybuf2_off ? smem_e + p.ybuf2_off : ybuf_a);
    // this thread's TMEM address (lane quadrant, 32-column slice) and its slice of the per-channel coefficients
    const uint32_t tmem_t = keep_u32(tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(cslice * 32));
    const float* const es_t = keep_ptr(p.epi_scale ? p.epi_scale + n0 + cslice * 32 : nullptr);
    const float* const eb_t = keep_ptr(p.epi_shift ? p.epi_shift + n0 + cslice * 32 : nullptr);
    // fused BatchNorm backward: channel of this warp's 8-column group in each chunk (pixel-folded launches: j % bn_ld)
    uint32_t cbv[NCH];
#pragma unroll
    for (int c = 0; c < NCH; ++c) cbv[c] = bnred ? keep_u32((uint32_t)((n0 + c * CH + (warp - 2) * 8) % p.bn_ld)) : 0u;
    const bool ydeep = p.ybuf2_off != 0;
    uint32_t ycons = 0;  // side tiles consumed so far: tile i sits in buffer i & 1 (one buffer: always the first)
    bool store_pending = false;
    int it = 0;
    PatchWalk pw;
    pw.init(patch0, patch_step, p);
    // side-tile requests (thread te == 0): a cursor over this CTA's chunks in the order the epilogue visits them; one or
    // two requests are in flight ahead of the chunk being staged
    PatchWalk pf = pw;
    int pf_patch = patch0, pf_ch = 0;
    uint32_t yreq = 0;
    auto request_side_tile = [&]() {
      if (pf_patch >= patch_lim) return;
      const int col = n0 + pf_ch * CH;
      const int nsub = (min(CH, p.Co - col) + ocols - 1) >> olog;
      uint64_t* bar = &y_bar[ydeep ? (yreq & 1) : 0];
      const uint32_t dst = (yreq & 1) ? ybuf_b : ybuf_a;
      mbar_arrive_expect_tx(bar, (uint32_t)nsub * (uint32_t)rows * rb);
      for (int gsub = 0; gsub < nsub; ++gsub)
        tma_load_5d(dst + gsub * sub_bytes, &tmap_y, bar, col + gsub * ocols, pf.tw * p.Wt, pf.th * p.Ht, pf.tt * p.Tt, pf.n);
      ++yreq;
      if (++pf_ch > last_ch) {
        pf_ch = 0;
        pf_patch += patch_step;
        pf.next(p);
      }
    };
    if (tile && te == 0) {
      request_side_tile();
      if (ydeep) request_side_tile();
    }
    for (int patch = patch0; patch < patch_lim; patch += patch_step, ++it, pw.next(p)) {
      const int tw = pw.tw, th = pw.th, tt = pw.tt, n = pw.n;
      const int w0 = tw * p.Wt, h0 = th * p.Ht, t0 = tt * p.Tt;
      const int wo = w0 + ww, ho = h0 + hh, to = t0 + tq;
      const bool mv = row_live && r < rows && wo < p.Wo && ho < p.Ho && to < p.To && n < p.N;
      const long long m = (p.res && mv) ? (((long long)n * p.To + to) * p.Ho + ho) * p.Wo + wo : 0;
      const __nv_bfloat16* rrow = p.res ? p.res + (size_t)m * p.res_ld + p.res_coff : nullptr;
      const int buf = it & 1;
      mbar_wait(&tfull_bar[buf], (uint32_t)((it >> 1) & 1), 14);
      tc_fence_after();
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        if (ch <= last_ch) {
          const int col0 = n0 + ch * CH;
          // (A) the TMA store that last used this staging tile must have finished reading it
          const uint32_t stg = (dbl && (nstaged & 1)) ? stg_b : stg_a;
          ++nstaged;
          //     -- one staging tile: wait and tell everybody here.  Two tiles: the wait sits in front of the barrier of
          //     (D) of the PREVIOUS chunk (the store that used this tile was issued a whole chunk before that), so this
          //     barrier and its skew disappear
          if (!dbl) {
            if (te == 0 && store_pending) tma_store_wait_read();
            named_bar_sync(1, ET);
          }
          // this chunk's side tile: buffer, barrier and the parity of that barrier's current phase
          const uint32_t ybuf = (ycons & 1) ? ybuf_b : ybuf_a;
          uint64_t* const ybar = &y_bar[ydeep ? (ycons & 1) : 0];
          const uint32_t yphase = ydeep ? ((ycons >> 1) & 1) : (ycons & 1);
          // (B) accumulator -> staging: this warp's 32-column slice
          {
            const int nb = col0 + cslice * 32;
            if constexpr (plain) {
              // training fprop / dgrad: convert and stage, nothing else (ragged last slice: only the live 8-column groups)
              if (nb < p.Co) {
                float v[32];
                tmem_ld32(tmem_t + (uint32_t)(buf * BN + ch * CH), v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  if (nb + j * 8 < p.Co) {
                    uint4 pk = pack8(v + j * 8);
                    if (!mv) pk = make_uint4(0, 0, 0, 0);
                    if (row_live) sts_v4(stg + soff[j], pk);
                  }
                }
              }
            } else if (nb < p.Co) {
              // inference epilogue (folded BatchNorm, residual, ReLU).  The eval forward is epilogue-bound on the
              // memory-bound layers, so: the residual's 64 bytes are requested before the accumulator is read, the affine
              // coefficients come as two 16-byte loads per 8 columns (they were 16 scalar loads), the staging offsets are
              // the precomputed ones.
              uint4 rq[4];
              const bool has_res = rrow != nullptr && mv;
              if (has_res && !tile) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  rq[j] = nb + j * 8 < p.Co ? *reinterpret_cast<const uint4*>(rrow + nb + j * 8) : make_uint4(0, 0, 0, 0);
              }
              float v[32];
              tmem_ld32(tmem_t + (uint32_t)(buf * BN + ch * CH), v);
              tmem_ld_wait();
              if (tile) {
                // residual tile (same swizzled layout as the staging tile: this thread's four 16-byte groups sit at soff[])
                mbar_wait(ybar, yphase, 17);
#pragma unroll
                for (int j = 0; j < 4; ++j) rq[j] = lds_v4(ybuf + soff[j]);
              }
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int nn = nb + j * 8;
                if (nn < p.Co) {
                  float* f = v + j * 8;
                  if (es_t) {
                    float es[8], eb[8];
                    load8f(es_t + ch * CH + j * 8, es);
                    load8f(eb_t + ch * CH + j * 8, eb);
#pragma unroll
                    for (int i = 0; i < 8; ++i) f[i] = fmaf(f[i], es[i], eb[i]);
                  }
                  if (has_res) {
                    float rf[8];
                    unpack8(rq[j], rf);
#pragma unroll
                    for (int i = 0; i < 8; ++i) f[i] += rf[i];
                  }
                  if (p.epi_relu) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) f[i] = fmaxf(f[i], 0.f);
                  }
                  uint4 pk = pack8(f);
                  if (!mv) pk = make_uint4(0, 0, 0, 0);
                  if (row_live) sts_v4(stg + soff[j], pk);
                }
              }
            }
          }
          // (C) last chunk read: hand the accumulator buffer back to the MMA warp
          if (ch == last_ch) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (PAIR && rank != 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[buf]), 0));
              else mbar_arrive(&tempty_bar[buf]);
            }
          }
          // (C') fused BatchNorm(+ReLU) backward reduction: epilogue warp e owns the 8-column group e of the chunk, its
          //      lanes split the 128 rows (as in (E)); the staged gradient is masked IN PLACE before it is stored
          if (bnred) {
            named_bar_sync(1, ET);                       // the whole gradient tile is staged
            mbar_wait(ybar, yphase, 16);
            const int cg = col0 + (warp - 2) * 8;
            if (cg < p.Co) {
              const int cb = (int)cbv[ch];               // channel of this thread's first column
              float sc[8], sh[8], mu[8];
              if (cb < p.bn_C) {                         // C is a multiple of 8: the group is real or pad as a whole
                load8f(p.bn_scale + cb, sc);
                load8f(p.bn_shift + cb, sh);
                load8f(p.bn_mean + cb, mu);
              } else {
#pragma unroll
                for (int c = 0; c < 8; ++c) { sc[c] = 0.f; sh[c] = 0.f; mu[c] = 0.f; }
              }
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                float g[8], yv[8];
                unpack8(lds_v4(stg + toff[i]), g);
                unpack8(lds_v4(ybuf + toff[i]), yv);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                  g[c] = fmaf(yv[c], sc[c], sh[c]) > 0.f ? g[c] : 0.f;
                  ssum[ch][c] += g[c];
                  ssq[ch][c] = fmaf(g[c], yv[c] - mu[c], ssq[ch][c]);
                }
                sts_v4(stg + toff[i], pack8(g));
              }
            }
          }
          // (D) staged tile -> global memory (TMA clips rows/columns outside the tensor)
          fence_proxy_async_smem();
          if (dbl && te == 0 && store_pending) tma_store_wait_read();  // the other tile is free for the next chunk
          named_bar_sync(1, ET);
          if (te == 0) {
            const int nsub = (min(CH, p.Co - col0) + ocols - 1) >> olog;
            for (int gsub = 0; gsub < nsub; ++gsub) {
              if (p.reduce_add) tma_reduce_add_5d(&tmap_o, stg + gsub * sub_bytes, col0 + gsub * ocols, w0, h0, t0, n);
              else tma_store_5d(&tmap_o, stg + gsub * sub_bytes, col0 + gsub * ocols, w0, h0, t0, n);
            }
            tma_store_commit();
            store_pending = true;
            // every epilogue thread is past its reads of this chunk's side tile (barrier above): its buffer takes the
            // next request, which travels during the following chunk(s)' store wait, barrier and conversion
            if (tile) request_side_tile();
          }
          if (tile) ++ycons;
          // (E) BatchNorm statistics from the staged tile (rows outside the tensor were staged as zeros): epilogue
          //     warp e owns the 8-column group e of the chunk, its lanes split the 128 rows (16-byte reads)
          if (want_stats) {
            if (col0 + (warp - 2) * 8 < p.Co) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                float x[8];
                unpack8(lds_v4(stg + toff[i]), x);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                  ssum[ch][c] += x[c];
                  ssq[ch][c] = fmaf(x[c], x[c], ssq[ch][c]);
                }
              }
            }
          }
        }
      }
    }
    if (te == 0 && store_pending) tma_store_wait_all();
    if (bnred) {
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        const int col = n0 + ch * CH + (warp - 2) * 8;
        const int cb = (int)cbv[ch];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          float a = ssum[ch][c], b = ssq[ch][c];
#pragma unroll
          for (int off = 16; off >= 1; off >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, off);
            b += __shfl_xor_sync(0xffffffffu, b, off);
          }
          if (lane == c && ch <= last_ch && col < p.Co && cb < p.bn_C) {
            atomicAdd(p.bn_dbeta + cb + c, a);
            atomicAdd(p.bn_dgamma + cb + c, b * __ldg(p.bn_invstd + cb + c));
          }
        }
      }
    }
    if (p.stats) {
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        const int col = n0 + ch * CH + (warp - 2) * 8;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          float a = ssum[ch][c], b = ssq[ch][c];
#pragma unroll
          for (int off = 16; off >= 1; off >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, off);
            b += __shfl_xor_sync(0xffffffffu, b, off);
          }
          if (lane == c && ch <= last_ch && col < p.Co) {
            atomicAdd(p.stats + col + c, a);
            atomicAdd(p.stats + p.Co + col + c, b);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the leader's MMAs and commits touch the peer's shared memory and TMEM until here
  if (warp == 2) tmem_dealloc(tmem_base, L::kTmemCols);
}

}  // namespace sfb
