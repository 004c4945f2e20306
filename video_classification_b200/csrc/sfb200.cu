// sfb200.cu -- C ABI of libsfb200.so (see include/sfb200.h): argument validation, TMA tensor-map
// cache, kernel dispatch.  Single translation unit; links only cudart (static).  The driver entry point
// cuTensorMapEncodeTiled is resolved at run time so the library also loads on a box without a GPU.
#include "conv_simt.cuh"
#include "conv_tma.cuh"
#include "wgrad_tma.cuh"
#include "stem_win.cuh"
#include "elementwise.cuh"
#include "pack_batch.cuh"
#include "stage_u8.cuh"
#include "scores.cuh"
#include "adam.cuh"

#include <algorithm>
#include <cmath>
#include <array>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

using namespace sfb;

namespace {

thread_local char g_cuda_err[256] = "";
std::atomic<long long> g_launches{0};

int cuda_fail(cudaError_t e, const char* where) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s", where, cudaGetErrorString(e));
  (void)cudaGetLastError();  // non-sticky errors must not leak into the next call's launch check
  return SFB_ERR_CUDA;
}
#define SFB_CHECK_LAUNCH(where)                         \
  do {                                                  \
    g_launches.fetch_add(1, std::memory_order_relaxed); \
    cudaError_t e__ = cudaGetLastError();               \
    if (e__ != cudaSuccess) return cuda_fail(e__, where); \
  } while (0)

inline int grid_for(long long total, int threads, int cap = 148 * 16) {
  long long b = (total + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > cap) b = cap;
  return (int)b;
}
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---- N tile / padding policy (shared with sfb_packed_weight_dims) ----
int pick_bn(int Co) {
  if (Co <= 32) return 32;
  if (Co <= 64) return 64;
  if (Co <= 128) return 128;
  return 256;
}

// ---- TMA tensor maps for packed weights ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
std::mutex g_mu;
EncodeTiledFn g_encode = nullptr;
std::map<std::tuple<const void*, int, int, int>, CUtensorMap> g_tmaps;

int get_weight_tmap(const void* wp, int rows_pad, int k_pad, int bn, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto key = std::make_tuple(wp, rows_pad, k_pad, bn);
  auto it = g_tmaps.find(key);
  if (it != g_tmaps.end()) {
    *out = it->second;
    return SFB_OK;
  }
  if (!g_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
      snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaGetDriverEntryPoint(cuTensorMapEncodeTiled) failed");
      return SFB_ERR_DRIVER;
    }
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  }
  CUtensorMap tm;
  cuuint64_t gdim[2] = {(cuuint64_t)k_pad, (cuuint64_t)rows_pad};
  cuuint64_t gstr[1] = {(cuuint64_t)k_pad * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)bn};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(wp), gdim, gstr, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "cuTensorMapEncodeTiled failed: %d", (int)r);
    return SFB_ERR_DRIVER;
  }
  if (g_tmaps.size() > 4096) g_tmaps.clear();
  g_tmaps[key] = tm;
  *out = tm;
  return SFB_OK;
}

int check_geom(const SfbConvGeom* g) {
  if (!g) return SFB_ERR_BAD_ARG;
  if (g->N <= 0 || g->Ti <= 0 || g->Hi <= 0 || g->Wi <= 0 || g->To <= 0 || g->Ho <= 0 || g->Wo <= 0) return SFB_ERR_BAD_ARG;
  if (g->Cg <= 0 || (g->Cg & 7) || g->Co <= 0 || (g->Co & 7)) return SFB_ERR_BAD_ARG;
  if (g->kT <= 0 || g->kH <= 0 || g->kW <= 0 || g->sT <= 0 || g->sH <= 0 || g->sW <= 0) return SFB_ERR_BAD_ARG;
  if (g->pT < 0 || g->pH < 0 || g->pW < 0) return SFB_ERR_BAD_ARG;
  if ((g->g_ld & 7) || (g->g_coff & 7) || (g->o_ld & 7) || (g->o_coff & 7)) return SFB_ERR_BAD_ARG;
  if (g->g_coff + g->Cg > g->g_ld || g->o_coff + g->Co > g->o_ld) return SFB_ERR_BAD_ARG;
  const long long M = (long long)g->N * g->To * g->Ho * g->Wo;
  const long long Pin = (long long)g->N * g->Ti * g->Hi * g->Wi;
  if (M >= (1ll << 31) - 256 || Pin >= (1ll << 31)) return SFB_ERR_UNSUPPORTED;
  return SFB_OK;
}

// ---- 5-D activation tensor maps (C, W, H, T, N) with traversal strides = conv strides ----
std::map<std::array<long long, 16>, CUtensorMap> g_act_tmaps;

int ensure_encode() {
  if (g_encode) return SFB_OK;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaGetDriverEntryPoint(cuTensorMapEncodeTiled) failed");
    return SFB_ERR_DRIVER;
  }
  g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  return SFB_OK;
}

int get_act_tmap(const void* base, int C, int W, int H, int T, int N, int ld, int box_c, int bw, int bh, int bt, int sW,
                 int sH, int sT, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  std::array<long long, 16> key = {(long long)(uintptr_t)base, C, W, H, T, N, ld, box_c, bw, bh, bt, sW, sH, sT, 0, 0};
  auto it = g_act_tmaps.find(key);
  if (it != g_act_tmaps.end()) {
    *out = it->second;
    return SFB_OK;
  }
  int rc = ensure_encode();
  if (rc) return rc;
  CUtensorMap tm;
  cuuint64_t gdim[5] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)T, (cuuint64_t)N};
  cuuint64_t gstr[4] = {(cuuint64_t)ld * 2, (cuuint64_t)W * ld * 2, (cuuint64_t)H * W * ld * 2,
                        (cuuint64_t)T * H * W * ld * 2};
  cuuint32_t box[5] = {(cuuint32_t)box_c, (cuuint32_t)(bw * sW), (cuuint32_t)(bh * sH), (cuuint32_t)(bt * sT), 1};
  cuuint32_t estr[5] = {1, (cuuint32_t)sW, (cuuint32_t)sH, (cuuint32_t)sT, 1};
  const CUtensorMapSwizzle sw = box_c == 64 ? CU_TENSOR_MAP_SWIZZLE_128B
                                            : (box_c == 32 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                           : (box_c == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE));
  CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), gdim, gstr, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "cuTensorMapEncodeTiled(5d act C=%d W=%d H=%d T=%d N=%d ld=%d box=%d,%d,%d,%d) failed: %d",
             C, W, H, T, N, ld, box_c, bw * sW, bh * sH, bt * sT, (int)r);
    return SFB_ERR_DRIVER;
  }
  if (g_act_tmaps.size() > 8192) g_act_tmaps.clear();
  g_act_tmaps[key] = tm;
  *out = tm;
  return SFB_OK;
}


// ---- generic tiled tensor map (rank <= 5, element strides 1), cached ----
std::map<std::array<long long, 20>, CUtensorMap> g_gen_tmaps;
int get_tmap_generic(const void* base, int rank, const long long* dims, const long long* strides_bytes, const int* box,
                     int swizzle_bytes, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  std::array<long long, 20> key;
  key.fill(0);
  key[0] = (long long)(uintptr_t)base; key[1] = rank; key[2] = swizzle_bytes;
  for (int i = 0; i < rank; ++i) { key[3 + i] = dims[i]; key[8 + i] = i + 1 < rank ? strides_bytes[i] : 0; key[13 + i] = box[i]; }
  auto it = g_gen_tmaps.find(key);
  if (it != g_gen_tmaps.end()) { *out = it->second; return SFB_OK; }
  int rc = ensure_encode();
  if (rc) return rc;
  cuuint64_t gdim[5]; cuuint64_t gstr[4]; cuuint32_t bx[5]; cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = (cuuint64_t)dims[i]; bx[i] = (cuuint32_t)box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = (cuuint64_t)strides_bytes[i];
  const CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUtensorMap tm;
  CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "cuTensorMapEncodeTiled(generic rank %d, box0 %d, swizzle %d) failed: %d", rank,
             box[0], swizzle_bytes, (int)r);
    return SFB_ERR_DRIVER;
  }
  if (g_gen_tmaps.size() > 4096) g_gen_tmaps.clear();
  g_gen_tmaps[key] = tm;
  *out = tm;
  return SFB_OK;
}

constexpr size_t kMaxSmem = 227 * 1024;
int stem_bn(int Co) { return Co <= 16 ? 16 : (Co <= 32 ? 32 : 64); }
// Measured on B200 (tests/test_kernels_gpu.py::test_stem_s2d_*): tcgen05.mma applies the 32/64/128-byte swizzle XOR
// to ABSOLUTE shared-memory address bits, so a window whose start is shifted by whole pixels (not a multiple of the
// 8-row swizzle period) needs NO descriptor "base offset"; setting base_offset = (addr >> 7) & 7 gives wrong results.
int stem_base_off_mode() { return 0; }

template <int BN, int CS>
int launch_stem_fprop(const CUtensorMap& tx, const CUtensorMap& tw, const StemParams& p, int grid, cudaStream_t st) {
  const size_t smem = StemFpropCfg<BN>::smem_bytes(p.Cs);
  if (smem > kMaxSmem) return SFB_ERR_UNSUPPORTED;
  static size_t attr_bytes = 0;
  if (smem > attr_bytes) {
    cudaError_t e = cudaFuncSetAttribute(stem_fprop_kernel<BN, CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(stem_fprop)");
    attr_bytes = smem;
  }
  stem_fprop_kernel<BN, CS><<<grid, 192, smem, st>>>(tx, tw, p);
  SFB_CHECK_LAUNCH("stem_fprop_kernel");
  return SFB_OK;
}

template <int BR, int CS>
int launch_stem_fprop_rows(const CUtensorMap& tx, const CUtensorMap& tw, const StemParams& p, int grid, cudaStream_t st) {
  const size_t smem = StemFpropRowsCfg<BR>::smem_bytes(p.Cs);
  if (smem > kMaxSmem) return SFB_ERR_UNSUPPORTED;
  static size_t attr_bytes = 0;
  if (smem > attr_bytes) {
    cudaError_t e = cudaFuncSetAttribute(stem_fprop_rows_kernel<BR, CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(stem_fprop_rows)");
    attr_bytes = smem;
  }
  stem_fprop_rows_kernel<BR, CS><<<grid, 192, smem, st>>>(tx, tw, p);
  SFB_CHECK_LAUNCH("stem_fprop_rows_kernel");
  return SFB_OK;
}

template <int BN, int CS>
int launch_stem_wgrad(const CUtensorMap& tx, const CUtensorMap& tdy, const StemParams& p, int grid, cudaStream_t st) {
  const size_t smem = 1024 + (size_t)kStemSlots * 128 * p.Cs * 2 + 1024 + (size_t)StemDyRing<BN>::kSlots * 128 * BN * 2 + 512;
  if (smem > kMaxSmem || 4 * (p.Cs / 32) * BN > 256) return SFB_ERR_UNSUPPORTED;
  static size_t attr_bytes = 0;
  if (smem > attr_bytes) {
    cudaError_t e = cudaFuncSetAttribute(stem_wgrad_kernel<BN, CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(stem_wgrad)");
    attr_bytes = smem;
  }
  stem_wgrad_kernel<BN, CS><<<grid, 192, smem, st>>>(tx, tdy, p);
  SFB_CHECK_LAUNCH("stem_wgrad_kernel");
  return SFB_OK;
}

int stem_check(int32_t N, int32_t T, int32_t H2, int32_t W2, int32_t Ci, int32_t Cs, int32_t Co) {
  if (N <= 0 || T <= 0 || H2 <= 0 || W2 <= 0 || Ci <= 0 || Co <= 0 || (Co & 7)) return SFB_ERR_BAD_ARG;
  if (Cs != 32 && Cs != 64) return SFB_ERR_BAD_ARG;
  if (4 * Ci > Cs) return SFB_ERR_BAD_ARG;
  if (W2 > 125 || Co > 64) return SFB_ERR_UNSUPPORTED;
  if (Cs == 64 && Co > 32) return SFB_ERR_UNSUPPORTED;  // ring + weights would not fit 227 KB of shared memory
  if ((long long)N * T * H2 >= (1ll << 31)) return SFB_ERR_UNSUPPORTED;
  return SFB_OK;
}


// ---- strided dgrad classes (see elementwise.cuh) ----
// k/s/p in (T,H,W) order; O = dX dims (only needed for amin / na; pass nullptr when packing)
int build_dgrad_classes(const int* k, const int* s, const int* p, const int* O, int Cg, DgradClassTable* tb, int* k_pad) {
  if (s[0] * s[1] * s[2] > 16) return SFB_ERR_UNSUPPORTED;
  memset(tb, 0, sizeof(*tb));
  for (int d = 0; d < 3; ++d) { tb->s[d] = s[d]; tb->k[d] = k[d]; }
  int k0 = 0;
  for (int rt = 0; rt < s[0]; ++rt)
    for (int rh = 0; rh < s[1]; ++rh)
      for (int rw = 0; rw < s[2]; ++rw) {
        DgradClass& c = tb->c[tb->n++];
        const int r[3] = {rt, rh, rw};
        c.taps = 1;
        for (int d = 0; d < 3; ++d) {
          c.r[d] = r[d];
          c.nj[d] = r[d] < k[d] ? (k[d] - r[d] + s[d] - 1) / s[d] : 0;
          c.taps *= c.nj[d];
          if (p && O) {
            const int num = p[d] - r[d];
            c.amin[d] = num > 0 ? (num + s[d] - 1) / s[d] : 0;
            const int top = O[d] - 1 + p[d] - r[d];
            c.na[d] = top >= 0 ? top / s[d] - c.amin[d] + 1 : 0;
            if (c.na[d] < 0) c.na[d] = 0;
          }
        }
        c.k0 = k0;
        k0 += (c.taps * Cg + 63) / 64 * 64;
      }
  if (k_pad) *k_pad = k0 > 0 ? k0 : 64;
  return SFB_OK;
}

int g_num_sms = 0;
int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    g_num_sms = n > 0 ? n : 148;
  }
  return g_num_sms;
}

// output patch (Wt, Ht, Tt) with Wt*Ht*Tt <= 128 that wastes the fewest MMA rows; ties -> longest W run
void pick_patch(int Wo, int Ho, int To, int sW, int sH, int sT, int* Wt, int* Ht, int* Tt, int max_rows = 128) {
  double best = -1.0;
  int bw = 1, bh = 1, bt = 1;
  for (int w = 1; w <= Wo && w <= max_rows && w * sW <= 256; ++w) {
    const int tw = (Wo + w - 1) / w;
    for (int h = 1; h <= Ho && w * h <= max_rows && h * sH <= 256; ++h) {
      const int th = (Ho + h - 1) / h;
      int t = max_rows / (w * h);
      if (t > To) t = To;
      while (t * sT > 256) --t;
      // prefer a Tt that divides To
      int tbest = t;
      for (int c = t; c >= 1 && c > t / 2; --c)
        if (To % c == 0) { tbest = c; break; }
      for (int pass = 0; pass < 2; ++pass) {
        const int tcur = pass == 0 ? t : tbest;
        const int ttiles = (To + tcur - 1) / tcur;
        const double util = (double)Wo * Ho * To / ((double)tw * th * ttiles * (double)max_rows);
        const double score = util + 1e-4 * w / (double)max_rows;
        if (score > best) { best = score; bw = w; bh = h; bt = tcur; }
      }
    }
  }
  *Wt = bw; *Ht = bh; *Tt = bt;
}

// bytes between the pipeline stages of a halo-mode launch: 168 pixel rows of 128 bytes (a box of <= 10 x 16 pixels plus
// the two rows the last taps read past it)
constexpr uint32_t kHaloStride = 168 * 128;
inline bool halo_enabled() {  // SFB_HALO=0: tap-by-tap boxes for those layers too (read per call: tests compare both)
  const char* e = getenv("SFB_HALO");
  return !(e && *e == '0');
}

inline int side_depth() {  // SFB_SIDE_DEPTH=1: one side tile (read per call, like SFB_HALO)
  const char* e = getenv("SFB_SIDE_DEPTH");
  return e && *e ? atoi(e) : 2;
}

inline bool merge_hw_enabled() {
  const char* e = getenv("SFB_MERGE_HW");
  return !(e && *e == '0');
}
inline bool pair_enabled() {  // SFB_PAIR=0: single-CTA MMAs for the streamed-weight 256-wide layers too
  const char* e = getenv("SFB_PAIR");
  return !(e && *e == '0');
}

template <int BN, bool BRES, int MODE, bool PAIR = false>
int launch_conv_tma_impl(const CUtensorMap& ta, const CUtensorMap& tw, const CUtensorMap& to, const CUtensorMap& ty,
                         const ConvTmaParams& p, cudaStream_t st) {
  using L = ConvTmaCfg<BN, BRES, PAIR>;
  const size_t smem = L::smem_bytes(p.nk);
  // Shared memory behind the kernel's fixed layout holds up to three more swizzled tiles (1024-byte aligned): a second
  // staging tile (TMA stores double-buffered) and one or two side tiles (y of the fused BatchNorm backward / residual
  // of the inference epilogue).  What does not fit there takes the shared memory of the last pipeline stages.
  ConvTmaParams pp = p;
  size_t smem_total = smem;
  const size_t base = (smem - 1024 + 1023) / 1024 * 1024;
  const size_t tile_bytes = L::kStagingBytes;
  const int n_tail = 1024 + base >= 227 * 1024 ? 0 : (int)((227 * 1024 - 1024 - base) / tile_bytes);
  pp.nstages = L::kStages;
  pp.stg2_off = pp.ybuf_off = pp.ybuf2_off = 0;
  const bool side_tile = p.bn_mode || p.res_tile;
  int used = 0;
  auto tail_slot = [&]() -> uint32_t { return used < n_tail ? (uint32_t)(base + (size_t)(used++) * tile_bytes) : 0u; };
  auto steal = [&](int min_left) -> uint32_t {
    const int drop = (int)((tile_bytes + L::kStage - 1) / L::kStage);
    if (pp.nstages - drop < min_left) return 0u;
    pp.nstages -= drop;
    return (uint32_t)(pp.nstages * L::kStage);
  };
  if (p.halo) {
    // three stages of kHaloStride bytes at the front of the pipeline area; the rest of it holds the extra tiles
    if (!BRES || 3 * (size_t)kHaloStride + 3 * tile_bytes > (size_t)L::kStagingOff) return SFB_ERR_UNSUPPORTED;
    pp.nstages = 3;
    pp.stg2_off = 3 * kHaloStride;
    if (side_tile) {
      pp.ybuf_off = pp.stg2_off + (uint32_t)tile_bytes;
      if (side_depth() >= 2) pp.ybuf2_off = pp.ybuf_off + (uint32_t)tile_bytes;
    }
  } else {
  pp.stg2_off = tail_slot();
  if (side_tile) {
    // with two side tiles the requests run two chunks ahead (ConvTmaParams::ybuf2_off); the fused dgrads keep at least
    // two tiles of operands in flight (4 stages).  SFB_SIDE_DEPTH=1: one side tile, the arrangement of earlier builds.
    const bool deep = side_depth() >= 2;
    // the residual convs of the inference epilogue write 4x the columns they read: a tile's epilogue outlasts the
    // operand loads of the next one several times over, so one tile of operands in flight is enough there
    const int keep = MODE == kEpiAffine ? std::max(2, p.nk) : std::max(4, 2 * p.nk);
    if (deep) {
      pp.ybuf_off = tail_slot();
      if (!pp.ybuf_off) pp.ybuf_off = steal(keep);
      if (pp.ybuf_off) {
        pp.ybuf2_off = tail_slot();
        if (!pp.ybuf2_off) pp.ybuf2_off = steal(keep);
      }
    } else {
      pp.ybuf_off = n_tail >= 2 ? tail_slot() : 0u;
    }
    if (!pp.ybuf_off) {
      if (pp.stg2_off) {
        pp.ybuf_off = pp.stg2_off;  // the second staging tile holds the side tile (stores are single-buffered then)
      } else {
        pp.ybuf_off = steal(2);
        if (!pp.ybuf_off) return SFB_ERR_UNSUPPORTED;
      }
    }
  }
  }
  if (used) smem_total = 1024 + base + (size_t)used * tile_bytes;
  // one N tile per CTA: grid = (CTAs per N tile) * n_tiles, at most one CTA per SM
  // (pair: one N tile per CTA PAIR, the pair walks neighbouring patches)
  const int units = PAIR ? num_sms() / 2 : num_sms();
  const int unit_patches = PAIR ? (p.total_patches + 1) / 2 : p.total_patches;
  long long per_nt = units / p.n_tiles;
  if (per_nt < 1) per_nt = 1;
  if (per_nt > unit_patches) per_nt = unit_patches;
  const int grid = (int)per_nt * p.n_tiles * (PAIR ? 2 : 1);
  static size_t attr_bytes = 0;
  if (smem_total > attr_bytes) {
    cudaError_t e = cudaFuncSetAttribute(conv_tma_kernel<BN, BRES, MODE, PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_total);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_tma)");
    attr_bytes = smem_total;
  }
  if (PAIR) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid, 1, 1);
    cfg.blockDim = dim3((unsigned)L::kThreads, 1, 1);
    cfg.dynamicSmemBytes = smem_total;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv_tma_kernel<BN, BRES, MODE, PAIR>, ta, tw, to, ty, pp);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(conv_tma_kernel, CTA pair)");
    return SFB_OK;
  }
  conv_tma_kernel<BN, BRES, MODE, PAIR><<<grid, L::kThreads, smem_total, st>>>(ta, tw, to, ty, pp);
  SFB_CHECK_LAUNCH("conv_tma_kernel");
  return SFB_OK;
}

template <int BN>
int launch_conv_tma(const CUtensorMap& ta, const CUtensorMap& tw, const CUtensorMap& to, const CUtensorMap& ty,
                    const ConvTmaParams& p, cudaStream_t st) {
  // weights resident in shared memory when the CTA's [BN][Ktot] slice fits next to the pipeline (see ConvTmaCfg)
  const bool bres = ConvTmaCfg<BN, true>::smem_bytes(p.nk) <= 227 * 1024;
  if constexpr (BN == 256) {
    if (p.pair && !bres) {
      if (p.bn_mode) {
        if (p.epi_scale || p.res || p.epi_relu || p.stats) return SFB_ERR_UNSUPPORTED;
        return launch_conv_tma_impl<256, false, kEpiBnBwd, true>(ta, tw, to, ty, p, st);
      }
      if (p.epi_scale || p.res || p.epi_relu) {
        if (p.stats) return SFB_ERR_UNSUPPORTED;
        return launch_conv_tma_impl<256, false, kEpiAffine, true>(ta, tw, to, ty, p, st);
      }
      return launch_conv_tma_impl<256, false, kEpiPlain, true>(ta, tw, to, ty, p, st);
    }
  }
  if (p.bn_mode) {
    if (p.epi_scale || p.res || p.epi_relu || p.stats) return SFB_ERR_UNSUPPORTED;
    return bres ? launch_conv_tma_impl<BN, true, kEpiBnBwd>(ta, tw, to, ty, p, st)
                : launch_conv_tma_impl<BN, false, kEpiBnBwd>(ta, tw, to, ty, p, st);
  }
  if (p.epi_scale || p.res || p.epi_relu) {
    if (p.stats) return SFB_ERR_UNSUPPORTED;  // forward statistics go with the raw training output only
    return bres ? launch_conv_tma_impl<BN, true, kEpiAffine>(ta, tw, to, ty, p, st)
                : launch_conv_tma_impl<BN, false, kEpiAffine>(ta, tw, to, ty, p, st);
  }
  return bres ? launch_conv_tma_impl<BN, true, kEpiPlain>(ta, tw, to, ty, p, st)
              : launch_conv_tma_impl<BN, false, kEpiPlain>(ta, tw, to, ty, p, st);
}

// ConvTmaParams fields of the fused BatchNorm-backward epilogue (sfb_conv_dgrad_bn)
void set_bn_fuse(ConvTmaParams& q, const SfbBnBwdFuse* bn, int Co) {
  if (!bn) return;
  q.bn_mode = 1;
  q.bn_scale = bn->scale; q.bn_shift = bn->shift; q.bn_mean = bn->mean; q.bn_invstd = bn->invstd;
  q.bn_dgamma = bn->dgamma; q.bn_dbeta = bn->dbeta;
  q.bn_C = bn->C;
  q.bn_ld = bn->fold_ld > 0 ? bn->fold_ld : Co;
}

// Can the TMA kernel take this conv?  (everything else stays on the register-gather kernel)
bool tma_eligible(const SfbConvGeom* g, int taps) {
  if (g->Cg % 16) return false;
  if (g->Cg > 64 && (g->Cg % 64) && taps != 1) return false;
  if (g->Cg < 64 && g->Cg != 16 && g->Cg != 32) return false;
  if (g->transposed && !(g->sT == 1 && g->sH == 1 && g->sW == 1)) return false;
  if (g->Co > 2048) return false;
  return true;
}

template <int BN, int STAGES>
int launch_conv_tc(const CUtensorMap& tm, const ConvTcParams& p, cudaStream_t st) {
  using L = ConvTcSmem<BN, STAGES>;
  const size_t smem = L::bytes(p.g.Cg, p.pro_scale != nullptr);
  static bool attr_done = false;
  static size_t attr_bytes = 0;
  if (!attr_done || smem > attr_bytes) {
    cudaError_t e = cudaFuncSetAttribute(conv_gemm_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(L::bytes(2048, true)));
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_gemm_tc)");
    attr_done = true;
    attr_bytes = L::bytes(2048, true);
  }
  dim3 grid((p.M + kTileM - 1) / kTileM, (p.g.Co + BN - 1) / BN);
  conv_gemm_tc_kernel<BN, STAGES><<<grid, kProducerThreads + 64, smem, st>>>(tm, p);
  SFB_CHECK_LAUNCH("conv_gemm_tc_kernel");
  return SFB_OK;
}

}  // namespace

extern "C" {

const char* sfb_strerror(int status) {
  switch (status) {
    case SFB_OK: return "ok";
    case SFB_ERR_BAD_ARG: return "bad argument (shape, alignment or null pointer)";
    case SFB_ERR_CUDA: return "CUDA error (see sfb_last_cuda_error)";
    case SFB_ERR_UNSUPPORTED: return "unsupported configuration";
    case SFB_ERR_DRIVER: return "driver entry point unavailable (no CUDA driver?)";
  }
  return "unknown status";
}
const char* sfb_last_cuda_error(void) { return g_cuda_err; }
int sfb_version(void) { return 100; }
int64_t sfb_launch_count(void) { return (int64_t)g_launches.load(); }
int64_t sfb_launch_count_add(int64_t n) { return (int64_t)(g_launches.fetch_add(n > 0 ? n : 0) + (n > 0 ? n : 0)); }

int sfb_packed_weight_dims(int32_t Co, int32_t Cg, int32_t taps, int32_t* rows_pad, int32_t* k_pad) {
  if (Co <= 0 || Cg <= 0 || (Cg & 7) || taps <= 0 || !rows_pad || !k_pad) return SFB_ERR_BAD_ARG;
  const int bn = pick_bn(Co);
  *rows_pad = (Co + bn - 1) / bn * bn;
  *k_pad = (taps * Cg + 63) / 64 * 64;
  return SFB_OK;
}

int sfb_pack_weight(const float* w, int32_t Co, int32_t Ci, int32_t taps, int32_t Cg, int32_t transposed, void* wp,
                    int32_t rows_pad, int32_t k_pad, void* stream) {
  if (!w || !wp || Co <= 0 || Ci <= 0 || taps <= 0 || (Cg & 7) || (k_pad & 63) || k_pad < taps * Cg) return SFB_ERR_BAD_ARG;
  if (!transposed && (Cg < Ci || rows_pad < Co)) return SFB_ERR_BAD_ARG;
  if (transposed && (Cg < Co || rows_pad < Ci)) return SFB_ERR_BAD_ARG;
  const long long total = (long long)rows_pad * k_pad;
  pack_weight_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(w, Co, Ci, taps, Cg, transposed,
                                                                             (__nv_bfloat16*)wp, rows_pad, k_pad);
  SFB_CHECK_LAUNCH("pack_weight_kernel");
  return SFB_OK;
}

int sfb_layout_input(const float* x, int64_t sn, int64_t sc, int64_t st, int64_t sh, int64_t sw, int32_t N, int32_t C,
                     int32_t T, int32_t H, int32_t W, void* out, int32_t out_ld, void* stream) {
  if (!x || !out || N <= 0 || C <= 0 || T <= 0 || H <= 0 || W <= 0 || (out_ld & 7) || out_ld < C || !aligned16(out))
    return SFB_ERR_BAD_ARG;
  const long long total = (long long)N * T * H * W;
  layout_input_kernel<<<grid_for(total, 256, 148 * 32), 256, 0, (cudaStream_t)stream>>>(
      x, sn, sc, st, sh, sw, N, C, T, H, W, (__nv_bfloat16*)out, out_ld);
  SFB_CHECK_LAUNCH("layout_input_kernel");
  return SFB_OK;
}



int sfb_packed_weight_dims_dgrad(int32_t Ci, int32_t Cg, int32_t kT, int32_t kH, int32_t kW, int32_t sT, int32_t sH,
                                 int32_t sW, int32_t* rows_pad, int32_t* k_pad) {
  if (Ci <= 0 || Cg <= 0 || (Cg & 7) || kT <= 0 || kH <= 0 || kW <= 0 || sT <= 0 || sH <= 0 || sW <= 0 || !rows_pad || !k_pad)
    return SFB_ERR_BAD_ARG;
  const int k[3] = {kT, kH, kW}, s[3] = {sT, sH, sW};
  DgradClassTable tb;
  int kp = 0;
  int rc = build_dgrad_classes(k, s, nullptr, nullptr, Cg, &tb, &kp);
  if (rc) return rc;
  const int bn = pick_bn(Ci);
  *rows_pad = (Ci + bn - 1) / bn * bn;
  *k_pad = kp;
  return SFB_OK;
}

int sfb_pack_weight_dgrad(const float* w, int32_t Co, int32_t Ci, int32_t kT, int32_t kH, int32_t kW, int32_t sT,
                          int32_t sH, int32_t sW, int32_t Cg, void* wp, int32_t rows_pad, int32_t k_pad, void* stream) {
  if (!w || !wp || Co <= 0 || Ci <= 0 || (Cg & 7) || Cg < Co || rows_pad < Ci) return SFB_ERR_BAD_ARG;
  const int k[3] = {kT, kH, kW}, s[3] = {sT, sH, sW};
  DgradClassTable tb;
  int kp = 0;
  int rc = build_dgrad_classes(k, s, nullptr, nullptr, Cg, &tb, &kp);
  if (rc) return rc;
  if (k_pad != kp) return SFB_ERR_BAD_ARG;
  const long long total = (long long)rows_pad * k_pad;
  pack_weight_dgrad_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(w, Co, Ci, Cg, tb, (__nv_bfloat16*)wp,
                                                                                   rows_pad, k_pad);
  SFB_CHECK_LAUNCH("pack_weight_dgrad_kernel");
  return SFB_OK;
}


static bool fold_args_ok(int Co, int Ci, int kT, int kH, int kW, int f, int ldo, int ldi) {
  return Co > 0 && Ci > 0 && kT > 0 && kH > 0 && (kW == 1 || kW == 3) && f >= 2 && f <= 8 && ldo >= Co && ldi >= Ci &&
         !(ldo & 7) && !(ldi & 7);
}

int sfb_fold_weight(const float* w, int32_t Co, int32_t Ci, int32_t kT, int32_t kH, int32_t kW, int32_t f, int32_t ldo,
                    int32_t ldi, float* wf, void* stream) {
  if (!w || !wf || !fold_args_ok(Co, Ci, kT, kH, kW, f, ldo, ldi)) return SFB_ERR_BAD_ARG;
  const long long total = (long long)f * ldo * f * ldi * kT * kH * (kW == 1 ? 1 : 3);
  fold_weight_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(w, Co, Ci, kT, kH, kW, f, ldo, ldi, wf);
  SFB_CHECK_LAUNCH("fold_weight_kernel");
  return SFB_OK;
}

int sfb_unfold_wgrad(const float* dwf, int32_t Co, int32_t Ci, int32_t kT, int32_t kH, int32_t kW, int32_t f, int32_t ldo,
                     int32_t ldi, float* dw, void* stream) {
  if (!dwf || !dw || !fold_args_ok(Co, Ci, kT, kH, kW, f, ldo, ldi)) return SFB_ERR_BAD_ARG;
  const long long total = (long long)Co * Ci * kT * kH * kW;
  unfold_wgrad_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(dwf, Co, Ci, kT, kH, kW, f, ldo, ldi, dw);
  SFB_CHECK_LAUNCH("unfold_wgrad_kernel");
  return SFB_OK;
}

int sfb_unfold_stats(const float* stats_folded, int32_t f, int32_t ld, int32_t C, float* stats, void* stream) {
  if (!stats_folded || !stats || f < 2 || f > 8 || C <= 0 || ld < C) return SFB_ERR_BAD_ARG;
  unfold_stats_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(stats_folded, f, ld, C, stats);
  SFB_CHECK_LAUNCH("unfold_stats_kernel");
  return SFB_OK;
}

int sfb_stem_s2d_channels(int32_t Ci) {
  if (Ci <= 0 || 4 * Ci > 64) return SFB_ERR_UNSUPPORTED;
  return 4 * Ci <= 32 ? 32 : 64;
}

int sfb_layout_input_s2d(const float* x, int64_t sn, int64_t sc, int64_t st, int64_t sh, int64_t sw, int32_t N, int32_t C,
                         int32_t T, int32_t H, int32_t W, void* out, int32_t Cs, void* stream) {
  if (!x || !out || N <= 0 || C <= 0 || T <= 0 || H <= 0 || W <= 0 || (H & 1) || (W & 1) || !aligned16(out)) return SFB_ERR_BAD_ARG;
  if ((Cs != 32 && Cs != 64) || 4 * C > Cs) return SFB_ERR_BAD_ARG;
  const long long total = (long long)N * T * (H / 2) * (W / 2);
  // W-contiguous clips (what _prepare_slowfast_data hands over) with the reference's channel counts: vector path
  const bool vec = sw == 1 && ((uintptr_t)x & 7) == 0 && !((sn | sc | st | sh) & 1);
  const int grid = grid_for(total, 128, 148 * 64);
  if (vec && C == 5 && Cs == 32)
    layout_input_s2d_vec_kernel<5, 32><<<grid, 128, 0, (cudaStream_t)stream>>>(x, sn, sc, st, sh, N, T, H / 2, W / 2, (__nv_bfloat16*)out);
  else if (vec && C == 15 && Cs == 64)
    layout_input_s2d_vec_kernel<15, 64><<<grid, 128, 0, (cudaStream_t)stream>>>(x, sn, sc, st, sh, N, T, H / 2, W / 2, (__nv_bfloat16*)out);
  else
    layout_input_s2d_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(
        x, sn, sc, st, sh, sw, N, C, T, H / 2, W / 2, (__nv_bfloat16*)out, Cs);
  SFB_CHECK_LAUNCH("layout_input_s2d_kernel");
  return SFB_OK;
}

int sfb_stage_clips_u8(const uint8_t* frames, int32_t N, int32_t T, int32_t H, int32_t W, int32_t Ct,
                       const int32_t* crop_yx, float mean, float std, const SfbStagePathway* pw, int32_t n_pw,
                       void* stream) {
  if (!frames || !pw || N <= 0 || T <= 0 || H <= 0 || W <= 0 || Ct <= 0 || n_pw <= 0) return SFB_ERR_BAD_ARG;
  if (n_pw > kStageMaxPw || Ct > 64 || !(std > 0.f) || !aligned16(frames)) return SFB_ERR_BAD_ARG;
  if ((H & 1) || (W & 1)) return SFB_ERR_UNSUPPORTED;  // the s2d stems need even frames (as sfb_layout_input_s2d)
  StageParams p;
  memset(&p, 0, sizeof(p));
  for (int q = 0; q < n_pw; ++q) {
    const SfbStagePathway& s = pw[q];
    if (!s.out || !aligned16(s.out) || s.C <= 0 || s.c0 < 0 || s.c0 + s.C > Ct || s.t_stride <= 0 || (T % s.t_stride))
      return SFB_ERR_BAD_ARG;
    if ((s.Cs != 8 && s.Cs != 16 && s.Cs != 32 && s.Cs != 64) || 4 * s.C > s.Cs) return SFB_ERR_BAD_ARG;
    p.pw[q].c0 = s.c0; p.pw[q].C = s.C; p.pw[q].t_stride = s.t_stride; p.pw[q].Cs = s.Cs;
    p.pw[q].out = (__nv_bfloat16*)s.out;
  }
  p.frames = frames;
  p.total_bytes = (long long)N * T * H * W * Ct;
  p.N = N; p.T = T; p.H = H; p.W = W; p.Ct = Ct;
  p.crop_yx = crop_yx; p.mean = mean; p.std = std; p.n_pw = n_pw;
  // rows of a pair 16 bytes off a multiple of 128: the byte gathers of the two rows then fall on interleaved banks
  // (measured 1.83 -> 1.76 ms on the 64 x 32 x 224^2 x 21 batch)
  p.row_bytes = ((W * Ct + 32 + 127) & ~127) + 16;
  const size_t smem = 256 * 32 * 4 + kStageMaxPw * 8 * 16 + 4 * (size_t)p.row_bytes;  // two buffers of a row pair
  if (smem > kMaxSmem) return SFB_ERR_UNSUPPORTED;
  static size_t attr_bytes = 48 * 1024;
  if (smem > attr_bytes) {
    cudaError_t e = cudaFuncSetAttribute(stage_clips_u8_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(stage_clips_u8_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(stage_clips_u8)");
    attr_bytes = smem;
  }
  const long long n_rows = (long long)N * T * (H / 2);
  if (n_rows + 148 * 8 >= (1ll << 31)) return SFB_ERR_UNSUPPORTED;
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(4, (220 * 1024) / (smem + 1024)));
  const int grid = (int)std::min<long long>(n_rows, (long long)num_sms() * per_sm);
  if (crop_yx)
    stage_clips_u8_kernel<true><<<grid, kStageThreads, smem, (cudaStream_t)stream>>>(p);
  else
    stage_clips_u8_kernel<false><<<grid, kStageThreads, smem, (cudaStream_t)stream>>>(p);
  SFB_CHECK_LAUNCH("stage_clips_u8_kernel");
  return SFB_OK;
}

int sfb_pack_weight_stem(const float* w, int32_t Co, int32_t Ci, int32_t Cs, void* wp, int32_t rows_pad, void* stream) {
  if (!w || !wp || Co <= 0 || Ci <= 0 || (Cs != 32 && Cs != 64) || 4 * Ci > Cs || rows_pad < Co) return SFB_ERR_BAD_ARG;
  const long long total = (long long)rows_pad * 16 * Cs;
  pack_weight_stem_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(w, Co, Ci, Cs, (__nv_bfloat16*)wp, rows_pad);
  SFB_CHECK_LAUNCH("pack_weight_stem_kernel");
  return SFB_OK;
}

int64_t sfb_pack_table_bytes(int32_t n_jobs) { return n_jobs > 0 ? (int64_t)n_jobs * (int64_t)sizeof(PackJobDev) : 0; }

int sfb_pack_plan(const SfbPackJob* jobs, int32_t n_jobs, void* table_dev, int64_t table_bytes, int32_t* total_blocks) {
  if (!jobs || n_jobs <= 0 || !table_dev || !total_blocks || table_bytes < sfb_pack_table_bytes(n_jobs)) return SFB_ERR_BAD_ARG;
  std::vector<PackJobDev> tab((size_t)n_jobs);
  int block0 = 0;
  for (int i = 0; i < n_jobs; ++i) {
    const SfbPackJob& j = jobs[i];
    PackJobDev& d = tab[(size_t)i];
    memset(&d, 0, sizeof(d));
    if (!j.src || !j.dst || j.Co <= 0 || j.Ci <= 0) return SFB_ERR_BAD_ARG;
    d.kind = j.kind; d.Co = j.Co; d.Ci = j.Ci; d.kT = j.kT; d.kH = j.kH; d.kW = j.kW; d.Cg = j.Cg;
    d.transposed = j.transposed; d.f = j.f; d.ldo = j.ldo; d.ldi = j.ldi; d.rows_pad = j.rows_pad; d.k_pad = j.k_pad;
    d.src = j.src; d.dst = j.dst;
    long long total = 0;
    switch (j.kind) {
      case 0:
        if (j.kT <= 0 || j.kH <= 0 || j.kW <= 0 || j.Cg <= 0 || j.rows_pad <= 0 || j.k_pad <= 0) return SFB_ERR_BAD_ARG;
        total = (long long)j.rows_pad * j.k_pad;
        break;
      case 1: {
        if ((j.Cg & 7) || j.Cg < j.Co || j.rows_pad < j.Ci) return SFB_ERR_BAD_ARG;
        const int k[3] = {j.kT, j.kH, j.kW}, s[3] = {j.sT, j.sH, j.sW};
        int kp = 0;
        int rc = build_dgrad_classes(k, s, nullptr, nullptr, j.Cg, &d.tb, &kp);
        if (rc) return rc;
        if (kp != j.k_pad) return SFB_ERR_BAD_ARG;
        total = (long long)j.rows_pad * j.k_pad;
        break;
      }
      case 2:
        if (!fold_args_ok(j.Co, j.Ci, j.kT, j.kH, j.kW, j.f, j.ldo, j.ldi)) return SFB_ERR_BAD_ARG;
        total = (long long)j.f * j.ldo * j.f * j.ldi * j.kT * j.kH * (j.kW == 1 ? 1 : 3);
        break;
      case 3:
        if ((j.Cg != 32 && j.Cg != 64) || 4 * j.Ci > j.Cg || j.rows_pad < j.Co) return SFB_ERR_BAD_ARG;
        total = (long long)j.rows_pad * 16 * j.Cg;
        break;
      default:
        return SFB_ERR_BAD_ARG;
    }
    long long nb = (total + 256ll * 8 - 1) / (256ll * 8);  // ~8 elements per thread
    if (nb < 1) nb = 1;
    if (nb > 512) nb = 512;
    d.block0 = block0;
    d.nblocks = (int)nb;
    block0 += (int)nb;
  }
  cudaError_t e = cudaMemcpy(table_dev, tab.data(), (size_t)n_jobs * sizeof(PackJobDev), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy(pack table)");
  *total_blocks = block0;
  return SFB_OK;
}

int sfb_pack_run(const void* table_dev, int32_t n_jobs, int32_t total_blocks, void* stream) {
  if (!table_dev || n_jobs <= 0 || total_blocks <= 0) return SFB_ERR_BAD_ARG;
  pack_batch_kernel<<<total_blocks, 256, 0, (cudaStream_t)stream>>>((const PackJobDev*)table_dev, n_jobs);
  SFB_CHECK_LAUNCH("pack_batch_kernel");
  return SFB_OK;
}

int64_t sfb_adam_table_bytes(int32_t n_jobs) { return n_jobs > 0 ? (int64_t)n_jobs * (int64_t)sizeof(AdamJobDev) : 0; }

int sfb_adam_plan(const SfbAdamJob* jobs, int32_t n_jobs, void* table_dev, int64_t table_bytes, int32_t* total_blocks) {
  if (!jobs || n_jobs <= 0 || !table_dev || !total_blocks || table_bytes < sfb_adam_table_bytes(n_jobs)) return SFB_ERR_BAD_ARG;
  std::vector<AdamJobDev> tab((size_t)n_jobs);
  long long block0 = 0;
  for (int i = 0; i < n_jobs; ++i) {
    const SfbAdamJob& j = jobs[i];
    if (!j.param || !j.grad || !j.exp_avg || !j.exp_avg_sq || j.n <= 0) return SFB_ERR_BAD_ARG;
    if (((uintptr_t)j.param | (uintptr_t)j.grad | (uintptr_t)j.exp_avg | (uintptr_t)j.exp_avg_sq) & 3) return SFB_ERR_BAD_ARG;
    AdamJobDev& d = tab[(size_t)i];
    d.p = j.param; d.g = j.grad; d.m = j.exp_avg; d.v = j.exp_avg_sq; d.n = j.n;
    d.vec = aligned16(j.param) && aligned16(j.grad) && aligned16(j.exp_avg) && aligned16(j.exp_avg_sq);
    d.block0 = (int)block0;
    block0 += (j.n + kAdamChunk - 1) / kAdamChunk;
    if (block0 > 0x7fffffffll) return SFB_ERR_BAD_ARG;
  }
  cudaError_t e = cudaMemcpy(table_dev, tab.data(), (size_t)n_jobs * sizeof(AdamJobDev), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy(adam table)");
  *total_blocks = (int)block0;
  return SFB_OK;
}

int sfb_adam_step(const void* table_dev, int32_t n_jobs, int32_t total_blocks, int64_t step, double lr, double beta1,
                  double beta2, double eps, double weight_decay, double grad_scale, void* stream) {
  if (!table_dev || n_jobs <= 0 || total_blocks <= 0 || step < 1) return SFB_ERR_BAD_ARG;
  if (!(lr >= 0.0) || !(eps >= 0.0) || !(beta1 >= 0.0 && beta1 < 1.0) || !(beta2 >= 0.0 && beta2 < 1.0) || !(weight_decay >= 0.0))
    return SFB_ERR_BAD_ARG;
  // scalars in double, rounded to fp32 once (torch passes Python doubles to its fp32 kernels the same way)
  const double bc1 = 1.0 - std::pow(beta1, (double)step), bc2 = 1.0 - std::pow(beta2, (double)step);
  AdamHyper h;
  h.one_minus_beta1 = (float)(1.0 - beta1);
  h.beta2 = (float)beta2;
  h.one_minus_beta2 = (float)(1.0 - beta2);
  h.eps = (float)eps;
  h.weight_decay = (float)weight_decay;
  h.neg_step_size = (float)(-(lr / bc1));
  h.bc2_sqrt = (float)std::sqrt(bc2);
  h.grad_scale = (float)grad_scale;
  adam_step_kernel<<<total_blocks, kAdamThreads, 0, (cudaStream_t)stream>>>((const AdamJobDev*)table_dev, n_jobs, h);
  SFB_CHECK_LAUNCH("adam_step_kernel");
  return SFB_OK;
}

int sfb_stem_conv_fprop(const void* xs2d, int32_t N, int32_t T, int32_t H2, int32_t W2, int32_t Ci, int32_t Cs,
                        const void* wp, int32_t rows_pad, void* out, int32_t o_ld, int32_t o_coff, int32_t Co,
                        float* stats, void* stream) {
  int rc = stem_check(N, T, H2, W2, Ci, Cs, Co);
  if (rc) return rc;
  if (!xs2d || !wp || !out || !aligned16(xs2d) || !aligned16(wp) || !aligned16(out)) return SFB_ERR_BAD_ARG;
  if ((o_ld & 7) || (o_coff & 7) || o_coff + Co > o_ld) return SFB_ERR_BAD_ARG;
  const int bn = stem_bn(Co);
  if (rows_pad != bn) return SFB_ERR_BAD_ARG;
  const long long F = (long long)N * T;
  CUtensorMap tx, tw;
  const long long xd[4] = {Cs, W2, H2, F};
  const long long xs[3] = {(long long)Cs * 2, (long long)W2 * Cs * 2, (long long)H2 * W2 * Cs * 2};
  const int xb[4] = {Cs, 128, 1, 1};
  rc = get_tmap_generic(xs2d, 4, xd, xs, xb, Cs * 2, &tx);
  if (rc) return rc;
  // narrow outputs: four output rows per tile (stem_fprop_rows_kernel), row blocks of br channels
  const int br = Co <= 8 ? 8 : (Co <= 16 && Cs == 32 ? 16 : 0);
  const long long wd[2] = {16ll * Cs, bn};
  const long long ws[1] = {16ll * Cs * 2};
  const int wb[2] = {64, br ? br : bn};
  rc = get_tmap_generic(wp, 2, wd, ws, wb, 128, &tw);
  if (rc) return rc;
  StemParams p;
  memset(&p, 0, sizeof(p));
  p.F = (int)F; p.Ho = H2; p.Wo = W2; p.Cs = Cs; p.Co = Co; p.o_ld = o_ld; p.o_coff = o_coff; p.Ci = Ci;
  p.total_tiles = br ? (int)(F * ((H2 + kStemRowsF - 1) / kStemRowsF)) : (int)(F * H2);
  int grid = num_sms();
  if (grid > p.total_tiles) grid = p.total_tiles;
  p.tiles_per_cta = (p.total_tiles + grid - 1) / grid;
  grid = (p.total_tiles + p.tiles_per_cta - 1) / p.tiles_per_cta;
  p.base_off_mode = stem_base_off_mode();
  p.out = (__nv_bfloat16*)out; p.stats = stats;
  cudaStream_t st = (cudaStream_t)stream;
  if (br == 8) return Cs == 64 ? launch_stem_fprop_rows<8, 64>(tx, tw, p, grid, st) : launch_stem_fprop_rows<8, 32>(tx, tw, p, grid, st);
  if (br == 16) return launch_stem_fprop_rows<16, 32>(tx, tw, p, grid, st);
  if (Cs == 64) return bn == 16 ? launch_stem_fprop<16, 64>(tx, tw, p, grid, st) : launch_stem_fprop<32, 64>(tx, tw, p, grid, st);
  switch (bn) {
    case 16: return launch_stem_fprop<16, 32>(tx, tw, p, grid, st);
    case 32: return launch_stem_fprop<32, 32>(tx, tw, p, grid, st);
    default: return launch_stem_fprop<64, 32>(tx, tw, p, grid, st);
  }
}

int sfb_stem_conv_wgrad(const void* xs2d, int32_t N, int32_t T, int32_t H2, int32_t W2, int32_t Ci, int32_t Cs,
                        const void* dy, int32_t dy_ld, int32_t dy_coff, int32_t Co, float* dw, void* stream) {
  int rc = stem_check(N, T, H2, W2, Ci, Cs, Co);
  if (rc) return rc;
  if (!xs2d || !dy || !dw || !aligned16(xs2d) || !aligned16(dy)) return SFB_ERR_BAD_ARG;
  if ((dy_ld & 7) || (dy_coff & 7) || dy_coff + Co > dy_ld) return SFB_ERR_BAD_ARG;
  const int bn = stem_bn(Co);
  const long long F = (long long)N * T;
  CUtensorMap tx, tdy;
  const long long xd[4] = {Cs, W2, H2, F};
  const long long xs[3] = {(long long)Cs * 2, (long long)W2 * Cs * 2, (long long)H2 * W2 * Cs * 2};
  const int xb[4] = {Cs, 128, 1, 1};
  rc = get_tmap_generic(xs2d, 4, xd, xs, xb, Cs * 2, &tx);
  if (rc) return rc;
  const long long dd[4] = {Co, W2, H2, F};
  const long long ds[3] = {(long long)dy_ld * 2, (long long)W2 * dy_ld * 2, (long long)H2 * W2 * dy_ld * 2};
  const int db[4] = {bn, 128, 1, 1};
  rc = get_tmap_generic((const __nv_bfloat16*)dy + dy_coff, 4, dd, ds, db, bn * 2, &tdy);
  if (rc) return rc;
  StemParams p;
  memset(&p, 0, sizeof(p));
  p.F = (int)F; p.Ho = H2; p.Wo = W2; p.Cs = Cs; p.Co = Co; p.Ci = Ci;
  p.total_tiles = (int)(F * H2);
  int grid = num_sms();
  if (grid > p.total_tiles) grid = p.total_tiles;
  p.tiles_per_cta = (p.total_tiles + grid - 1) / grid;
  grid = (p.total_tiles + p.tiles_per_cta - 1) / p.tiles_per_cta;
  p.ksteps = (W2 + 15) / 16;
  p.base_off_mode = stem_base_off_mode();
  p.dW = dw;
  cudaStream_t st = (cudaStream_t)stream;
  if (Cs == 64) return bn == 16 ? launch_stem_wgrad<16, 64>(tx, tdy, p, grid, st) : launch_stem_wgrad<32, 64>(tx, tdy, p, grid, st);
  switch (bn) {
    case 16: return launch_stem_wgrad<16, 32>(tx, tdy, p, grid, st);
    case 32: return launch_stem_wgrad<32, 32>(tx, tdy, p, grid, st);
    default: return launch_stem_wgrad<64, 32>(tx, tdy, p, grid, st);
  }
}

}  // extern "C"
namespace {
int conv_gemm_impl(const SfbConvGeom* geom, const void* g, const SfbPrologue* pro, const void* wp, int32_t rows_pad,
                   int32_t k_pad, void* out, const SfbEpilogue* epi, int32_t impl, void* stream, const SfbBnBwdFuse* bn);
}
extern "C" {
int sfb_conv_gemm(const SfbConvGeom* geom, const void* g, const SfbPrologue* pro, const void* wp, int32_t rows_pad,
                  int32_t k_pad, void* out, const SfbEpilogue* epi, int32_t impl, void* stream) {
  return conv_gemm_impl(geom, g, pro, wp, rows_pad, k_pad, out, epi, impl, stream, nullptr);
}

int sfb_conv_dgrad_bn(const SfbConvGeom* geom, const void* g, const void* wp, int32_t rows_pad, int32_t k_pad, void* out,
                      const SfbBnBwdFuse* bn, void* stream) {
  if (!geom || !bn || !geom->transposed) return SFB_ERR_BAD_ARG;
  if (!bn->y || !aligned16(bn->y) || !bn->scale || !bn->shift || !bn->mean || !bn->invstd || !bn->dgamma || !bn->dbeta)
    return SFB_ERR_BAD_ARG;
  if (bn->C <= 0 || (bn->C & 7) || (bn->y_ld & 7) || (bn->y_coff & 7) || bn->y_coff + geom->Co > bn->y_ld) return SFB_ERR_BAD_ARG;
  if (bn->fold_ld < 0 || (bn->fold_ld & 7) || (bn->fold_ld > 0 && (geom->Co % bn->fold_ld || bn->C > bn->fold_ld)))
    return SFB_ERR_BAD_ARG;
  if (bn->fold_ld == 0 && bn->C > geom->Co) return SFB_ERR_BAD_ARG;
  return conv_gemm_impl(geom, g, nullptr, wp, rows_pad, k_pad, out, nullptr, SFB_IMPL_TC, stream, bn);
}
}  // extern "C"

namespace {
int conv_gemm_impl(const SfbConvGeom* geom, const void* g, const SfbPrologue* pro, const void* wp, int32_t rows_pad,
                   int32_t k_pad, void* out, const SfbEpilogue* epi, int32_t impl, void* stream, const SfbBnBwdFuse* bn) {
  int rc = check_geom(geom);
  if (rc) return rc;
  if (!g || !wp || !out || !aligned16(g) || !aligned16(wp) || !aligned16(out)) return SFB_ERR_BAD_ARG;
  const int taps = geom->kT * geom->kH * geom->kW;
  if ((k_pad & 63) || k_pad < taps * geom->Cg || rows_pad < geom->Co) return SFB_ERR_BAD_ARG;
  ConvTcParams p;
  memset(&p, 0, sizeof(p));
  p.g = *geom;
  p.G = (const __nv_bfloat16*)g;
  if (pro && pro->scale) {
    if (!pro->shift) return SFB_ERR_BAD_ARG;
    p.pro_scale = pro->scale; p.pro_shift = pro->shift; p.pro_relu = pro->relu;
  }
  p.out = (__nv_bfloat16*)out;
  if (epi) {
    if (epi->scale && !epi->shift) return SFB_ERR_BAD_ARG;
    p.epi_scale = epi->scale; p.epi_shift = epi->shift;
    p.res = (const __nv_bfloat16*)epi->residual; p.res_ld = epi->res_ld; p.res_coff = epi->res_coff;
    if (p.res && ((p.res_ld & 7) || (p.res_coff & 7) || !aligned16(p.res))) return SFB_ERR_BAD_ARG;
    p.epi_relu = epi->relu;
    p.stats = epi->stats;
  }
  p.M = geom->N * geom->To * geom->Ho * geom->Wo;
  p.Ktot = taps * geom->Cg;
  p.nk = k_pad / 64;
  cudaStream_t st = (cudaStream_t)stream;
  if (bn && impl != SFB_IMPL_TC) return SFB_ERR_UNSUPPORTED;
  if (impl == SFB_IMPL_SIMT) {
    ConvSimtParams q;
    q.c = p; q.wp = (const __nv_bfloat16*)wp; q.k_pad = k_pad;
    const long long total = (long long)p.M * geom->Co;
    const long long blocks = (total + 255) / 256;
    if (blocks >= (1ll << 31)) return SFB_ERR_UNSUPPORTED;
    conv_gemm_simt_kernel<<<(unsigned)blocks, 256, 0, st>>>(q);
    SFB_CHECK_LAUNCH("conv_gemm_simt_kernel");
    return SFB_OK;
  }
  if (impl != SFB_IMPL_TC && impl != SFB_IMPL_TC_GATHER) return SFB_ERR_BAD_ARG;
  if (geom->Cg > 2048) return SFB_ERR_UNSUPPORTED;
  const int bnt = pick_bn(geom->Co);
  if (rows_pad % bnt) return SFB_ERR_BAD_ARG;
  CUtensorMap tm;
  rc = get_weight_tmap(wp, rows_pad, k_pad, bnt, &tm);
  if (rc) return rc;
  const bool strided_dgrad = geom->transposed && (geom->sT != 1 || geom->sH != 1 || geom->sW != 1);
  if (impl == SFB_IMPL_TC && strided_dgrad) {
    // ---- strided dgrad: one stride-1 launch of the persistent TMA kernel per parity class; the operand `wp` must be
    //      the class-major packing of sfb_pack_weight_dgrad ----
    if (p.pro_scale || p.epi_scale || p.epi_relu || p.stats) return SFB_ERR_UNSUPPORTED;
    if ((geom->Cg % 16) || (geom->Cg > 64 && (geom->Cg % 64)) || (geom->Cg < 64 && geom->Cg != 16 && geom->Cg != 32) ||
        geom->Co > 2048)
      return SFB_ERR_UNSUPPORTED;
    const bool accumulate = p.res != nullptr;
    if (accumulate && !(p.res == p.out && p.res_ld == geom->o_ld && p.res_coff == geom->o_coff)) return SFB_ERR_UNSUPPORTED;
    const int k[3] = {geom->kT, geom->kH, geom->kW}, s[3] = {geom->sT, geom->sH, geom->sW};
    const int pd[3] = {geom->pT, geom->pH, geom->pW}, O[3] = {geom->To, geom->Ho, geom->Wo};
    DgradClassTable tb;
    int kp = 0;
    rc = build_dgrad_classes(k, s, pd, O, geom->Cg, &tb, &kp);
    if (rc) return rc;
    if (kp != k_pad) return SFB_ERR_BAD_ARG;
    bool any_empty = false;
    for (int ci = 0; ci < tb.n; ++ci) any_empty |= tb.c[ci].taps == 0;
    if (any_empty && !accumulate) {
      ViewW vo{p.out, geom->o_ld, geom->o_coff};
      zero_view_kernel<<<grid_for((long long)p.M * (geom->Co >> 3), 256), 256, 0, st>>>(vo, p.M, geom->Co);
      SFB_CHECK_LAUNCH("zero_view_kernel");
    }
    CUtensorMap ta;
    for (int ci = 0; ci < tb.n; ++ci) {
      const DgradClass& c = tb.c[ci];
      if (c.taps == 0 || c.na[0] <= 0 || c.na[1] <= 0 || c.na[2] <= 0) continue;
      ConvTmaParams q;
      memset(&q, 0, sizeof(q));
      q.N = geom->N; q.To = c.na[0]; q.Ho = c.na[1]; q.Wo = c.na[2];
      q.Co = geom->Co; q.o_ld = geom->o_ld; q.o_coff = geom->o_coff;
      q.kT = c.nj[0]; q.kH = c.nj[1]; q.kW = c.nj[2];
      q.sT = q.sH = q.sW = 1;
      q.pT = c.amin[0]; q.pH = c.amin[1]; q.pW = c.amin[2];  // transposed box origin = a0 + amin - j
      q.transposed = 1;
      q.Cg = geom->Cg;
      q.box_k = geom->Cg >= 64 ? 64 : geom->Cg;
      q.cchunks = geom->Cg >= 64 ? geom->Cg / 64 : 1;
      q.taps = c.taps; q.Ktot = c.taps * geom->Cg; q.nk = (q.Ktot + 63) / 64; q.wk0 = c.k0;
      pick_patch(q.Wo, q.Ho, q.To, 1, 1, 1, &q.Wt, &q.Ht, &q.Tt);
      q.tilesW = (q.Wo + q.Wt - 1) / q.Wt; q.tilesH = (q.Ho + q.Ht - 1) / q.Ht; q.tilesT = (q.To + q.Tt - 1) / q.Tt;
      q.n_tiles = (geom->Co + bnt - 1) / bnt;
      q.total_patches = q.N * q.tilesT * q.tilesH * q.tilesW;
      q.ocols = geom->Co >= 64 ? 64 : (geom->Co >= 32 ? 32 : (geom->Co >= 16 ? 16 : 8));
      q.out = p.out;
      q.reduce_add = accumulate ? 1 : 0;
      rc = get_act_tmap((const __nv_bfloat16*)g + geom->g_coff, geom->Cg, geom->Wi, geom->Hi, geom->Ti, geom->N, geom->g_ld,
                        q.box_k, q.Wt, q.Ht, q.Tt, 1, 1, 1, &ta);
      if (rc) return rc;
      // output: the class's sub-lattice of dX as a strided 5-D tensor
      const int o0[3] = {s[0] * c.amin[0] + c.r[0] - pd[0], s[1] * c.amin[1] + c.r[1] - pd[1], s[2] * c.amin[2] + c.r[2] - pd[2]};
      const long long ld2 = (long long)geom->o_ld * 2;
      const __nv_bfloat16* obase = (const __nv_bfloat16*)out + geom->o_coff +
                                   (((long long)o0[0] * geom->Ho + o0[1]) * geom->Wo + o0[2]) * geom->o_ld;
      const long long od[5] = {geom->Co, c.na[2], c.na[1], c.na[0], geom->N};
      const long long os[4] = {s[2] * ld2, (long long)s[1] * geom->Wo * ld2, (long long)s[0] * geom->Ho * geom->Wo * ld2,
                               (long long)geom->To * geom->Ho * geom->Wo * ld2};
      const int ob[5] = {q.ocols, q.Wt, q.Ht, q.Tt, 1};
      CUtensorMap to;
      rc = get_tmap_generic(obase, 5, od, os, ob, q.ocols * 2 >= 128 ? 128 : q.ocols * 2, &to);
      if (rc) return rc;
      CUtensorMap ty = to;
      if (bn) {
        // the class's sub-lattice of y: the same strided 5-D tensor on the y view
        if (accumulate) return SFB_ERR_UNSUPPORTED;
        const long long yld2 = (long long)bn->y_ld * 2;
        const __nv_bfloat16* ybase = (const __nv_bfloat16*)bn->y + bn->y_coff +
                                     (((long long)o0[0] * geom->Ho + o0[1]) * geom->Wo + o0[2]) * bn->y_ld;
        const long long ys[4] = {s[2] * yld2, (long long)s[1] * geom->Wo * yld2, (long long)s[0] * geom->Ho * geom->Wo * yld2,
                                 (long long)geom->To * geom->Ho * geom->Wo * yld2};
        rc = get_tmap_generic(ybase, 5, od, ys, ob, q.ocols * 2 >= 128 ? 128 : q.ocols * 2, &ty);
        if (rc) return rc;
        set_bn_fuse(q, bn, geom->Co);
      }
      switch (bnt) {
        case 32: rc = launch_conv_tma<32>(ta, tm, to, ty, q, st); break;
        case 64: rc = launch_conv_tma<64>(ta, tm, to, ty, q, st); break;
        case 128: rc = launch_conv_tma<128>(ta, tm, to, ty, q, st); break;
        default: rc = launch_conv_tma<256>(ta, tm, to, ty, q, st); break;
      }
      if (rc) return rc;
    }
    return SFB_OK;
  }
  if (impl == SFB_IMPL_TC && !p.pro_scale && tma_eligible(geom, taps)) {
    // ---- persistent TMA / tcgen05 kernel ----
    ConvTmaParams q;
    memset(&q, 0, sizeof(q));
    const bool pointwise = taps == 1 && geom->sT == 1 && geom->sH == 1 && geom->sW == 1 && geom->pT == 0 &&
                           geom->pH == 0 && geom->pW == 0;
    int gW = geom->Wi, gH = geom->Hi, gT = geom->Ti, gN = geom->N;
    q.N = geom->N; q.To = geom->To; q.Ho = geom->Ho; q.Wo = geom->Wo;
    if (pointwise) {  // a 1x1x1 stride-1 conv is a plain GEMM over the flattened pixel list: full 128-row tiles
      gW = geom->N * geom->Ti * geom->Hi * geom->Wi; gH = gT = gN = 1;
      q.N = 1; q.To = 1; q.Ho = 1; q.Wo = p.M;
    } else if (geom->kH == 1 && geom->kW == 1 && geom->sH == 1 && geom->sW == 1 && geom->pH == 0 && geom->pW == 0 &&
               geom->Hi == geom->Ho && geom->Wi == geom->Wo && merge_hw_enabled()) {
      // temporal convs ((3,1,1), the laterals): H and W carry no taps, so a frame is ONE axis of H*W pixels and a patch
      // may cut it anywhere -- a 14 x 14 map tiles as 13 patches of (16 pixels x 8 frames) per clip instead of 14 of
      // (14 x 2 x 4), which on 148 SMs is 6 rounds of tiles instead of 7 for the res4 layers (SFB_MERGE_HW=0: off)
      gW = geom->Wi * geom->Hi; gH = 1;
      q.Wo = geom->Wo * geom->Ho; q.Ho = 1;
    }
    q.Co = geom->Co; q.o_ld = geom->o_ld; q.o_coff = geom->o_coff;
    q.kT = geom->kT; q.kH = geom->kH; q.kW = geom->kW;
    q.sT = geom->sT; q.sH = geom->sH; q.sW = geom->sW;
    q.pT = geom->pT; q.pH = geom->pH; q.pW = geom->pW;
    q.transposed = geom->transposed;
    q.Cg = geom->Cg;
    q.box_k = geom->Cg >= 64 ? 64 : geom->Cg;
    q.cchunks = geom->Cg >= 64 ? (geom->Cg + 63) / 64 : 1;
    q.taps = taps; q.Ktot = p.Ktot; q.nk = p.nk;
    if (geom->Cg >= 64 && q.cchunks * taps != q.nk) return SFB_ERR_BAD_ARG;
    pick_patch(q.Wo, q.Ho, q.To, q.sW, q.sH, q.sT, &q.Wt, &q.Ht, &q.Tt);
    // halo mode (ConvTmaParams::halo): (1,3,3) / stride 1 / pad (0,1,1) on 64-channel pixels with <= 64 outputs -- the
    // conv_b of Slow res2 and of every pixel-folded Fast stage; the patch is <= 14 pixels wide so that a line of the box
    // with its border is <= 16 pixel rows, and as tall as 128 MMA rows allow
    int halo_gw = 0, halo_gh = 0;
    if (halo_enabled() && taps == 9 && geom->kT == 1 && geom->kH == 3 && geom->kW == 3 && geom->sT == 1 && geom->sH == 1 &&
        geom->sW == 1 && geom->pT == 0 && geom->pH == 1 && geom->pW == 1 && geom->Cg == 64 && bnt == 64 && p.nk == 9 &&
        ConvTmaCfg<64, true>::smem_bytes(p.nk) <= 227 * 1024) {
      const int wt = q.Wo + 2 <= 16 ? q.Wo : 14, pitch = wt + 2;
      const int ht = std::min(q.Ho, 128 / pitch);
      if ((ht + 2) * pitch <= (int)(kHaloStride / 128) && 2 * pitch + 2 + 128 <= (int)(kHaloStride / 128) && ht * wt >= 96) {
        q.halo = 1; q.pitch = pitch; q.Wt = wt; q.Ht = ht; q.Tt = 1;
        q.halo_stride = kHaloStride;
        q.halo_bytes = (uint32_t)(pitch * (ht + 2) * 128);
        halo_gw = pitch; halo_gh = ht + 2;
      }
    }
    q.tilesW = (q.Wo + q.Wt - 1) / q.Wt; q.tilesH = (q.Ho + q.Ht - 1) / q.Ht; q.tilesT = (q.To + q.Tt - 1) / q.Tt;
    q.n_tiles = (geom->Co + bnt - 1) / bnt;
    const long long tt = (long long)q.N * q.tilesT * q.tilesH * q.tilesW;
    if (tt >= (1ll << 31)) return SFB_ERR_UNSUPPORTED;
    q.total_patches = (int)tt;
    q.ocols = geom->Co >= 64 ? 64 : (geom->Co >= 32 ? 32 : (geom->Co >= 16 ? 16 : 8));
    q.out = p.out; q.epi_scale = p.epi_scale; q.epi_shift = p.epi_shift; q.res = p.res; q.res_ld = p.res_ld;
    q.res_coff = p.res_coff; q.epi_relu = p.epi_relu; q.stats = p.stats;
    // in-place accumulation (dX += dgrad): no residual read in the SM, the TMA store becomes a bf16 reduce-add in L2
    if (q.res == p.out && q.res_ld == geom->o_ld && q.res_coff == geom->o_coff && !q.epi_scale && !q.epi_relu && !q.stats) {
      q.reduce_add = 1;
      q.res = nullptr;
    }
    // CTA pairs for the 256-wide layers whose weights are streamed (not resident) and whose tiles are long enough to be
    // bound by the MMA operand traffic: the (3,1,1) / (1,3,3) convs of res3..res5 and the strided projections.  The
    // stride-1 pointwise convs (conv_c and its dgrad) and short-K tiles are bound by their epilogue and HBM; coupling
    // two CTAs per accumulator made them 10 % slower (profiles/r4b_ops_train_p0/p1.txt).
    if (pair_enabled() && bnt == 256 && geom->Cg >= 64 && ConvTmaCfg<256, true>::smem_bytes(p.nk) > 227 * 1024 &&
        q.n_tiles <= num_sms() / 2 &&
        ((p.nk >= 10 && !pointwise) || (!bn && (q.epi_scale || q.epi_relu || (q.res && !q.reduce_add)) && p.nk >= 4))) {
      // (the inference epilogue is lighter than the training one: there every streamed-weight layer gains, 15.25 -> 14.85 ms)
      q.pair = 1;
      rc = get_weight_tmap(wp, rows_pad, k_pad, 128, &tm);  // each CTA of a pair loads 128 of the tile's 256 rows
      if (rc) return rc;
    }
    CUtensorMap ta;
    const int sW = geom->transposed ? 1 : q.sW, sH = geom->transposed ? 1 : q.sH, sT = geom->transposed ? 1 : q.sT;
    if (q.halo)
      rc = get_act_tmap((const __nv_bfloat16*)g + geom->g_coff, geom->Cg, gW, gH, gT, gN, geom->g_ld, 64, halo_gw, halo_gh, 1,
                        1, 1, 1, &ta);
    else
      rc = get_act_tmap((const __nv_bfloat16*)g + geom->g_coff, geom->Cg, gW, gH, gT, gN, geom->g_ld, q.box_k, q.Wt, q.Ht,
                        q.Tt, sW, sH, sT, &ta);
    if (rc) return rc;
    CUtensorMap to;
    rc = get_act_tmap((const __nv_bfloat16*)out + geom->o_coff, geom->Co, q.Wo, q.Ho, q.To, q.N, geom->o_ld, q.ocols, q.Wt,
                      q.Ht, q.Tt, 1, 1, 1, &to);
    if (rc) return rc;
    CUtensorMap ty = to;
    if (!bn && q.res && !q.reduce_add) {
      // inference epilogue with a residual: its tiles travel by TMA, one chunk ahead of the epilogue (same boxes as the store)
      rc = get_act_tmap(q.res + q.res_coff, geom->Co, q.Wo, q.Ho, q.To, q.N, q.res_ld, q.ocols, q.Wt, q.Ht, q.Tt, 1, 1, 1, &ty);
      if (rc) return rc;
      q.res_tile = 1;
    }
    if (bn) {
      if (q.reduce_add || q.res || q.epi_scale || q.epi_relu || q.stats) return SFB_ERR_UNSUPPORTED;
      rc = get_act_tmap((const __nv_bfloat16*)bn->y + bn->y_coff, geom->Co, q.Wo, q.Ho, q.To, q.N, bn->y_ld, q.ocols, q.Wt,
                        q.Ht, q.Tt, 1, 1, 1, &ty);
      if (rc) return rc;
      set_bn_fuse(q, bn, geom->Co);
    }
    switch (bnt) {
      case 32: return launch_conv_tma<32>(ta, tm, to, ty, q, st);
      case 64: return launch_conv_tma<64>(ta, tm, to, ty, q, st);
      case 128: return launch_conv_tma<128>(ta, tm, to, ty, q, st);
      default: return launch_conv_tma<256>(ta, tm, to, ty, q, st);
    }
  }
  if (bn) return SFB_ERR_UNSUPPORTED;
  switch (bnt) {
    case 32: return launch_conv_tc<32, 4>(tm, p, st);
    case 64: return launch_conv_tc<64, 4>(tm, p, st);
    case 128: return launch_conv_tc<128, 3>(tm, p, st);
    default: return launch_conv_tc<256, 4>(tm, p, st);
  }
}
}  // namespace
extern "C" {

namespace {
struct WgradFold { int f, ldo, ldi, kW, Co, Ci; };
int conv_wgrad_impl(const SfbConvGeom* geom, const void* x, const SfbPrologue* pro, const void* dy, int32_t dy_ld,
                    int32_t dy_coff, float* dw, int32_t Ci_real, int32_t impl, void* stream, const WgradFold* fold);
}  // namespace

int sfb_conv_wgrad(const SfbConvGeom* geom, const void* x, const SfbPrologue* pro, const void* dy, int32_t dy_ld,
                   int32_t dy_coff, float* dw, int32_t Ci_real, int32_t impl, void* stream) {
  return conv_wgrad_impl(geom, x, pro, dy, dy_ld, dy_coff, dw, Ci_real, impl, stream, nullptr);
}

int sfb_conv_wgrad_folded(const SfbConvGeom* geom, const void* x, const void* dy, int32_t dy_ld, int32_t dy_coff, float* dw,
                          int32_t Co, int32_t Ci, int32_t kW, int32_t f, int32_t ldo, int32_t ldi, void* stream) {
  if (!geom || !fold_args_ok(Co, Ci, geom->kT, geom->kH, kW, f, ldo, ldi)) return SFB_ERR_BAD_ARG;
  if (geom->Co != f * ldo || geom->Cg != f * ldi || geom->kW != (kW == 1 ? 1 : 3)) return SFB_ERR_BAD_ARG;
  if (geom->sT != 1 || geom->sH != 1 || geom->sW != 1 || (geom->Cg % 16)) return SFB_ERR_UNSUPPORTED;
  const WgradFold fold{f, ldo, ldi, kW, Co, Ci};
  return conv_wgrad_impl(geom, x, nullptr, dy, dy_ld, dy_coff, dw, geom->Cg, SFB_IMPL_TC, stream, &fold);
}

namespace {
int conv_wgrad_impl(const SfbConvGeom* geom, const void* x, const SfbPrologue* pro, const void* dy, int32_t dy_ld,
                    int32_t dy_coff, float* dw, int32_t Ci_real, int32_t impl, void* stream, const WgradFold* fold) {
  int rc = check_geom(geom);
  if (rc) return rc;
  if (geom->transposed) return SFB_ERR_BAD_ARG;
  if (!x || !dy || !dw || !aligned16(x) || !aligned16(dy)) return SFB_ERR_BAD_ARG;
  if ((dy_ld & 7) || (dy_coff & 7) || dy_coff + geom->Co > dy_ld || Ci_real <= 0 || Ci_real > geom->Cg) return SFB_ERR_BAD_ARG;
  WgradTcParams p;
  memset(&p, 0, sizeof(p));
  p.g = *geom;
  p.X = (const __nv_bfloat16*)x;
  if (pro && pro->scale) {
    if (!pro->shift) return SFB_ERR_BAD_ARG;
    p.pro_scale = pro->scale; p.pro_shift = pro->shift; p.pro_relu = pro->relu;
  }
  p.dY = (const __nv_bfloat16*)dy; p.dy_ld = dy_ld; p.dy_coff = dy_coff;
  p.dW = dw; p.Ci_real = Ci_real;
  p.taps = geom->kT * geom->kH * geom->kW;
  p.M = geom->N * geom->To * geom->Ho * geom->Wo;
  p.Ktot = p.taps * geom->Cg;
  p.chunks_total = (p.M + 63) / 64;
  cudaStream_t st = (cudaStream_t)stream;
  if (impl == SFB_IMPL_SIMT) {
    WgradSimtParams q;
    q.w = p;
    const long long total = (long long)geom->Co * Ci_real * p.taps;
    int splits = (int)((148ll * 8 * 256 + total - 1) / total);
    if (splits < 1) splits = 1;
    if (splits > 1024) splits = 1024;
    q.m_per_split = (p.M + splits - 1) / splits;
    splits = (p.M + q.m_per_split - 1) / q.m_per_split;
    dim3 grid((unsigned)((total + 255) / 256), splits);
    conv_wgrad_simt_kernel<<<grid, 256, 0, st>>>(q);
    SFB_CHECK_LAUNCH("conv_wgrad_simt_kernel");
    return SFB_OK;
  }
  if (impl != SFB_IMPL_TC && impl != SFB_IMPL_TC_GATHER) return SFB_ERR_BAD_ARG;
  if (geom->Cg > 2048) return SFB_ERR_UNSUPPORTED;
  if (fold && !(impl == SFB_IMPL_TC && !p.pro_scale && (geom->Cg % 16) == 0)) return SFB_ERR_UNSUPPORTED;
  if (impl == SFB_IMPL_TC && !p.pro_scale && (geom->Cg % 16) == 0) {
    // ---- TMA-fed tcgen05 kernel ----
    WgradTmaParams q;
    memset(&q, 0, sizeof(q));
    q.N = geom->N; q.To = geom->To; q.Ho = geom->Ho; q.Wo = geom->Wo;
    q.Co = geom->Co; q.Cg = geom->Cg; q.Ci_real = Ci_real; q.taps = p.taps;
    q.kT = geom->kT; q.kH = geom->kH; q.kW = geom->kW; q.sT = geom->sT; q.sH = geom->sH; q.sW = geom->sW;
    q.pT = geom->pT; q.pH = geom->pH; q.pW = geom->pW;
    q.box_k = geom->Cg >= 64 ? 64 : geom->Cg;
    if (q.box_k != 64 && q.box_k != 32 && q.box_k != 16) q.box_k = 16;  // Cg = 48: three 16-channel... (not used)
    q.cchunks = geom->Cg >= 64 ? (geom->Cg + 63) / 64 : geom->Cg / q.box_k;
    q.n_units = p.taps * q.cchunks;
    q.upc = 256 / q.box_k;
    if (q.upc > q.n_units) q.upc = q.n_units;
    q.upc = (q.n_units + (q.n_units + q.upc - 1) / q.upc - 1) / ((q.n_units + q.upc - 1) / q.upc);  // same CTA count, even split (9 units: 3+3+3, not 4+4+1)
    int gW = geom->Wi, gH = geom->Hi, gT = geom->Ti, gN = geom->N;
    const bool pointwise = p.taps == 1 && geom->sT == 1 && geom->sH == 1 && geom->sW == 1 && geom->pT == 0 &&
                           geom->pH == 0 && geom->pW == 0;
    if (pointwise) {
      gW = p.M; gH = gT = gN = 1;
      q.N = 1; q.To = 1; q.Ho = 1; q.Wo = p.M;
    } else if (geom->kH == 1 && geom->kW == 1 && geom->sH == 1 && geom->sW == 1 && geom->pH == 0 && geom->pW == 0 &&
               geom->Hi == geom->Ho && geom->Wi == geom->Wo && merge_hw_enabled()) {
      // temporal convs: one H*W axis, as in conv_gemm_impl (26 patches of 16 pixels x 4 frames per 14 x 14 x 8 clip
      // instead of 32 of 14 x 4 x 1)
      gW = geom->Wi * geom->Hi; gH = 1;
      q.Wo = geom->Wo * geom->Ho; q.Ho = 1;
    }
    pick_patch(q.Wo, q.Ho, q.To, q.sW, q.sH, q.sT, &q.Wt, &q.Ht, &q.Tt, kWtRows);
    q.tilesW = (q.Wo + q.Wt - 1) / q.Wt; q.tilesH = (q.Ho + q.Ht - 1) / q.Ht; q.tilesT = (q.To + q.Tt - 1) / q.Tt;
    const long long tp = (long long)q.N * q.tilesT * q.tilesH * q.tilesW;
    if (tp >= (1ll << 31)) return SFB_ERR_UNSUPPORTED;
    q.total_patches = (int)tp;
    q.dW = dw;
    if (fold) {
      q.fold_f = fold->f; q.fold_ldo = fold->ldo; q.fold_ldi = fold->ldi; q.fold_kW = fold->kW;
      q.fold_Co = fold->Co; q.fold_Ci = fold->Ci;
    }
    const int gx = (q.n_units + q.upc - 1) / q.upc;
    // second accumulator only where the X boxes dominate the stage traffic (many tap/channel units) and enough CTAs remain
    q.mh = (geom->Co > 128 && q.n_units >= 2 * q.upc) ? 2 : 1;
    const int gy = (geom->Co + 128 * q.mh - 1) / (128 * q.mh);
    // Split of the pixel reduction over blockIdx.z: one CTA per SM at a time, so the launch runs in
    // rounds = ceil(CTAs / SMs) and every CTA pays a fixed cost (operand-stage zeroing, TMEM allocation, pipeline fill,
    // the atomic epilogue ~ `fixed` patches of MMA time).  Pick the split with the least rounds * (patches + fixed).
    const int fixed = q.mh == 2 ? 32 : 48;  // measured insensitive between 16 and 96
    const int gxy = gx * gy;
    long long splits = 1;
    double best_cost = 1e30;
    const int max_splits = (int)std::min<long long>(q.total_patches, (4ll * num_sms() + gxy - 1) / gxy);
    for (int sp = 1; sp <= max_splits; ++sp) {
      const long long rounds = ((long long)gxy * sp + num_sms() - 1) / num_sms();
      const double cost = (double)rounds * ((double)((q.total_patches + sp - 1) / sp) + fixed);
      if (cost < best_cost) { best_cost = cost; splits = sp; }
    }
    q.patches_per_split = (int)((q.total_patches + splits - 1) / splits);
    splits = (q.total_patches + q.patches_per_split - 1) / q.patches_per_split;
    CUtensorMap tx, tdy;
    // X boxes: when Cg < 64 a "64-channel chunk" is a box_k-channel chunk at channel offset chunk*box_k
    rc = get_act_tmap((const __nv_bfloat16*)x + geom->g_coff, geom->Cg, gW, gH, gT, gN, geom->g_ld, q.box_k, q.Wt, q.Ht,
                      q.Tt, q.sW, q.sH, q.sT, &tx);
    if (rc) return rc;
    rc = get_act_tmap((const __nv_bfloat16*)dy + dy_coff, geom->Co, q.Wo, q.Ho, q.To, q.N, dy_ld, 64, q.Wt, q.Ht, q.Tt, 1, 1,
                      1, &tdy);
    if (rc) return rc;
    static bool attr_done2 = false;
    if (!attr_done2) {
      cudaError_t e = cudaFuncSetAttribute(wgrad_tma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWtSmem);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(wgrad_tma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWtSmem);
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(wgrad_tma)");
      attr_done2 = true;
    }
    dim3 grid(gx, gy, (unsigned)splits);
    if (q.mh == 2) wgrad_tma_kernel<2><<<grid, 192, kWtSmem, st>>>(tx, tdy, q);
    else wgrad_tma_kernel<1><<<grid, 192, kWtSmem, st>>>(tx, tdy, q);
    SFB_CHECK_LAUNCH("wgrad_tma_kernel");
    return SFB_OK;
  }
  const int kpad = (p.Ktot + 63) / 64 * 64;
  const int gx = (kpad + 255) / 256;
  const int gy = (geom->Co + 127) / 128;
  int splits = (148 * 2 + gx * gy - 1) / (gx * gy);
  if (splits < 1) splits = 1;
  if (splits > p.chunks_total) splits = p.chunks_total;
  p.chunks_per_split = (p.chunks_total + splits - 1) / splits;
  splits = (p.chunks_total + p.chunks_per_split - 1) / p.chunks_per_split;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(conv_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)wgrad_smem_bytes(2048, true));
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_wgrad_tc)");
    attr_done = true;
  }
  dim3 grid(gx, gy, splits);
  conv_wgrad_tc_kernel<<<grid, kProducerThreads + 32, wgrad_smem_bytes(geom->Cg, p.pro_scale != nullptr), st>>>(p);
  SFB_CHECK_LAUNCH("conv_wgrad_tc_kernel");
  return SFB_OK;
}
}  // namespace

int sfb_bn_finalize(const float* stats, int32_t C, int64_t count, const float* gamma, const float* beta, float eps,
                    float momentum, float* running_mean, float* running_var, float* scale, float* shift,
                    float* save_mean, float* save_invstd, void* stream) {
  if (!stats || !gamma || !beta || !scale || !shift || !save_mean || !save_invstd || C <= 0 || count <= 0)
    return SFB_ERR_BAD_ARG;
  if ((running_mean == nullptr) != (running_var == nullptr)) return SFB_ERR_BAD_ARG;
  const double unbias = count > 1 ? (double)count / (double)(count - 1) : 1.0;
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(stats, C, 1.0 / (double)count, unbias, gamma,
                                                                        beta, eps, momentum, running_mean, running_var,
                                                                        scale, shift, save_mean, save_invstd);
  SFB_CHECK_LAUNCH("bn_finalize_kernel");
  return SFB_OK;
}

int sfb_bn_fold(const float* gamma, const float* beta, const float* running_mean, const float* running_var, float eps,
                int32_t C, float* scale, float* shift, void* stream) {
  if (!gamma || !beta || !running_mean || !running_var || !scale || !shift || C <= 0) return SFB_ERR_BAD_ARG;
  bn_fold_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(gamma, beta, running_mean, running_var, eps, C,
                                                                    scale, shift);
  SFB_CHECK_LAUNCH("bn_fold_kernel");
  return SFB_OK;
}

static bool view_ok(const void* p, int ld, int coff, int C) {
  return p && aligned16(p) && !(ld & 7) && !(coff & 7) && coff + C <= ld;
}

int sfb_bn_add_relu(const void* y, int32_t y_ld, int32_t y_coff, const float* scale, const float* shift, const void* r,
                    int32_t r_ld, int32_t r_coff, const float* r_scale, const float* r_shift, void* out, int32_t o_ld,
                    int32_t o_coff, uint8_t* mask_bits, int64_t M, int32_t C, int32_t relu, void* stream) {
  if (M <= 0 || C <= 0 || (C & 7) || C > 2048 || !scale || !shift) return SFB_ERR_BAD_ARG;
  if (!view_ok(y, y_ld, y_coff, C) || !view_ok(out, o_ld, o_coff, C)) return SFB_ERR_BAD_ARG;
  if (r && !view_ok(r, r_ld, r_coff, C)) return SFB_ERR_BAD_ARG;
  if (r_scale && (!r || !r_shift)) return SFB_ERR_BAD_ARG;
  View vy{(const __nv_bfloat16*)y, y_ld, y_coff};
  View vr{(const __nv_bfloat16*)r, r_ld, r_coff};
  ViewW vo{(__nv_bfloat16*)out, o_ld, o_coff};
  const int RL = kTailThreads / (C >> 3);
  const int unroll = r ? 4 : 8;
  const int grid = grid_for((M + (long long)RL * unroll - 1) / ((long long)RL * unroll), 1, 148 * 8);
  BnRawDev none;
  memset(&none, 0, sizeof(none));
  if (r)
    bn_add_relu_kernel<true, false><<<grid, kTailThreads, 0, (cudaStream_t)stream>>>(vy, scale, shift, vr, r_scale, r_shift,
                                                                                     vo, mask_bits, M, C, relu, none, none);
  else
    bn_add_relu_kernel<false, false><<<grid, kTailThreads, 0, (cudaStream_t)stream>>>(vy, scale, shift, vr, r_scale, r_shift,
                                                                                      vo, mask_bits, M, C, relu, none, none);
  SFB_CHECK_LAUNCH("bn_add_relu_kernel");
  return SFB_OK;
}

static int bn_raw_dev(const SfbBnRaw* q, int C, BnRawDev* d) {
  memset(d, 0, sizeof(*d));
  if (!q) return SFB_OK;
  if (!q->stats || q->count <= 0 || !q->gamma || !q->beta || !q->scale || !q->shift || !q->save_mean || !q->save_invstd)
    return SFB_ERR_BAD_ARG;
  if ((q->running_mean == nullptr) != (q->running_var == nullptr)) return SFB_ERR_BAD_ARG;
  if (q->fold_f > 1 && (q->fold_f > 8 || q->fold_ld < C)) return SFB_ERR_BAD_ARG;
  d->stats = q->stats; d->gamma = q->gamma; d->beta = q->beta;
  d->running_mean = q->running_mean; d->running_var = q->running_var;
  d->scale = q->scale; d->shift = q->shift; d->save_mean = q->save_mean; d->save_invstd = q->save_invstd;
  d->inv_count = 1.0 / (double)q->count;
  d->unbias = q->count > 1 ? (double)q->count / (double)(q->count - 1) : 1.0;
  d->eps = q->eps; d->momentum = q->momentum;
  d->fold_f = q->fold_f; d->fold_ld = q->fold_ld;
  return SFB_OK;
}

int sfb_bn_add_relu_raw(const void* y, int32_t y_ld, int32_t y_coff, const SfbBnRaw* y_bn, const void* r, int32_t r_ld,
                        int32_t r_coff, const SfbBnRaw* r_bn, void* out, int32_t o_ld, int32_t o_coff, uint8_t* mask_bits,
                        int64_t M, int32_t C, int32_t relu, void* stream) {
  if (M <= 0 || C <= 0 || (C & 7) || C > 2048 || !y_bn) return SFB_ERR_BAD_ARG;
  if (!view_ok(y, y_ld, y_coff, C) || !view_ok(out, o_ld, o_coff, C)) return SFB_ERR_BAD_ARG;
  if (r && !view_ok(r, r_ld, r_coff, C)) return SFB_ERR_BAD_ARG;
  if (r_bn && !r) return SFB_ERR_BAD_ARG;
  BnRawDev dy, dr;
  int rc = bn_raw_dev(y_bn, C, &dy);
  if (rc) return rc;
  rc = bn_raw_dev(r_bn, C, &dr);
  if (rc) return rc;
  View vy{(const __nv_bfloat16*)y, y_ld, y_coff};
  View vr{(const __nv_bfloat16*)r, r_ld, r_coff};
  ViewW vo{(__nv_bfloat16*)out, o_ld, o_coff};
  const int RL = kTailThreads / (C >> 3);
  const int unroll = r ? 4 : 8;
  const int grid = grid_for((M + (long long)RL * unroll - 1) / ((long long)RL * unroll), 1, 148 * 8);
  if (r)
    bn_add_relu_kernel<true, true><<<grid, kTailThreads, 0, (cudaStream_t)stream>>>(vy, nullptr, nullptr, vr, nullptr, nullptr,
                                                                                    vo, mask_bits, M, C, relu, dy, dr);
  else
    bn_add_relu_kernel<false, true><<<grid, kTailThreads, 0, (cudaStream_t)stream>>>(vy, nullptr, nullptr, vr, nullptr, nullptr,
                                                                                     vo, mask_bits, M, C, relu, dy, dr);
  SFB_CHECK_LAUNCH("bn_add_relu_kernel");
  return SFB_OK;
}

int sfb_bn_bwd_reduce(const void* dz, int32_t dz_ld, int32_t dz_coff, const void* y, int32_t y_ld, int32_t y_coff,
                      const void* mask_src, int32_t m_ld, int32_t m_coff, const uint8_t* mask_bits, const float* scale,
                      const float* shift, const float* mean, const float* invstd, int32_t relu, int64_t M, int32_t C,
                      float* dgamma, float* dbeta, void* stream) {
  if (M <= 0 || C <= 0 || (C & 7) || C > 2048 || !mean || !invstd || !dgamma || !dbeta) return SFB_ERR_BAD_ARG;
  if (!view_ok(dz, dz_ld, dz_coff, C) || !view_ok(y, y_ld, y_coff, C)) return SFB_ERR_BAD_ARG;
  if (mask_src && !view_ok(mask_src, m_ld, m_coff, C)) return SFB_ERR_BAD_ARG;
  if (relu && !mask_src && !mask_bits && (!scale || !shift)) return SFB_ERR_BAD_ARG;
  View vdz{(const __nv_bfloat16*)dz, dz_ld, dz_coff};
  View vy{(const __nv_bfloat16*)y, y_ld, y_coff};
  View vm{(const __nv_bfloat16*)mask_src, m_ld, m_coff};
  const int RL = kRedThreads / (C >> 3);
  const int grid = grid_for((M + (long long)RL * kRedUnroll - 1) / ((long long)RL * kRedUnroll), 1, 148 * 8);
  const int mode = !relu ? 3 : (mask_bits ? 2 : (mask_src ? 1 : 0));
#define SFB_REDUCE(MD) bn_bwd_reduce_kernel<MD><<<grid, kRedThreads, 0, (cudaStream_t)stream>>>(       vdz, vy, vm, mask_bits, scale, shift, mean, invstd, M, C, dgamma, dbeta)
  if (mode == 0) SFB_REDUCE(0);
  else if (mode == 1) SFB_REDUCE(1);
  else if (mode == 2) SFB_REDUCE(2);
  else SFB_REDUCE(3);
#undef SFB_REDUCE
  SFB_CHECK_LAUNCH("bn_bwd_reduce_kernel");
  return SFB_OK;
}

int sfb_bn_bwd_apply(const void* dz, int32_t dz_ld, int32_t dz_coff, const void* y, int32_t y_ld, int32_t y_coff,
                     const void* mask_src, int32_t m_ld, int32_t m_coff, const uint8_t* mask_bits, const float* scale,
                     const float* shift, const float* mean, const float* invstd, const float* gamma, const float* dgamma,
                     const float* dbeta, int32_t relu, int64_t M, int32_t C, void* dy, int32_t dy_ld, int32_t dy_coff,
                     void* g_out, int32_t g_ld, int32_t g_coff, int32_t g_accumulate, void* stream) {
  if (M <= 0 || C <= 0 || (C & 7) || C > 2048 || !mean || !invstd || !gamma || !dgamma || !dbeta) return SFB_ERR_BAD_ARG;
  if (!view_ok(dz, dz_ld, dz_coff, C) || !view_ok(y, y_ld, y_coff, C) || !view_ok(dy, dy_ld, dy_coff, C))
    return SFB_ERR_BAD_ARG;
  if (mask_src && !view_ok(mask_src, m_ld, m_coff, C)) return SFB_ERR_BAD_ARG;
  if (g_out && !view_ok(g_out, g_ld, g_coff, C)) return SFB_ERR_BAD_ARG;
  if (relu && !mask_src && !mask_bits && (!scale || !shift)) return SFB_ERR_BAD_ARG;
  View vdz{(const __nv_bfloat16*)dz, dz_ld, dz_coff};
  View vy{(const __nv_bfloat16*)y, y_ld, y_coff};
  View vm{(const __nv_bfloat16*)mask_src, m_ld, m_coff};
  ViewW vdy{(__nv_bfloat16*)dy, dy_ld, dy_coff};
  ViewW vg{(__nv_bfloat16*)g_out, g_ld, g_coff};
  const int RLa = kRedThreads / (C >> 3);
  const int grid_a = grid_for((M + (long long)RLa * kBnUnroll - 1) / ((long long)RLa * kBnUnroll), 1, 148 * 8);
  const int mode = !relu ? 0 : (mask_bits ? 2 : (mask_src ? 1 : 0));
  const bool hg = g_out != nullptr;
#define SFB_APPLY(MK, GO) bn_bwd_apply_kernel<MK, GO><<<grid_a, kRedThreads, 0, (cudaStream_t)stream>>>(       vdz, vy, vm, mask_bits, scale, shift, mean, invstd, gamma, dgamma, dbeta, relu, M, C, 1.0f / (float)M, vdy, vg, g_accumulate)
  if (mode == 2 && hg) SFB_APPLY(2, true);
  else if (mode == 2) SFB_APPLY(2, false);
  else if (mode == 1 && hg) SFB_APPLY(1, true);
  else if (mode == 1) SFB_APPLY(1, false);
  else if (hg) SFB_APPLY(0, true);
  else SFB_APPLY(0, false);
#undef SFB_APPLY
  SFB_CHECK_LAUNCH("bn_bwd_apply_kernel");
  return SFB_OK;
}

int sfb_bn_bwd_tail2(const void* dz, int32_t dz_ld, int32_t dz_coff, const uint8_t* mask_bits, const SfbBnTailSide* c,
                     const SfbBnTailSide* s, int64_t M, int32_t C, void* stream) {
  if (M <= 0 || C <= 0 || (C & 7) || C > 2048 || !mask_bits || !c || !s) return SFB_ERR_BAD_ARG;
  if (!view_ok(dz, dz_ld, dz_coff, C)) return SFB_ERR_BAD_ARG;
  const SfbBnTailSide* sides[2] = {c, s};
  for (const SfbBnTailSide* q : sides) {
    if (!q->mean || !q->invstd || !q->gamma || !q->dgamma || !q->dbeta) return SFB_ERR_BAD_ARG;
    if (!view_ok(q->y, q->y_ld, q->y_coff, C) || !view_ok(q->dy, q->dy_ld, q->dy_coff, C)) return SFB_ERR_BAD_ARG;
  }
  View vdz{(const __nv_bfloat16*)dz, dz_ld, dz_coff};
  BnTail2 t;
  t.yc = View{(const __nv_bfloat16*)c->y, c->y_ld, c->y_coff};
  t.ys = View{(const __nv_bfloat16*)s->y, s->y_ld, s->y_coff};
  t.mean_c = c->mean; t.invstd_c = c->invstd; t.gamma_c = c->gamma; t.dgamma_c = c->dgamma; t.dbeta_c = c->dbeta;
  t.mean_s = s->mean; t.invstd_s = s->invstd; t.gamma_s = s->gamma; t.dgamma_s = s->dgamma; t.dbeta_s = s->dbeta;
  ViewW vc{(__nv_bfloat16*)c->dy, c->dy_ld, c->dy_coff};
  ViewW vs{(__nv_bfloat16*)s->dy, s->dy_ld, s->dy_coff};
  const int RL = kRedThreads / (C >> 3);
  const int grid_r = grid_for((M + (long long)RL * kTail2Unroll - 1) / ((long long)RL * kTail2Unroll), 1, 148 * 4);
  bn_bwd_tail2_reduce_kernel<<<grid_r, kRedThreads, 0, (cudaStream_t)stream>>>(vdz, mask_bits, t, M, C);
  SFB_CHECK_LAUNCH("bn_bwd_tail2_reduce_kernel");
  const int grid_a = grid_for((M + (long long)RL * 2 - 1) / ((long long)RL * 2), 1, 148 * 8);
  bn_bwd_tail2_apply_kernel<<<grid_a, kRedThreads, 0, (cudaStream_t)stream>>>(vdz, mask_bits, t, M, C, 1.0f / (float)M, vc, vs);
  SFB_CHECK_LAUNCH("bn_bwd_tail2_apply_kernel");
  return SFB_OK;
}

int sfb_maxpool_fwd(const void* y, int32_t y_ld, int32_t y_coff, const float* scale, const float* shift, int32_t N,
                    int32_t T, int32_t H, int32_t W, int32_t C, void* out, int32_t o_ld, int32_t o_coff, uint8_t* idx,
                    void* stream) {
  if (N <= 0 || T <= 0 || H <= 0 || W <= 0 || C <= 0 || (C & 7)) return SFB_ERR_BAD_ARG;
  if (!view_ok(y, y_ld, y_coff, C) || !view_ok(out, o_ld, o_coff, C)) return SFB_ERR_BAD_ARG;
  if ((scale == nullptr) != (shift == nullptr)) return SFB_ERR_BAD_ARG;
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  View vy{(const __nv_bfloat16*)y, y_ld, y_coff};
  ViewW vo{(__nv_bfloat16*)out, o_ld, o_coff};
  if ((long long)N * T > 65535 || (long long)H * W * (C >> 3) >= (1ll << 31)) return SFB_ERR_UNSUPPORTED;
  const long long per_frame = (long long)Ho * Wo * (C >> 3);
  dim3 grid((unsigned)grid_for(per_frame, 256, 64), (unsigned)(N * T));
  maxpool_fwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(vy, scale, shift, N, T, H, W, C, vo, idx);
  SFB_CHECK_LAUNCH("maxpool_fwd_kernel");
  return SFB_OK;
}

int sfb_maxpool_bwd(const void* dout, int32_t d_ld, int32_t d_coff, const uint8_t* idx, int32_t N, int32_t T, int32_t H,
                    int32_t W, int32_t C, void* dz, int32_t dz_ld, int32_t dz_coff, void* stream) {
  if (N <= 0 || T <= 0 || H <= 0 || W <= 0 || C <= 0 || (C & 7) || !idx) return SFB_ERR_BAD_ARG;
  if (!view_ok(dout, d_ld, d_coff, C) || !view_ok(dz, dz_ld, dz_coff, C)) return SFB_ERR_BAD_ARG;
  View vd{(const __nv_bfloat16*)dout, d_ld, d_coff};
  ViewW vz{(__nv_bfloat16*)dz, dz_ld, dz_coff};
  if ((long long)N * T > 65535 || (long long)H * W * (C >> 3) >= (1ll << 31)) return SFB_ERR_UNSUPPORTED;
  const int Hb = (H - 1) / 2 + 1, Wb = (W - 1) / 2 + 1;  // 2x2 blocks of input pixels = pooling windows
  const long long per_frame = (long long)Hb * Wb * (C >> 3);
  dim3 grid((unsigned)grid_for(per_frame, 256, 64), (unsigned)(N * T));
  maxpool_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(vd, idx, N, T, H, W, C, vz);
  SFB_CHECK_LAUNCH("maxpool_bwd_kernel");
  return SFB_OK;
}

int sfb_head_pool_fwd(const void* x, int32_t x_ld, int32_t x_coff, int32_t N, int32_t T, int32_t H, int32_t W, int32_t C,
                      int32_t kT, int32_t kH, int32_t kW, float dropout_p, uint64_t seed, const uint64_t* seed_dev,
                      float* feat, int32_t f_ld, int32_t f_off, void* stream) {
  if (N <= 0 || C <= 0 || (C & 7) || kT <= 0 || kH <= 0 || kW <= 0 || kT > T || kH > H || kW > W || !feat) return SFB_ERR_BAD_ARG;
  if (!view_ok(x, x_ld, x_coff, C) || f_off + C > f_ld || dropout_p < 0.f || dropout_p >= 1.f) return SFB_ERR_BAD_ARG;
  View vx{(const __nv_bfloat16*)x, x_ld, x_coff};
  const int CV = C >> 3;
  int cvb = 1;
  while (cvb < 32 && cvb < CV) cvb <<= 1;  // channel vectors per block (power of two <= 32); the other threads split positions
  // dropout acts per pooled position: one block per window, accumulated into feat; without it the windows collapse
  const long long npos = dropout_p > 0.f ? (long long)(T - kT + 1) * (H - kH + 1) * (W - kW + 1) : 1;
  if ((long long)N * npos > 0x7fffffffll) return SFB_ERR_UNSUPPORTED;
  if (npos > 1) {
    cudaError_t e = cudaMemset2DAsync(feat + f_off, (size_t)f_ld * 4, 0, (size_t)C * 4, (size_t)N, (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemset2DAsync(head_pool_fwd)");
  }
  dim3 grid((unsigned)(N * npos), (unsigned)((CV + cvb - 1) / cvb));
  head_pool_fwd_kernel<<<grid, kHeadThreads, 0, (cudaStream_t)stream>>>(vx, N, T, H, W, C, kT, kH, kW, dropout_p, seed,
                                                                        seed_dev, feat, f_ld, f_off, cvb, (int)npos);
  SFB_CHECK_LAUNCH("head_pool_fwd_kernel");
  return SFB_OK;
}

int sfb_head_pool_bwd(const float* dfeat, int32_t f_ld, int32_t f_off, int32_t N, int32_t T, int32_t H, int32_t W,
                      int32_t C, int32_t kT, int32_t kH, int32_t kW, float dropout_p, uint64_t seed,
                      const uint64_t* seed_dev, void* dx, int32_t dx_ld, int32_t dx_coff, void* stream) {
  if (N <= 0 || C <= 0 || (C & 7) || kT <= 0 || kH <= 0 || kW <= 0 || kT > T || kH > H || kW > W || !dfeat) return SFB_ERR_BAD_ARG;
  if (!view_ok(dx, dx_ld, dx_coff, C) || f_off + C > f_ld || dropout_p < 0.f || dropout_p >= 1.f) return SFB_ERR_BAD_ARG;
  ViewW vx{(__nv_bfloat16*)dx, dx_ld, dx_coff};
  const long long total = (long long)N * T * H * W * (C >> 3);
  head_pool_bwd_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(dfeat, f_ld, f_off, N, T, H, W, C, kT, kH,
                                                                               kW, dropout_p, seed, seed_dev, vx);
  SFB_CHECK_LAUNCH("head_pool_bwd_kernel");
  return SFB_OK;
}

int sfb_fc_fwd(const float* feat, const float* w, const float* b, int32_t N, int32_t K, int32_t C, float* logits,
               void* stream) {
  if (!feat || !w || !logits || N <= 0 || K <= 0 || C <= 0) return SFB_ERR_BAD_ARG;
  const long long threads = (long long)N * C * 32;
  fc_fwd_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(feat, w, b, N, K, C, logits);
  SFB_CHECK_LAUNCH("fc_fwd_kernel");
  return SFB_OK;
}

int sfb_fc_bwd(const float* dlogits, const float* feat, const float* w, int32_t N, int32_t K, int32_t C, float* dfeat,
               float* dw, float* db, void* stream) {
  if (!dlogits || !feat || !w || !dfeat || !dw || N <= 0 || K <= 0 || C <= 0) return SFB_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  fc_bwd_dfeat_kernel<<<(N * K + 255) / 256, 256, 0, st>>>(dlogits, w, N, K, C, dfeat);
  SFB_CHECK_LAUNCH("fc_bwd_dfeat_kernel");
  fc_bwd_dw_kernel<<<(C * K + 255) / 256, 256, 0, st>>>(dlogits, feat, N, K, C, dw, db);
  SFB_CHECK_LAUNCH("fc_bwd_dw_kernel");
  return SFB_OK;
}

int sfb_video_scores(const float* scores, int32_t n_clips, int32_t C, const int32_t* seg, int32_t n_videos,
                     int32_t softmax, float* clip_probs, float* video_mean, int32_t* top1, void* stream) {
  if (!scores || !seg || n_clips <= 0 || C <= 0 || n_videos <= 0 || (!video_mean && !top1 && !clip_probs))
    return SFB_ERR_BAD_ARG;
  if (C > kScoreThreads * kScoreMaxPer) return SFB_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (softmax)
    video_scores_kernel<true><<<n_videos, kScoreThreads, 0, st>>>(scores, C, seg, clip_probs, video_mean, top1);
  else
    video_scores_kernel<false><<<n_videos, kScoreThreads, 0, st>>>(scores, C, seg, clip_probs, video_mean, top1);
  SFB_CHECK_LAUNCH("video_scores_kernel");
  return SFB_OK;
}

int sfb_sparse_fuse(const float* x, const float* w, const float* b, int32_t N, int32_t P, int32_t C, float* y,
                    void* stream) {
  if (!x || !w || !b || !y || N <= 0 || P <= 0 || C <= 0) return SFB_ERR_BAD_ARG;
  const long long total = (long long)N * C;
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)num_sms() * 8);
  sparse_fuse_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, w, b, N, P, C, y);
  SFB_CHECK_LAUNCH("sparse_fuse_kernel");
  return SFB_OK;
}

int sfb_sparse_fuse_bwd(const float* dy, const float* x, const float* w, int32_t N, int32_t P, int32_t C, float* dw,
                        float* db, float* dx, void* stream) {
  if (!dy || !x || !w || !dw || !db || N <= 0 || P <= 0 || C <= 0) return SFB_ERR_BAD_ARG;
  sparse_fuse_bwd_kernel<<<(C + 7) / 8, 256, 0, (cudaStream_t)stream>>>(dy, x, w, N, P, C, dw, db, dx);
  SFB_CHECK_LAUNCH("sparse_fuse_bwd_kernel");
  return SFB_OK;
}

}  // extern "C"

// ---- BatchNorm(+ReLU) backward without any allocation (see sfb_bn_bwd_noalloc in include/sfb200.h) -----------------
// The per-channel sums the reduce and apply phases share live in this file-scope device array: [0, kNoAllocMaxC) sum g,
// [kNoAllocMaxC, 2*kNoAllocMaxC) sum g*xhat.  16 KB, held by every GPU that loads the library until the process exits.
// In host code the name is only a handle for cudaGetSymbolAddress; the kernels get the address that call returns.
namespace {
constexpr int kNoAllocMaxC = SFB_NOALLOC_MAX_C;
constexpr int kNoAllocMaxDevices = 64;
}  // namespace
static __device__ __align__(256) float g_bn_bwd_sums[2 * kNoAllocMaxC];
namespace {
float* g_bn_bwd_sums_addr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mu;

// device address of g_bn_bwd_sums on the CURRENT device, looked up once per device
int noalloc_sums(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "sfb_bn_bwd_noalloc: cudaGetDevice");
  if (dev < 0 || dev >= kNoAllocMaxDevices) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "sfb_bn_bwd_noalloc: device ordinal %d outside [0, %d)", dev, kNoAllocMaxDevices);
    return SFB_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_bn_bwd_sums_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_bn_bwd_sums);
    if (e != cudaSuccess || !p) return cuda_fail(e, "sfb_bn_bwd_noalloc: cudaGetSymbolAddress(g_bn_bwd_sums)");
    g_bn_bwd_sums_addr[dev] = static_cast<float*>(p);
  }
  *out = g_bn_bwd_sums_addr[dev];
  return SFB_OK;
}
}  // namespace

extern "C" int sfb_bn_bwd_noalloc(const void* dz, int32_t dz_ld, int32_t dz_coff, const void* y, int32_t y_ld, int32_t y_coff,
                                  const float* scale, const float* shift, const float* mean, const float* invstd,
                                  const float* gamma, int32_t relu, int64_t M, int32_t C, void* dy, int32_t dy_ld,
                                  int32_t dy_coff, float* dgamma_out, float* dbeta_out, void* stream) {
  if (C > kNoAllocMaxC) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "sfb_bn_bwd_noalloc: C = %d exceeds SFB_NOALLOC_MAX_C = %d (static scratch)",
             (int)C, kNoAllocMaxC);
    return SFB_ERR_UNSUPPORTED;
  }
  if (M <= 0 || C <= 0 || (C & 7) || !mean || !invstd || !gamma) return SFB_ERR_BAD_ARG;
  if (!view_ok(dz, dz_ld, dz_coff, C) || !view_ok(y, y_ld, y_coff, C) || !view_ok(dy, dy_ld, dy_coff, C))
    return SFB_ERR_BAD_ARG;
  if (relu && (!scale || !shift)) return SFB_ERR_BAD_ARG;
  float* sums = nullptr;
  int rc = noalloc_sums(&sums);
  if (rc != SFB_OK) return rc;
  float* dbeta = sums;                   // sum g
  float* dgamma = sums + kNoAllocMaxC;   // sum g * xhat
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the previous call's sums: the accumulators start from zero on the caller's stream
  cudaError_t e = cudaMemsetAsync(sums, 0, sizeof(float) * 2 * kNoAllocMaxC, st);
  if (e != cudaSuccess) return cuda_fail(e, "sfb_bn_bwd_noalloc: cudaMemsetAsync");
  View vdz{(const __nv_bfloat16*)dz, dz_ld, dz_coff};
  View vy{(const __nv_bfloat16*)y, y_ld, y_coff};
  View vm{nullptr, 0, 0};
  ViewW vdy{(__nv_bfloat16*)dy, dy_ld, dy_coff};
  ViewW vg{nullptr, 0, 0};
  const int RL = kRedThreads / (C >> 3);
  const int grid_r = grid_for((M + (long long)RL * kRedUnroll - 1) / ((long long)RL * kRedUnroll), 1, 148 * 8);
  if (relu)
    bn_bwd_reduce_kernel<0><<<grid_r, kRedThreads, 0, st>>>(vdz, vy, vm, nullptr, scale, shift, mean, invstd, M, C, dgamma, dbeta);
  else
    bn_bwd_reduce_kernel<3><<<grid_r, kRedThreads, 0, st>>>(vdz, vy, vm, nullptr, scale, shift, mean, invstd, M, C, dgamma, dbeta);
  SFB_CHECK_LAUNCH("bn_bwd_reduce_kernel (noalloc)");
  const int grid_a = grid_for((M + (long long)RL * kBnUnroll - 1) / ((long long)RL * kBnUnroll), 1, 148 * 8);
  bn_bwd_apply_kernel<0, false><<<grid_a, kRedThreads, 0, st>>>(vdz, vy, vm, nullptr, scale, shift, mean, invstd, gamma, dgamma,
                                                                dbeta, relu, M, C, 1.0f / (float)M, vdy, vg, 0);
  SFB_CHECK_LAUNCH("bn_bwd_apply_kernel (noalloc)");
  if (dgamma_out) {
    e = cudaMemcpyAsync(dgamma_out, dgamma, sizeof(float) * C, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_fail(e, "sfb_bn_bwd_noalloc: copy of dgamma");
  }
  if (dbeta_out) {
    e = cudaMemcpyAsync(dbeta_out, dbeta, sizeof(float) * C, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_fail(e, "sfb_bn_bwd_noalloc: copy of dbeta");
  }
  return SFB_OK;
}
