// stage_u8.cuh -- input staging of raw clips: what the reference does on the host in
// dataset/chalearn_dataset.py:41-46,113-118 (transforms.ToTensor: HWC uint8 -> CHW fp32 / 255; transforms.Normalize:
// (x - mean) / std), dataset/chalearn_dataset.py:79-88 (RandomCrop(size, padding): one window per clip, zero padding
// applied AFTER normalisation) and train.py:136-140 (.cuda(), permute NTCHW -> NCTHW, channel split 0:5 / 5:20), done
// here in one pass from the uint8 frames straight into the bf16 space-to-depth operands of the two stem convolutions.
// The host -> device copy shrinks 4x (uint8 instead of fp32) and the fp32 clip never exists in HBM.
//
// HBM-bound byte shuffling: per frame row pair the block reads 2*W*Ct bytes (16-byte loads) and writes
// W/2 * Cs * 2 bytes per pathway as fully coalesced 16-byte chunks.  The value map has 256 inputs, so it is a table:
// entry v = bf16_rn(((float)v / 255.0f - mean) / std), evaluated with the reference's own fp32 operation order, which
// makes the output bit-identical to sfb_layout_input_s2d applied to the reference's fp32 clip.  The table is
// replicated per lane (32 KB, word [v][lane]) so that the random lookups of a warp never collide on a bank.
#pragma once
#include "common.cuh"

namespace sfb {

constexpr int kStageMaxPw = 4;
constexpr int kStageThreads = 256;
constexpr int kStageNever = -(1 << 30);  // offset of a channel / row that is padding

struct StagePathwayDev {
  int c0, C, t_stride, Cs;
  __nv_bfloat16* out;
};

struct StageParams {
  const uint8_t* frames;   // [N][T][H][W][Ct]
  long long total_bytes;
  int N, T, H, W, Ct;
  const int* crop_yx;      // [N][2] window offset (dy, dx) in source pixels, or null
  float mean, std;
  int n_pw;
  int row_bytes;           // shared-memory bytes per staged row (W*Ct + 32, multiple of 16)
  StagePathwayDev pw[kStageMaxPw];
};

__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// one row pair of one frame, as staged in shared memory
struct StageRow {
  int n, t, h2;
  int xlo, xhi;       // output x whose source pixel exists
  int base0, base1;   // shared-memory byte of output pixel 0, channel 0 of either row, or kStageNever
};

template <bool CROP>
__device__ __forceinline__ StageRow stage_issue_row(const StageParams& p, int rowi, uint8_t* buf, int tid) {
  // 32-bit index arithmetic (the launcher checks N*T*H/2 < 2^31): 64-bit divisions here once cost more
  // instructions per row than the shuffle itself (ncu r2c: 1.3 G warp instructions, 0.75 G of them row set-up)
  StageRow R;
  const int H2 = p.H >> 1;
  const int fi = rowi / H2;
  R.h2 = rowi - fi * H2;
  R.n = fi / p.T;
  R.t = fi - R.n * p.T;
  const long long f = fi;
  int cdy = 0, cdx = 0;
  if (CROP) { cdy = p.crop_yx[2 * R.n]; cdx = p.crop_yx[2 * R.n + 1]; }
  R.xlo = max(0, -cdx);
  R.xhi = min(p.W, p.W - cdx);
  bool any = false;
  for (int q = 0; q < p.n_pw; ++q) any |= (R.t % p.pw[q].t_stride) == 0;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int y = 2 * R.h2 + r + cdy;
    int b = kStageNever;
    if (any && y >= 0 && y < p.H && R.xlo < R.xhi) {
      // source byte of output pixel x, channel c: g0 + x*Ct + c
      const long long g0 = ((f * p.H + y) * p.W + cdx) * (long long)p.Ct;
      const long long first = g0 + (long long)R.xlo * p.Ct, last = g0 + (long long)R.xhi * p.Ct;
      const long long a0 = first & ~15ll;
      b = (int)(g0 - a0) + r * p.row_bytes;  // g0 - a0 may be negative when xlo > 0; only x >= xlo is addressed
      const int nw = (int)((last - a0 + 15) >> 4);
      uint8_t* dst = buf + r * p.row_bytes;
      for (int i = tid; i < nw; i += kStageThreads) {
        const long long a = a0 + 16ll * i;
        if (a + 16 <= p.total_bytes) {
          cp_async16(smem_u32(dst + 16 * i), p.frames + a);
        } else {  // the last 16 bytes of the tensor may be partial
          for (int k = 0; k < 16; ++k) dst[16 * i + k] = a + k < p.total_bytes ? p.frames[a + k] : (uint8_t)0;
        }
      }
    }
    if (r == 0) R.base0 = b; else R.base1 = b;
  }
  return R;
}

template <bool CROP>
__global__ void __launch_bounds__(kStageThreads, 4)
stage_clips_u8_kernel(const StageParams p) {
  extern __shared__ __align__(16) uint8_t smem_stage[];
  uint32_t* lut = reinterpret_cast<uint32_t*>(smem_stage);            // [256][32]
  uint4* tab = reinterpret_cast<uint4*>(smem_stage + 256 * 32 * 4);   // [kStageMaxPw][8 chunks] of 8 u16 entries
  uint8_t* rows = smem_stage + 256 * 32 * 4 + kStageMaxPw * 8 * 16;   // [2 buffers][2 rows][row_bytes]
  const int tid = threadIdx.x, lane = tid & 31;
  const uint32_t* lutl = lut + lane;  // this lane's copy: word [v][lane]
  const int H2 = p.H >> 1, W2 = p.W >> 1;
  const int pix2 = 2 * p.Ct;  // bytes of one s2d pixel's two source pixels in a row
  const int n_rows = p.N * p.T * H2;
  int rowi = blockIdx.x;
  StageRow nxt;
  if (rowi < n_rows) nxt = stage_issue_row<CROP>(p, rowi, rows, tid);  // travels while the tables are built
  cp_async_commit();
  for (int i = tid; i < 256 * 32; i += kStageThreads) {
    const int v = i >> 5;
    const float x = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), p.mean), p.std);
    lut[i] = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(x));
  }
  // s2d channel -> (row of the pair << 15 | byte offset inside the 2-pixel group), 0xFFFF = zero padding
  if (tid < kStageMaxPw * 64) {
    const int q = tid >> 6, ch = tid & 63;
    uint16_t e = 0xFFFFu;
    if (q < p.n_pw && ch < p.pw[q].Cs) {
      const int C = p.pw[q].C, quad = ch / C, c = ch - quad * C;
      if (quad < 4) e = (uint16_t)(((quad >> 1) << 15) | ((quad & 1) * p.Ct + p.pw[q].c0 + c));
    }
    reinterpret_cast<uint16_t*>(tab)[tid] = e;
  }
  int bsel = 0;
  for (; rowi < n_rows; rowi += gridDim.x, bsel ^= 1) {
    const StageRow R = nxt;
    // the next row pair travels (cp.async) into the other buffer while this one is shuffled out
    if (rowi + (int)gridDim.x < n_rows) nxt = stage_issue_row<CROP>(p, rowi + (int)gridDim.x, rows + (bsel ^ 1) * 2 * p.row_bytes, tid);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();  // this pair has landed (and, first time round, the tables are written)
    const int boff = bsel * 2 * p.row_bytes;
    for (int q = 0; q < p.n_pw; ++q) {
      const StagePathwayDev& pw = p.pw[q];
      if (R.t % pw.t_stride) continue;
      const int KV = pw.Cs >> 3, kshift = 31 - __clz(KV);
      const int To = p.T / pw.t_stride;
      uint4* orow = reinterpret_cast<uint4*>(pw.out + ((((size_t)R.n * To + R.t / pw.t_stride) * H2 + R.h2) * W2) * pw.Cs);
      const int nchunks = W2 << kshift;
      // the thread's 8 channels (the same for every chunk it writes: kStageThreads is a multiple of KV).  The inner
      // loop is branch-free: every element is fetched (padding channels / rows from a harmless address) and padding
      // is cleared by masks, 5 instructions per element instead of 29 with per-element branches (ncu r2c).
      const uint4 eq = tab[q * 8 + (tid & (KV - 1))];
      const uint32_t ew[4] = {eq.x, eq.y, eq.z, eq.w};
      int addr[8];        // offset into rows[] of the element for x2 = 0 (negative for x < xlo under a crop)
      uint32_t pm[4];     // per bf16 pair: 0xFFFF per live element
      uint32_t sel1 = 0;  // CROP: bit i set when element i comes from the odd source pixel
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint32_t e = (ew[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu;
        const int o = (int)(e & 0x7FFFu);
        const int b = (e & 0x8000u) ? R.base1 : R.base0;
        const bool live = e != 0xFFFFu && b != kStageNever;
        addr[i] = live ? boff + b + o : 0;
        if ((i & 1) == 0) pm[i >> 1] = 0;
        pm[i >> 1] |= live ? (0xFFFFu << ((i & 1) * 16)) : 0u;
        sel1 |= (o >= p.Ct ? 1u : 0u) << i;
      }
      for (int ci = tid; ci < nchunks; ci += kStageThreads) {
        const int x2 = ci >> kshift;
        const int xb = x2 * pix2;
        uint32_t by[8], hv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) by[i] = rows[addr[i] + xb];      // 8 independent byte loads in flight
#pragma unroll
        for (int i = 0; i < 8; ++i) hv[i] = lutl[by[i] << 5];
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = (hv[2 * j] | (hv[2 * j + 1] << 16)) & pm[j];
        if (CROP) {
          // source pixel 2*x2 (even) / 2*x2+1 (odd) outside [xlo, xhi) is RandomCrop's zero padding
          const uint32_t m0 = (2 * x2 >= R.xlo && 2 * x2 < R.xhi) ? 0xFFFFu : 0u;
          const uint32_t m1 = (2 * x2 + 1 >= R.xlo && 2 * x2 + 1 < R.xhi) ? 0xFFFFu : 0u;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t lo = (sel1 >> (2 * j)) & 1u ? m1 : m0, hi = (sel1 >> (2 * j + 1)) & 1u ? m1 : m0;
            w[j] &= lo | (hi << 16);
          }
        }
        orow[ci] = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
    __syncthreads();  // everyone is done with this buffer before the next iteration's loads overwrite it
  }
  cp_async_wait<0>();
}

}  // namespace sfb
