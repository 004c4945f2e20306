// pack_batch.cuh -- one launch for all weight-operand packing jobs of a step (see include/sfb200.h, sfb_pack_plan).
#pragma once
#include "elementwise.cuh"
#include "stem_win.cuh"

namespace sfb {

// ---- batched operand packing -------------------------------------------------------------------------------------
// A training step re-derives every bf16 GEMM operand from the fp32 master weights the optimizer just changed: ~330
// launches of a few microseconds each (1.5 ms of a 72 ms step although they move 0.3 GB).  The engine describes them
// once as a device-resident job table; one launch walks it, each block finding its job by binary search on the
// jobs' first-block index.
struct PackJobDev {
  int kind;  // SfbPackJob kinds: 0 pack_weight, 1 pack_weight_dgrad, 2 fold_weight, 3 pack_weight_stem
  int Co, Ci, kT, kH, kW, Cg, transposed, f, ldo, ldi, rows_pad, k_pad;
  int block0, nblocks;
  const float* src;
  void* dst;
  DgradClassTable tb;
};
__global__ void __launch_bounds__(256) pack_batch_kernel(const PackJobDev* __restrict__ jobs, int n) {
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (jobs[mid].block0 <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const PackJobDev& j = jobs[lo];
  const long long first = (long long)(blockIdx.x - j.block0) * blockDim.x + threadIdx.x;
  const long long step = (long long)j.nblocks * blockDim.x;
  switch (j.kind) {
    case 0:
      pack_weight_body(j.src, j.Co, j.Ci, j.kT * j.kH * j.kW, j.Cg, j.transposed, (__nv_bfloat16*)j.dst, j.rows_pad,
                       j.k_pad, first, step);
      break;
    case 1:
      pack_weight_dgrad_body(j.src, j.Co, j.Ci, j.Cg, j.tb, (__nv_bfloat16*)j.dst, j.rows_pad, j.k_pad, first, step);
      break;
    case 2:
      fold_weight_body(j.src, j.Co, j.Ci, j.kT, j.kH, j.kW, j.f, j.ldo, j.ldi, (float*)j.dst, first, step);
      break;
    default:
      pack_weight_stem_body(j.src, j.Co, j.Ci, j.Cg, (__nv_bfloat16*)j.dst, j.rows_pad, first, step);
      break;
  }
}

}  // namespace sfb
