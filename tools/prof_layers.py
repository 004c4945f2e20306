"""Single-layer timing / ncu driver: launches selected kernels of the HYB training step (batch 64 shapes, SURVEY.md
Appendix A.1) in isolation through the C ABI, with CUDA-event timing and the algorithmic bytes / FLOPs of each, so a
kernel can be compared with its roofline without running the whole step (and so `ncu --set full -k regex:...` has a
short command to replay).

    python tools/prof_layers.py [--n 64] [--only name,name] [--reps 5]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from video_classification_b200 import _ffi, ops  # noqa: E402
from video_classification_b200.ops import Act  # noqa: E402

DEV = "cuda"
NOWARM = False  # --nowarm: no untimed first launch (ncu captures: one launch per layer)

# name, kind, T, H, W, Ci (real), Co, k, s     kind: fprop | fprop_stats | dgrad | dgrad_acc | wgrad
LAYERS = [
    ("stem_fast", "fprop_stats", 32, 224, 224, 15, 8, (1, 7, 7), (1, 2, 2)),
    ("stem_fast", "wgrad", 32, 224, 224, 15, 8, (1, 7, 7), (1, 2, 2)),
    ("stem_slow", "fprop_stats", 8, 224, 224, 5, 64, (1, 7, 7), (1, 2, 2)),
    ("stem_slow", "wgrad", 8, 224, 224, 5, 64, (1, 7, 7), (1, 2, 2)),
    ("stem_fast", "s2d_fprop", 32, 224, 224, 15, 8, (1, 7, 7), (1, 2, 2)),
    ("stem_fast", "s2d_wgrad", 32, 224, 224, 15, 8, (1, 7, 7), (1, 2, 2)),
    ("stem_slow", "s2d_fprop", 8, 224, 224, 5, 64, (1, 7, 7), (1, 2, 2)),
    ("stem_slow", "s2d_wgrad", 8, 224, 224, 5, 64, (1, 7, 7), (1, 2, 2)),
    ("slow_res2_c_64_256", "fprop_stats", 8, 56, 56, 64, 256, (1, 1, 1), (1, 1, 1)),
    ("slow_res2_c_64_256", "dgrad", 8, 56, 56, 64, 256, (1, 1, 1), (1, 1, 1)),
    ("slow_res2_c_64_256", "wgrad", 8, 56, 56, 64, 256, (1, 1, 1), (1, 1, 1)),
    ("slow_res2_a_256_64", "fprop_stats", 8, 56, 56, 256, 64, (1, 1, 1), (1, 1, 1)),
    ("slow_res2_a_256_64", "dgrad_acc", 8, 56, 56, 256, 64, (1, 1, 1), (1, 1, 1)),
    ("slow_res2_b_64_64", "fprop_stats", 8, 56, 56, 64, 64, (1, 3, 3), (1, 1, 1)),
    ("slow_res2_b_64_64", "dgrad", 8, 56, 56, 64, 64, (1, 3, 3), (1, 1, 1)),
    ("slow_res2_b_64_64", "wgrad", 8, 56, 56, 64, 64, (1, 3, 3), (1, 1, 1)),
    ("slow_res4_a_1280_256", "fprop_stats", 8, 14, 14, 1280, 256, (3, 1, 1), (1, 1, 1)),
    ("slow_res4_a_1280_256", "wgrad", 8, 14, 14, 1280, 256, (3, 1, 1), (1, 1, 1)),
    ("slow_res4_a_1024_256", "wgrad", 8, 14, 14, 1024, 256, (3, 1, 1), (1, 1, 1)),
    ("slow_res4_b_256_256", "wgrad", 8, 14, 14, 256, 256, (1, 3, 3), (1, 1, 1)),
    ("slow_res4_b_256_256", "fprop_stats", 8, 14, 14, 256, 256, (1, 3, 3), (1, 1, 1)),
    ("slow_res4_c_256_1024", "wgrad", 8, 14, 14, 256, 1024, (1, 1, 1), (1, 1, 1)),
    ("slow_res4_c_256_1024", "fprop_stats", 8, 14, 14, 256, 1024, (1, 1, 1), (1, 1, 1)),
    ("slow_res5_c_512_2048", "wgrad", 8, 7, 7, 512, 2048, (1, 1, 1), (1, 1, 1)),
    ("slow_res3_c_128_512", "fprop_stats", 8, 28, 28, 128, 512, (1, 1, 1), (1, 1, 1)),
    ("slow_res3_c_128_512", "dgrad", 8, 28, 28, 128, 512, (1, 1, 1), (1, 1, 1)),
    ("slow_res3_b_128_128_s2", "dgrad", 8, 56, 56, 128, 128, (1, 3, 3), (1, 2, 2)),
    ("slow_res3_s_320_512_s2", "dgrad_acc", 8, 56, 56, 320, 512, (1, 1, 1), (1, 2, 2)),
    ("fast_res2_b_8_8", "fprop_stats", 32, 56, 56, 8, 8, (1, 3, 3), (1, 1, 1)),
    ("fast_res2_b_8_8", "dgrad", 32, 56, 56, 8, 8, (1, 3, 3), (1, 1, 1)),
    ("fast_res2_b_8_8", "wgrad", 32, 56, 56, 8, 8, (1, 3, 3), (1, 1, 1)),
    ("fast_res2_a_32_8", "fprop_stats", 32, 56, 56, 32, 8, (3, 1, 1), (1, 1, 1)),
    ("fast_res2_c_8_32", "fprop_stats", 32, 56, 56, 8, 32, (1, 1, 1), (1, 1, 1)),
    ("fast_res2_c_8_32", "dgrad", 32, 56, 56, 8, 32, (1, 1, 1), (1, 1, 1)),
    ("fast_res3_b_16_16_s2", "dgrad", 32, 56, 56, 16, 16, (1, 3, 3), (1, 2, 2)),
    ("lateral1_32_64_s4", "dgrad_acc", 32, 56, 56, 32, 64, (3, 1, 1), (4, 1, 1)),
    ("layout_fast", "layout_s2d", 32, 224, 224, 15, 8, (1, 7, 7), (1, 2, 2)),
    ("stage_u8_hyb", "stage_u8", 32, 224, 224, 15, 8, (1, 7, 7), (1, 2, 2)),
    ("layout_slow", "layout_s2d", 8, 224, 224, 5, 64, (1, 7, 7), (1, 2, 2)),
    ("bn_bwd_256", "bn_bwd", 8, 56, 56, 256, 256, (1, 1, 1), (1, 1, 1)),
    ("bn_apply_256", "bn_apply", 8, 56, 56, 256, 256, (1, 1, 1), (1, 1, 1)),
    ("bn_tail_256", "bn_tail", 8, 56, 56, 256, 256, (1, 1, 1), (1, 1, 1)),
    ("bn_tail_bwd_256", "bn_tail_bwd", 8, 56, 56, 256, 256, (1, 1, 1), (1, 1, 1)),
]


def r16(c):
    return (c + 15) // 16 * 16


def run_layer(N, spec, reps):
    name, kind, T, H, W, Ci, Co, k, s = spec
    p = tuple(kk // 2 for kk in k)
    To, Ho, Wo = ops.conv_out_dims((T, H, W), k, s, p)
    g = torch.Generator(device=DEV).manual_seed(0)
    x = Act.empty(N, T, H, W, r16(Ci), C=Ci, zero=True)
    x.torch_view().copy_(torch.randn((N, T, H, W, Ci), device=DEV, generator=g))
    y = Act.empty(N, To, Ho, Wo, r16(Co), C=Co, zero=True)
    y.torch_view().copy_(torch.randn((N, To, Ho, Wo, Co), device=DEV, generator=g))
    w = torch.randn((Co, Ci) + k, device=DEV, generator=g) * 0.05
    M = N * To * Ho * Wo
    taps = k[0] * k[1] * k[2]
    flops = 2.0 * M * Co * Ci * taps
    in_b, out_b = N * T * H * W * Ci * 2, M * Co * 2
    xg = Act(x.buf, 0, r16(Ci))
    if kind == "layout_s2d":
        del x, y
        xf = torch.randn((N, Ci, T, H, W), device=DEV, generator=g)
        fn = lambda: ops.layout_input_s2d(xf)
        flops = 0.0
        bytes_ = xf.numel() * 4 + N * T * (H // 2) * (W // 2) * (32 if 4 * Ci <= 32 else 64) * 2
    elif kind == "stage_u8":
        del x, y
        fr = torch.randint(0, 256, (N, T, H, W, 21), device=DEV, dtype=torch.uint8)
        pws = ((0, 5, 4), (5, 15, 1))
        outs = ops.stage_clips_u8(fr, pws)
        fn = lambda: ops.stage_clips_u8(fr, pws, outs=outs)
        flops = 0.0
        bytes_ = fr.numel() + sum(o.numel() * 2 for o in outs)
    elif kind in ("s2d_fprop", "s2d_wgrad"):
        xf = torch.randn((N, Ci, T, H, W), device=DEV, generator=g)
        xs = ops.layout_input_s2d(xf)
        del xf
        yv = Act(y.buf[..., :Co].contiguous(), 0, Co)  # the engine keeps the stem output unpadded (ld = Co)
        if kind == "s2d_fprop":
            wp = ops.pack_weight_stem(w)
            stats = torch.zeros(2 * Co, device=DEV)
            fn = lambda: ops.stem_conv_fprop(xs, Ci, wp, yv, stats)
        else:
            dw = torch.zeros((Co, Ci) + k, device=DEV)
            fn = lambda: ops.stem_conv_wgrad(xs, Ci, yv, dw)
        bytes_ = in_b + out_b
    elif kind in ("fprop", "fprop_stats"):
        wp = ops.pack_weight(w, xg.C)
        stats = torch.zeros(2 * Co, device=DEV) if kind == "fprop_stats" else None
        fn = lambda: ops.conv_gemm(xg, wp, y, k, s, p, stats=stats)
        bytes_ = in_b + out_b
    elif kind in ("dgrad", "dgrad_acc"):
        yg = Act(y.buf, 0, r16(Co))
        wpt = ops.pack_weight_dgrad(w, yg.C, s)
        res = xg if kind == "dgrad_acc" else None
        fn = lambda: ops.conv_gemm(yg, wpt, xg, k, s, p, transposed=True, residual=res)
        bytes_ = in_b * (2 if res is not None else 1) + out_b
    elif kind == "wgrad":
        dw = torch.zeros((Co, Ci) + k, device=DEV)
        fn = lambda: ops.conv_wgrad(xg, y, dw, k, s, p)
        bytes_ = in_b + out_b
    elif kind == "bn_bwd":
        C_ = Co
        sc, sh = torch.rand(C_, device=DEV) + 0.5, torch.randn(C_, device=DEV)
        mean, inv = torch.randn(C_, device=DEV), torch.rand(C_, device=DEV) + 0.5
        gam = torch.rand(C_, device=DEV) + 0.5
        dga, dbe = torch.zeros(C_, device=DEV), torch.zeros(C_, device=DEV)
        dz = Act.empty(N, To, Ho, Wo, C_)
        dz.buf.normal_()
        dy = Act.empty(N, To, Ho, Wo, C_)

        def fn():
            ops.bn_bwd_reduce(dz, y, None, sc, sh, mean, inv, True, dga, dbe)
            ops.bn_bwd_apply(dz, y, None, sc, sh, mean, inv, gam, dga, dbe, True, dy)
        flops, bytes_ = 0.0, 5 * out_b  # reduce reads dz,y ; apply reads dz,y writes dy
    elif kind in ("bn_tail", "bn_tail_bwd"):
        # residual tail O = relu(bn(yc) + X) with the packed ReLU mask, and its backward (reduce + apply with g_out)
        C_ = Co
        sc, sh = torch.rand(C_, device=DEV) + 0.5, torch.randn(C_, device=DEV)
        mean, inv = torch.randn(C_, device=DEV), torch.rand(C_, device=DEV) + 0.5
        gam = torch.rand(C_, device=DEV) + 0.5
        dga, dbe = torch.zeros(C_, device=DEV), torch.zeros(C_, device=DEV)
        X = Act.empty(N, To, Ho, Wo, C_)
        X.buf.normal_()
        O = Act.empty(N, To, Ho, Wo, C_)
        bits = torch.zeros(M * C_ // 8, dtype=torch.uint8, device=DEV)
        if kind == "bn_tail":
            fn = lambda: ops.bn_add_relu(y, sc, sh, X, None, None, O, mask_bits=bits)
            flops, bytes_ = 0.0, 3 * out_b
        else:
            ops.bn_add_relu(y, sc, sh, X, None, None, O, mask_bits=bits)
            dO = X  # any gradient-like tensor
            dy = Act.empty(N, To, Ho, Wo, C_)
            gX = Act.empty(N, To, Ho, Wo, C_)

            def fn():
                ops.bn_bwd_reduce(dO, y, None, sc, sh, mean, inv, True, dga, dbe, mask_bits=bits)
                ops.bn_bwd_apply(dO, y, None, sc, sh, mean, inv, gam, dga, dbe, True, dy, gX, mask_bits=bits)
            flops, bytes_ = 0.0, 6 * out_b  # reduce reads dO,yc ; apply reads dO,yc writes dyc,gX
    elif kind == "bn_apply":
        C_ = Co
        sc, sh = torch.rand(C_, device=DEV) + 0.5, torch.randn(C_, device=DEV)
        z = Act.empty(N, To, Ho, Wo, C_)
        fn = lambda: ops.bn_add_relu(y, sc, sh, None, None, None, z)
        flops, bytes_ = 0.0, 2 * out_b
    else:
        raise ValueError(kind)
    if not NOWARM:
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    ts = []
    for _ in range(reps):
        flush.zero_()  # L2 flush between timed launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    return dict(name=name, kind=kind, ms=ms, tflops=flops / ms / 1e9, gbs=bytes_ / ms / 1e6, alg_mb=bytes_ / 1e6,
                alg_gflop=flops / 1e9)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=64)
    ap.add_argument("--only", default="")
    ap.add_argument("--kinds", default="")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--json", default="")
    ap.add_argument("--nowarm", action="store_true")
    a = ap.parse_args()
    global NOWARM
    NOWARM = a.nowarm
    only = set(filter(None, a.only.split(",")))
    kinds = set(filter(None, a.kinds.split(",")))
    rows = []
    for spec in LAYERS:
        if only and spec[0] not in only:
            continue
        if kinds and spec[1] not in kinds:
            continue
        r = run_layer(a.n, spec, a.reps)
        rows.append(r)
        print(f"{r['name']:26s} {r['kind']:12s} {r['ms']:8.3f} ms  {r['tflops']:7.1f} TF/s  {r['gbs']:7.0f} GB/s "
              f"(alg {r['alg_mb']:8.1f} MB, {r['alg_gflop']:8.1f} GF)", flush=True)
        torch.cuda.empty_cache()
    if a.json:
        json.dump(rows, open(a.json, "w"), indent=1)
    print(f"libsfb200 launches: {_ffi.launch_count()}")


if __name__ == "__main__":
    main()
