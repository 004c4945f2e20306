"""Shared kernel-level parity cases: libsfb200 kernels vs a plain PyTorch fp32 reference of the same op.

Used by tests/test_kernels_gpu.py (pytest -m gpu) and tools/kernel_diag.py (subprocess-isolated run).
Tolerances: operands are bf16, accumulation fp32 -> outputs compared after bf16 rounding with
max-abs error <= 1e-2 * max|ref| (conv/dgrad), 2e-2 (wgrad, long bf16 reductions).
"""
import torch
import torch.nn.functional as F

from video_classification_b200 import _ffi, ops
from video_classification_b200.ops import Act

DEV = "cuda"


def to_act(x, ld=None, coff=0):
    """fp32 NCTHW -> Act (bf16 NDHWC) inside a buffer with ld channels at offset coff (rest = garbage)."""
    N, Cc, T, H, W = x.shape
    ld = ld or Cc
    buf = torch.randn((N, T, H, W, ld), device=x.device).bfloat16()
    buf[..., coff:coff + Cc] = x.permute(0, 2, 3, 4, 1).bfloat16()
    return Act(buf, coff, Cc)


def from_act(a: Act):
    return a.torch_view().float().permute(0, 4, 1, 2, 3).contiguous()


def relerr(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


# (name, N, T, H, W, Cin, Cout, k, s)  -- every conv flavour of the SlowFast graph at small sizes
CONV_CASES = [
    ("1x1x1_64_256", 2, 4, 12, 12, 64, 256, (1, 1, 1), (1, 1, 1)),
    ("1x1x1_80_64_cat", 2, 4, 12, 12, 80, 64, (1, 1, 1), (1, 1, 1)),
    ("1x3x3_64_64", 2, 3, 14, 14, 64, 64, (1, 3, 3), (1, 1, 1)),
    ("1x3x3_128_128_s2", 1, 4, 14, 14, 128, 128, (1, 3, 3), (1, 2, 2)),
    ("3x1x1_256_128", 2, 8, 7, 7, 256, 128, (3, 1, 1), (1, 1, 1)),
    ("3x1x1_1024_256", 1, 8, 7, 7, 1024, 256, (3, 1, 1), (1, 1, 1)),
    ("1x1x1_320_512_s2", 2, 4, 14, 14, 320, 512, (1, 1, 1), (1, 2, 2)),
    ("stem_slow_8_64", 1, 2, 32, 32, 8, 64, (1, 7, 7), (1, 2, 2)),
    ("stem_fast_16_8", 1, 4, 32, 32, 16, 8, (1, 7, 7), (1, 2, 2)),
    ("fast_1x1x1_8_32", 2, 8, 12, 12, 8, 32, (1, 1, 1), (1, 1, 1)),
    ("fast_3x1x1_32_8", 2, 8, 12, 12, 32, 8, (3, 1, 1), (1, 1, 1)),
    ("fast_1x3x3_16_16", 2, 8, 12, 12, 16, 16, (1, 3, 3), (1, 1, 1)),
    ("lateral_3x1x1_32_64_s4", 2, 16, 8, 8, 32, 64, (3, 1, 1), (4, 1, 1)),
    ("1x1x1_512_2048", 1, 4, 7, 7, 512, 2048, (1, 1, 1), (1, 1, 1)),
    # larger / odd shapes: several persistent waves, partial patches, every A-box width of the TMA kernel
    ("big_1x3x3_64_64", 3, 8, 56, 56, 64, 64, (1, 3, 3), (1, 1, 1)),
    ("big_1x1x1_320_128", 2, 8, 56, 56, 320, 128, (1, 1, 1), (1, 1, 1)),
    ("big_3x1x1_640_256", 2, 8, 14, 14, 640, 256, (3, 1, 1), (1, 1, 1)),
    ("odd_1x3x3_32_32", 2, 5, 13, 11, 32, 32, (1, 3, 3), (1, 1, 1)),
    ("odd_3x1x1_16_48", 1, 7, 9, 10, 16, 48, (3, 1, 1), (1, 1, 1)),
    ("stem_slow_16_64", 2, 2, 64, 64, 16, 64, (1, 7, 7), (1, 2, 2)),
    ("ref_1x3x3_512_512_2x2", 2, 20, 2, 2, 512, 512, (1, 3, 3), (1, 1, 1)),
    # strided convs of the Fast pathway / first lateral (16- and 32-channel dgrad operands) and odd sizes
    ("fast_1x3x3_16_16_s2", 2, 8, 14, 14, 16, 16, (1, 3, 3), (1, 2, 2)),
    ("fast_1x3x3_32_32_s2", 2, 4, 14, 14, 32, 32, (1, 3, 3), (1, 2, 2)),
    ("lateral0_3x1x1_8_16_s4", 2, 16, 6, 6, 8, 16, (3, 1, 1), (4, 1, 1)),
    ("odd_1x3x3_64_64_s2", 1, 3, 13, 11, 64, 64, (1, 3, 3), (1, 2, 2)),
    ("fast_1x1x1_32_64_s2", 2, 4, 14, 14, 32, 64, (1, 1, 1), (1, 2, 2)),
    # halo-mode patches of the (1,3,3) 64-channel convs: 7-wide maps (pixel-folded Fast stages), a 15-wide map (second
    # patch column of one pixel), partial patch rows
    ("halo_1x3x3_64_64_w7", 2, 3, 28, 7, 64, 64, (1, 3, 3), (1, 1, 1)),
    ("halo_1x3x3_64_64_w15", 1, 2, 17, 15, 64, 64, (1, 3, 3), (1, 1, 1)),
]


def _pad(k):
    return tuple(kk // 2 for kk in k)


def run_conv_case(case, impl, variant="plain", seed=0):
    """variant: plain | pro (BN+ReLU prologue) | epi (affine+residual+relu epilogue, strided views) | stats"""
    name, N, T, H, W, Ci, Co, k, s = case
    g = torch.Generator(device=DEV).manual_seed(seed)
    x = torch.randn((N, Ci, T, H, W), device=DEV, generator=g)
    w = torch.randn((Co, Ci) + k, device=DEV, generator=g) / (Ci * k[0] * k[1] * k[2]) ** 0.5
    p = _pad(k)
    use_views = variant == "epi"
    xa = to_act(x, ld=Ci + 16 if use_views else None, coff=8 if use_views else 0)
    xq = from_act(xa)
    wq = w.bfloat16().float()
    To, Ho, Wo = ops.conv_out_dims((T, H, W), k, s, p)
    out = Act.empty(N, To, Ho, Wo, Co + (24 if use_views else 0), coff=16 if use_views else 0, C=Co, zero=True)
    kw = {}
    xin = xq
    if variant == "pro":
        sc = (torch.rand(Ci, device=DEV, generator=g) + 0.5)
        sh = torch.randn(Ci, device=DEV, generator=g) * 0.3
        kw["pro"] = (sc, sh, True)
        xin = torch.relu(xq * sc.view(1, -1, 1, 1, 1) + sh.view(1, -1, 1, 1, 1)).bfloat16().float()
    ref = F.conv3d(xin, wq, stride=s, padding=p)
    stats = None
    if variant == "epi":
        es = torch.rand(Co, device=DEV, generator=g) + 0.5
        eb = torch.randn(Co, device=DEV, generator=g) * 0.3
        r = torch.randn((N, Co, To, Ho, Wo), device=DEV, generator=g)
        ra = to_act(r, ld=Co + 8, coff=8)
        kw.update(epi_affine=(es, eb), residual=ra, relu=True)
        ref = torch.relu(ref * es.view(1, -1, 1, 1, 1) + eb.view(1, -1, 1, 1, 1) + from_act(ra))
    if variant == "stats":
        stats = torch.zeros(2 * Co, device=DEV)
        kw["stats"] = stats
    wp = ops.pack_weight(w.contiguous(), xa.C)
    ops.conv_gemm(xa, wp, out, k, s, p, impl=impl, **kw)
    torch.cuda.synchronize()
    got = from_act(out)
    res = {"err": relerr(got, ref)}
    if variant == "stats":
        cnt = ref.numel() / Co
        res["err_sum"] = relerr(stats[:Co] / cnt, ref.mean(dim=(0, 2, 3, 4))) if ref.mean(dim=(0, 2, 3, 4)).abs().max() > 1e-3 else 0.0
        res["err_sq"] = relerr(stats[Co:] / cnt, (ref * ref).mean(dim=(0, 2, 3, 4)))
    if use_views:  # untouched channels of the output buffer must stay zero
        res["clobber"] = float(out.buf[..., :16].abs().max().item() + out.buf[..., 16 + Co:].abs().max().item())
    return res


def run_dgrad_case(case, impl, seed=0, accumulate=False):
    name, N, T, H, W, Ci, Co, k, s = case
    g = torch.Generator(device=DEV).manual_seed(seed)
    p = _pad(k)
    To, Ho, Wo = ops.conv_out_dims((T, H, W), k, s, p)
    dy = torch.randn((N, Co, To, Ho, Wo), device=DEV, generator=g)
    w = torch.randn((Co, Ci) + k, device=DEV, generator=g) / (Co * k[0] * k[1] * k[2]) ** 0.5
    dya = to_act(dy)
    wq = w.bfloat16().float()
    x = torch.zeros((N, Ci, T, H, W), device=DEV, requires_grad=True)
    F.conv3d(x, wq, stride=s, padding=p).backward(from_act(dya))
    ref = x.grad
    # the tensor-core path takes the class-major operand (one K range per stride parity class); the SIMT / gather
    # cross-check kernels keep the tap-major layout
    wpt = ops.pack_weight_dgrad(w.contiguous(), Co, s) if impl == _ffi.IMPL_TC else ops.pack_weight(w.contiguous(), Co, transposed=True)
    dx = Act.empty(N, T, H, W, Ci, zero=True)
    kw = {}
    if accumulate:
        base = torch.randn((N, Ci, T, H, W), device=DEV, generator=g)
        dx = to_act(base)
        ref = ref + from_act(dx)
        kw["residual"] = dx
    ops.conv_gemm(dya, wpt, dx, k, s, p, transposed=True, impl=impl, **kw)
    torch.cuda.synchronize()
    return {"err": relerr(from_act(dx), ref)}


def run_wgrad_case(case, impl, seed=0, pro=False, ci_real=None):
    name, N, T, H, W, Ci, Co, k, s = case
    g = torch.Generator(device=DEV).manual_seed(seed)
    p = _pad(k)
    To, Ho, Wo = ops.conv_out_dims((T, H, W), k, s, p)
    x = torch.randn((N, Ci, T, H, W), device=DEV, generator=g)
    dy = torch.randn((N, Co, To, Ho, Wo), device=DEV, generator=g)
    xa, dya = to_act(x), to_act(dy, ld=Co + 8, coff=8)
    xin = from_act(xa)
    kw = {}
    if pro:
        sc = (torch.rand(Ci, device=DEV, generator=g) + 0.5)
        sh = torch.randn(Ci, device=DEV, generator=g) * 0.3
        kw["pro"] = (sc, sh, True)
        xin = torch.relu(xin * sc.view(1, -1, 1, 1, 1) + sh.view(1, -1, 1, 1, 1)).bfloat16().float()
    w = torch.zeros((Co, Ci) + k, device=DEV, requires_grad=True)
    F.conv3d(xin, w, stride=s, padding=p).backward(from_act(dya))
    ref = w.grad
    cr = ci_real or Ci
    dw = torch.zeros((Co, cr) + k, device=DEV)
    ops.conv_wgrad(xa, dya, dw, k, s, p, impl=impl, **kw)
    torch.cuda.synchronize()
    return {"err": relerr(dw, ref[:, :cr])}
