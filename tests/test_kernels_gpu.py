"""Kernel-level parity (B200): every libsfb200 kernel vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from video_classification_b200 import _ffi, ops  # noqa: E402
from video_classification_b200.ops import Act  # noqa: E402

import kernel_cases as kc  # noqa: E402

DEV = "cuda"


# ------------------------------------------------------------------ elementwise / pooling / head
def test_layout_input_strided():
    x = torch.randn(2, 6, 21, 10, 12, device=DEV)  # (N,T,C,H,W) like the reference loader
    v = x.permute(0, 2, 1, 3, 4)[:, 5:20]          # train.py:136-140: non-contiguous NCTHW view
    a = ops.layout_input(v, 16)
    got = a.buf.float()
    ref = torch.zeros_like(got)
    ref[..., :15] = v.permute(0, 2, 3, 4, 1).bfloat16().float()
    assert torch.equal(got, ref)


def test_pack_weight_roundtrip():
    w = torch.randn(24, 5, 1, 7, 7, device=DEV)
    wp = ops.pack_weight(w, 8)
    rows, kpad = wp.shape
    ref = torch.zeros(rows, kpad, device=DEV)
    wr = w.reshape(24, 5, 49).permute(0, 2, 1)  # (co, tap, ci)
    tmp = torch.zeros(24, 49, 8, device=DEV)
    tmp[:, :, :5] = wr
    ref[:24, :49 * 8] = tmp.reshape(24, -1)
    assert torch.equal(wp.float(), ref.bfloat16().float())
    wpt = ops.pack_weight(w, 24, transposed=True)
    tmp = w.reshape(24, 5, 49).permute(1, 2, 0)  # (ci, tap, co)
    ref = torch.zeros(wpt.shape, device=DEV)
    ref[:5, :49 * 24] = tmp.reshape(5, -1)
    assert torch.equal(wpt.float(), ref.bfloat16().float())


def _bn_ref(y, gamma, beta, eps=1e-5):
    mean = y.mean(dim=(0, 2, 3, 4))
    var = y.var(dim=(0, 2, 3, 4), unbiased=False)
    invstd = (var + eps).rsqrt()
    return mean, var, invstd


def test_bn_finalize_and_fold():
    C_, cnt = 64, 1000
    y = torch.randn(cnt, C_, device=DEV) * 2 + 0.5
    stats = torch.cat([y.sum(0), (y * y).sum(0)])
    gamma, beta = torch.rand(C_, device=DEV) + 0.5, torch.randn(C_, device=DEV)
    rm, rv = torch.zeros(C_, device=DEV), torch.ones(C_, device=DEV)
    sc, sh, mean, invstd = (torch.empty(C_, device=DEV) for _ in range(4))
    ops.bn_finalize(stats, cnt, gamma, beta, 1e-5, 0.1, rm, rv, sc, sh, mean, invstd)
    m = y.mean(0)
    v = y.var(0, unbiased=False)
    torch.testing.assert_close(mean, m, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(invstd, (v + 1e-5).rsqrt(), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(sc, gamma * (v + 1e-5).rsqrt(), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(sh, beta - m * sc, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(rm, 0.1 * m, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(rv, 0.9 + 0.1 * y.var(0, unbiased=True), rtol=1e-4, atol=1e-5)
    ops.bn_fold(gamma, beta, rm, rv, 1e-5, sc, sh)
    torch.testing.assert_close(sc, gamma / (rv + 1e-5).sqrt(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(sh, beta - rm * sc, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("C_,with_rbn", [(64, False), (256, True), (8, False), (1280, True)])
def test_bn_add_relu(C_, with_rbn):
    y = torch.randn(2, C_, 3, 6, 6, device=DEV)
    r = torch.randn(2, C_, 3, 6, 6, device=DEV)
    ya, ra = kc.to_act(y, ld=C_ + 8, coff=8), kc.to_act(r)
    s, b = torch.rand(C_, device=DEV) + 0.5, torch.randn(C_, device=DEV)
    rs, rb = (torch.rand(C_, device=DEV) + 0.5, torch.randn(C_, device=DEV)) if with_rbn else (None, None)
    out = Act.empty(2, 3, 6, 6, C_ + 16, coff=16, C=C_, zero=True)
    bits = torch.zeros(2 * 3 * 6 * 6, C_ // 8, dtype=torch.uint8, device=DEV)
    ops.bn_add_relu(ya, s, b, ra, rs, rb, out, mask_bits=bits)
    bc = lambda t: t.view(1, -1, 1, 1, 1)
    rr = kc.from_act(ra)
    if with_rbn:
        rr = rr * bc(rs) + bc(rb)
    ref = torch.relu(kc.from_act(ya) * bc(s) + bc(b) + rr)
    assert kc.relerr(kc.from_act(out), ref) < 6e-3
    assert out.buf[..., :16].abs().max() == 0
    # packed ReLU mask: bit i of byte cv = (out[..., 8*cv+i] > 0), exactly
    o = out.torch_view().reshape(-1, C_ // 8, 8) > 0
    w = (2 ** torch.arange(8, device=DEV, dtype=torch.int32)).view(1, 1, 8)
    assert torch.equal(bits, (o.to(torch.int32) * w).sum(-1).to(torch.uint8))


@pytest.mark.parametrize("C_,fold_f,with_r", [(64, 0, False), (256, 0, True), (16, 4, False), (32, 2, True)])
def test_bn_add_relu_raw_equals_finalize_then_apply(C_, fold_f, with_r):
    """the fused-finalize form gives the same coefficients, running statistics and output as
    sfb_bn_finalize (+ sfb_unfold_stats) followed by sfb_bn_add_relu"""
    N, T, H, W = 2, 3, 6, 8
    M = N * T * H * W
    y = torch.randn(N, C_, T, H, W, device=DEV) * 1.7 + 0.4
    r = torch.randn(N, C_, T, H, W, device=DEV)
    ya, ra = kc.to_act(y), kc.to_act(r)
    yq = kc.from_act(ya).permute(0, 2, 3, 4, 1).reshape(M, C_)
    sums = torch.cat([yq.sum(0), (yq * yq).sum(0)])

    def bn_state():
        g = torch.Generator(device=DEV).manual_seed(3)
        return dict(gamma=torch.rand(C_, device=DEV, generator=g) + 0.5, beta=torch.randn(C_, device=DEV, generator=g),
                    rm=torch.randn(C_, device=DEV, generator=g), rv=torch.rand(C_, device=DEV, generator=g) + 0.5,
                    sc=torch.empty(C_, device=DEV), sh=torch.empty(C_, device=DEV), mean=torch.empty(C_, device=DEV),
                    inv=torch.empty(C_, device=DEV))
    a, b = bn_state(), bn_state()
    ra_, rb_ = bn_state(), bn_state()
    # reference path
    ops.bn_finalize(sums, M, a["gamma"], a["beta"], 1e-5, 0.1, a["rm"], a["rv"], a["sc"], a["sh"], a["mean"], a["inv"])
    if with_r:
        rq = kc.from_act(ra).permute(0, 2, 3, 4, 1).reshape(M, C_)
        rsums = torch.cat([rq.sum(0), (rq * rq).sum(0)])
        ops.bn_finalize(rsums, M, ra_["gamma"], ra_["beta"], 1e-5, 0.1, ra_["rm"], ra_["rv"], ra_["sc"], ra_["sh"],
                        ra_["mean"], ra_["inv"])
    out_ref = Act.empty(N, T, H, W, C_)
    ops.bn_add_relu(ya, a["sc"], a["sh"], ra if with_r else None, ra_["sc"] if with_r else None,
                    ra_["sh"] if with_r else None, out_ref)
    # fused path; folded layout: the sums spread over fold_f column groups of width ld
    if fold_f:
        ld = C_ + 8
        raw = torch.zeros(2 * fold_f * ld, device=DEV)
        w = torch.rand(fold_f, 1, device=DEV)
        w = w / w.sum(0, keepdim=True)
        raw[:fold_f * ld].view(fold_f, ld)[:, :C_] = w * sums[:C_]
        raw[fold_f * ld:].view(fold_f, ld)[:, :C_] = w * sums[C_:]
    else:
        ld, raw = 0, sums
    qy = ops.bn_raw(raw, M, b["gamma"], b["beta"], 1e-5, 0.1, b["rm"], b["rv"], b["sc"], b["sh"], b["mean"], b["inv"],
                    fold_f, ld)
    qr = ops.bn_raw(rsums, M, rb_["gamma"], rb_["beta"], 1e-5, 0.1, rb_["rm"], rb_["rv"], rb_["sc"], rb_["sh"],
                    rb_["mean"], rb_["inv"]) if with_r else None
    out = Act.empty(N, T, H, W, C_)
    ops.bn_add_relu_raw(ya, qy, ra if with_r else None, qr, out)
    torch.cuda.synchronize()
    tol = dict(rtol=2e-5, atol=2e-6) if fold_f else dict(rtol=0, atol=0)  # folded: fp32 partial sums re-associated
    for k in ("sc", "sh", "mean", "inv", "rm", "rv"):
        torch.testing.assert_close(b[k], a[k], **tol)
        if with_r:
            torch.testing.assert_close(rb_[k], ra_[k], rtol=0, atol=0)
    if fold_f:
        assert kc.relerr(kc.from_act(out), kc.from_act(out_ref)) < 1e-2
    else:
        assert torch.equal(out.buf, out_ref.buf)


@pytest.mark.parametrize("C_,relu,use_mask", [(64, True, False), (256, True, True), (8, False, False), (1280, True, False),
                                              (256, True, "bits"), (32, True, "bits")])
def test_bn_bwd(C_, relu, use_mask):
    N, T, H, W = 2, 3, 6, 6
    y = (torch.randn(N, C_, T, H, W, device=DEV) * 1.5 + 0.3)
    ya = kc.to_act(y)
    yq = kc.from_act(ya).requires_grad_(True)
    gamma = (torch.rand(C_, device=DEV) + 0.5).requires_grad_(True)
    beta = (torch.randn(C_, device=DEV) * 0.2).requires_grad_(True)
    dz = torch.randn(N, C_, T, H, W, device=DEV)
    dza = kc.to_act(dz, ld=C_ + 8, coff=0)
    bn = F.batch_norm(yq, None, None, gamma, beta, True, 0.1, 1e-5)
    extra = torch.randn(N, C_, T, H, W, device=DEV) if use_mask else None
    pre = bn + extra if use_mask else bn
    z = torch.relu(pre) if relu else pre
    z.backward(kc.from_act(dza))
    mean, var, invstd = _bn_ref(yq.detach(), gamma, beta)
    scale = (gamma * invstd).detach()
    shift = (beta - mean * gamma * invstd).detach()
    mask = kc.to_act(z.detach()) if use_mask is True else None
    bits = None
    if use_mask == "bits":  # packed mask as the forward residual tail writes it
        o = z.detach().permute(0, 2, 3, 4, 1).reshape(-1, C_ // 8, 8) > 0
        w = (2 ** torch.arange(8, device=DEV, dtype=torch.int32)).view(1, 1, 8)
        bits = (o.to(torch.int32) * w).sum(-1).to(torch.uint8).contiguous()
    dgamma, dbeta = torch.zeros(C_, device=DEV), torch.zeros(C_, device=DEV)
    ops.bn_bwd_reduce(dza, ya, mask, scale, shift, mean, invstd, relu, dgamma, dbeta, mask_bits=bits)
    dy = Act.empty(N, T, H, W, C_)
    gout = Act.empty(N, T, H, W, C_)
    ops.bn_bwd_apply(dza, ya, mask, scale, shift, mean, invstd, gamma.detach(), dgamma, dbeta, relu, dy, gout, mask_bits=bits)
    torch.cuda.synchronize()
    assert kc.relerr(dgamma, gamma.grad) < 1e-2
    assert kc.relerr(dbeta, beta.grad) < 1e-2
    assert kc.relerr(kc.from_act(dy), yq.grad) < 1.5e-2
    gref = kc.from_act(dza) * ((z.detach() > 0).float() if relu else 1.0)
    assert kc.relerr(kc.from_act(gout), gref) < 1e-2


@pytest.mark.parametrize("C_,H,W", [(64, 16, 16), (8, 15, 13)])
def test_maxpool(C_, H, W):
    N, T = 2, 3
    y = torch.randn(N, C_, T, H, W, device=DEV)
    ya = kc.to_act(y)
    s, b = torch.rand(C_, device=DEV) + 0.5, torch.randn(C_, device=DEV) * 0.2
    s[0] = -s[0]
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    out = Act.empty(N, T, Ho, Wo, C_ + 16, coff=0, C=C_, zero=True)
    idx = torch.empty((N, T, Ho, Wo, C_), dtype=torch.uint8, device=DEV)
    ops.maxpool_fwd(ya, s, b, out, idx)
    z = torch.relu(kc.from_act(ya) * s.view(1, -1, 1, 1, 1) + b.view(1, -1, 1, 1, 1)).requires_grad_(True)
    ref = F.max_pool3d(z, (1, 3, 3), (1, 2, 2), (0, 1, 1))
    assert kc.relerr(kc.from_act(out), ref) < 5e-3
    assert out.buf[..., C_:].abs().max() == 0
    d = torch.randn_like(ref)
    da = kc.to_act(d)
    ref.backward(kc.from_act(da))
    dz = Act.empty(N, T, H, W, C_)
    ops.maxpool_bwd(da, idx, dz)
    got = kc.from_act(dz)
    # ties (bf16-equal maxima / all-zero windows) may pick a different arg-max than torch; compare where z>0
    m = (z.detach() > 0)
    diff = ((got - z.grad).abs() * m).max() / z.grad.abs().max()
    frac_bad = (((got - z.grad).abs() > 1e-2 * z.grad.abs().max()) & m).float().mean()
    assert frac_bad < 2e-3, (diff.item(), frac_bad.item())


@pytest.mark.parametrize("T,H,W,k", [(20, 2, 2, (4, 2, 2)), (8, 7, 7, (8, 7, 7)), (20, 4, 4, (4, 2, 2))])
def test_head_pool(T, H, W, k):
    N, C_ = 3, 64
    x = torch.relu(torch.randn(N, C_, T, H, W, device=DEV))
    xa = kc.to_act(x)
    xq = kc.from_act(xa).requires_grad_(True)
    feat = torch.zeros(N, C_ + 8, device=DEV)
    ops.head_pool_fwd(xa, k, 0.0, 0, feat, 8)
    ref = F.avg_pool3d(xq, k, 1).mean(dim=(2, 3, 4))
    torch.testing.assert_close(feat[:, 8:], ref, rtol=1e-4, atol=1e-5)
    df = torch.randn(N, C_ + 8, device=DEV)
    ref.backward(df[:, 8:])
    dx = Act.empty(N, T, H, W, C_)
    ops.head_pool_bwd(df, 8, k, 0.0, 0, dx)
    assert kc.relerr(kc.from_act(dx), xq.grad) < 6e-3
    # dropout: forward/backward consistency (d feat / d x == backward weights) and keep-rate
    feat_d = torch.full((N, C_ + 8), 7.0, device=DEV)  # stale contents: the kernel owns (and must reset) its columns
    ops.head_pool_fwd(xa, k, 0.5, 1234, feat_d, 8)
    assert bool((feat_d[:, :8] == 7.0).all())
    ones = torch.ones(N, C_ + 8, device=DEV)
    dxd = Act.empty(N, T, H, W, C_)
    ops.head_pool_bwd(ones, 8, k, 0.5, 1234, dxd)
    recon = (kc.from_act(dxd) * xq.detach()).sum(dim=(2, 3, 4))
    assert kc.relerr(recon, feat_d[:, 8:]) < 2e-2
    # the seed read from device memory (the form a CUDA-graph replay needs) draws the same mask as the by-value seed
    sd = torch.tensor([1234], dtype=torch.int64, device=DEV)
    feat_e = torch.zeros(N, C_ + 8, device=DEV)
    ops.head_pool_fwd(xa, k, 0.5, 0, feat_e, 8, seed_dev=sd)
    torch.testing.assert_close(feat_e[:, 8:], feat_d[:, 8:], rtol=1e-5, atol=1e-6)  # atomics order only
    dxe = Act.empty(N, T, H, W, C_)
    ops.head_pool_bwd(ones, 8, k, 0.5, 0, dxe, seed_dev=sd)
    assert torch.equal(kc.from_act(dxe), kc.from_act(dxd))
    sd.fill_(99)
    ops.head_pool_fwd(xa, k, 0.5, 0, feat_e, 8, seed_dev=sd)
    assert float((feat_e[:, 8:] - feat_d[:, 8:]).abs().max()) > 1e-2 * float(feat_d[:, 8:].abs().max())


def test_fc():
    N, K, C_ = 5, 2304, 249
    feat, w, b = torch.randn(N, K, device=DEV), torch.randn(C_, K, device=DEV) * 0.01, torch.randn(C_, device=DEV)
    logits = torch.empty(N, C_, device=DEV)
    ops.fc_fwd(feat, w, b, logits)
    torch.testing.assert_close(logits, feat @ w.t() + b, rtol=1e-4, atol=1e-4)
    dl = torch.randn(N, C_, device=DEV)
    dfeat, dw, db = torch.empty(N, K, device=DEV), torch.zeros(C_, K, device=DEV), torch.zeros(C_, device=DEV)
    ops.fc_bwd(dl, feat, w, dfeat, dw, db)
    torch.testing.assert_close(dfeat, dl @ w, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(dw, dl.t() @ feat, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(db, dl.sum(0), rtol=1e-4, atol=1e-4)


# ------------------------------------------------------------------ conv GEMM: SIMT cross-check first, then tcgen05
IMPLS = [("simt", _ffi.IMPL_SIMT), ("gather", _ffi.IMPL_TC_GATHER), ("tc", _ffi.IMPL_TC)]


@pytest.mark.parametrize("impl_name,impl", IMPLS)
@pytest.mark.parametrize("case", kc.CONV_CASES, ids=[c[0] for c in kc.CONV_CASES])
def test_conv_fprop(case, impl_name, impl):
    r = kc.run_conv_case(case, impl, "plain")
    assert r["err"] < 1e-2, r


@pytest.mark.parametrize("impl_name,impl", IMPLS)
@pytest.mark.parametrize("variant", ["pro", "epi", "stats"])
@pytest.mark.parametrize("case", [kc.CONV_CASES[i] for i in (1, 2, 3, 5, 8, 10, 12, 14, 17, 18, 19)], ids=lambda c: c[0])
def test_conv_fprop_fused(case, variant, impl_name, impl):
    r = kc.run_conv_case(case, impl, variant)
    assert r["err"] < 1e-2, r
    if variant == "stats":
        assert r["err_sum"] < 2e-2 and r["err_sq"] < 1e-2, r
    if variant == "epi":
        assert r["clobber"] == 0.0, r


@pytest.mark.parametrize("impl_name,impl", IMPLS)
@pytest.mark.parametrize("case", [c for c in kc.CONV_CASES if not c[0].startswith("stem")], ids=lambda c: c[0])
def test_conv_dgrad(case, impl_name, impl):
    r = kc.run_dgrad_case(case, impl, accumulate=(case[0] in ("1x1x1_64_256", "1x3x3_128_128_s2", "lateral0_3x1x1_8_16_s4", "fast_1x1x1_32_64_s2")))
    assert r["err"] < 1e-2, r


@pytest.mark.parametrize("impl_name,impl", IMPLS)
@pytest.mark.parametrize("case", kc.CONV_CASES, ids=[c[0] for c in kc.CONV_CASES])
def test_conv_wgrad(case, impl_name, impl):
    name = case[0]
    if impl == _ffi.IMPL_TC_GATHER:
        pytest.skip("one tensor-core wgrad kernel")
    ci_real = {"stem_slow_8_64": 5, "stem_fast_16_8": 15, "stem_slow_16_64": 5}.get(name)
    r = kc.run_wgrad_case(case, impl, pro=name in ("1x3x3_64_64", "3x1x1_256_128", "fast_3x1x1_32_8"), ci_real=ci_real)
    assert r["err"] < 2e-2, r


# ------------------------------------------------------------------ stem (1,7,7)/s(1,2,2) in space-to-depth form
# N, T, H, W, Ci, Co
STEM_CASES = [
    (1, 2, 32, 32, 5, 64),      # slow stem, REF 64^2 crop
    (2, 3, 64, 64, 15, 8),      # fast stem
    (1, 2, 224, 224, 15, 8),    # headline size: 112 output pixels per row
    (1, 1, 128, 128, 5, 64),
    (1, 3, 48, 250, 3, 16),     # widest supported row (125 pixels), CANON channel count
    (3, 1, 20, 36, 2, 24),      # reference's (5, 2) channel variant (new_feature_test.py:715)
]


def _stem_inputs(N, T, H, W, Ci, Co, seed=0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    base = torch.randn((N, T, Ci + 3, H, W), device=DEV, generator=g)
    x = base.permute(0, 2, 1, 3, 4)[:, 2:2 + Ci]  # non-contiguous NCTHW view like train.py:136-140
    w = torch.randn((Co, Ci, 1, 7, 7), device=DEV, generator=g) / (Ci * 49) ** 0.5
    return x, w


@pytest.mark.parametrize("N,T,H,W,Ci,Co", STEM_CASES)
def test_stem_s2d_layout_and_fprop(N, T, H, W, Ci, Co):
    x, w = _stem_inputs(N, T, H, W, Ci, Co)
    assert ops.stem_s2d_eligible(Ci, Co, (1, 7, 7), (1, 2, 2), (0, 3, 3), H, W)
    xs = ops.layout_input_s2d(x)
    cs = xs.shape[-1]
    # layout: channel (sh*2+sw)*Ci + c of s2d pixel (h, w) = x[c, 2h+sh, 2w+sw]
    ref_l = torch.zeros_like(xs, dtype=torch.float32)
    for sh in range(2):
        for sw in range(2):
            q = sh * 2 + sw
            ref_l[..., q * Ci:(q + 1) * Ci] = x[:, :, :, sh::2, sw::2].permute(0, 2, 3, 4, 1).bfloat16().float()
    assert torch.equal(xs.float(), ref_l)
    wp = ops.pack_weight_stem(w.contiguous())
    out = Act.empty(N, T, H // 2, W // 2, Co + 16, coff=8, C=Co, zero=True)
    stats = torch.zeros(2 * Co, device=DEV)
    ops.stem_conv_fprop(xs, Ci, wp, out, stats)
    torch.cuda.synchronize()
    ref = F.conv3d(x.bfloat16().float(), w.bfloat16().float(), stride=(1, 2, 2), padding=(0, 3, 3))
    got = kc.from_act(out)
    assert kc.relerr(got, ref) < 1e-2
    assert float(out.buf[..., :8].abs().max()) == 0.0 and float(out.buf[..., 8 + Co:].abs().max()) == 0.0
    cnt = ref.numel() / Co
    gq = got  # statistics are taken from the stored (bf16-rounded) values
    torch.testing.assert_close(stats[:Co] / cnt, gq.mean(dim=(0, 2, 3, 4)), rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(stats[Co:] / cnt, (gq * gq).mean(dim=(0, 2, 3, 4)), rtol=1e-3, atol=1e-4)


@pytest.mark.parametrize("N,T,H,W,Ci,Co", STEM_CASES)
def test_stem_s2d_wgrad(N, T, H, W, Ci, Co):
    x, w = _stem_inputs(N, T, H, W, Ci, Co, seed=1)
    xs = ops.layout_input_s2d(x)
    g = torch.Generator(device=DEV).manual_seed(5)
    dy = torch.randn((N, Co, T, H // 2, W // 2), device=DEV, generator=g)
    dya = kc.to_act(dy, ld=Co + 8, coff=8)
    wz = torch.zeros_like(w, requires_grad=True)
    F.conv3d(x.bfloat16().float(), wz, stride=(1, 2, 2), padding=(0, 3, 3)).backward(kc.from_act(dya))
    dw = torch.zeros_like(w)
    ops.stem_conv_wgrad(xs, Ci, dya, dw)
    ops.stem_conv_wgrad(xs, Ci, dya, dw)  # accumulates
    torch.cuda.synchronize()
    assert kc.relerr(dw, 2 * wz.grad) < 2e-2


# ------------------------------------------------------------------ pixel folding of narrow-channel layers
# N, T, H, W, Ci, Co, k
FOLD_CASES = [
    (2, 4, 12, 16, 8, 8, (1, 3, 3)),     # fast res2 conv_b
    (2, 6, 10, 12, 32, 8, (3, 1, 1)),    # fast res2 conv_a
    (2, 4, 10, 12, 8, 32, (1, 1, 1)),    # fast res2 conv_c
    (1, 4, 9, 8, 64, 16, (3, 1, 1)),     # fast res3 conv_a
    (2, 3, 7, 14, 32, 32, (1, 3, 3)),    # fast res4 conv_b (f = 2)
    (1, 3, 7, 14, 32, 128, (1, 1, 1)),   # fast res4 conv_c (f = 2)
]


@pytest.mark.parametrize("N,T,H,W,Ci,Co,k", FOLD_CASES)
def test_folded_conv_matches_reference(N, T, H, W, Ci, Co, k):
    """fprop (+ BatchNorm statistics), dgrad and wgrad run on the folded view of the same buffers"""
    g = torch.Generator(device=DEV).manual_seed(3)
    ldi, ldo = kc.ops.round_up(Ci, 16), kc.ops.round_up(Co, 16)
    s1, p = (1, 1, 1), tuple(kk // 2 for kk in k)
    f = ops.fold_factor(ldi, ldo, k, s1, W)
    assert f >= 2
    x = torch.randn((N, Ci, T, H, W), device=DEV, generator=g)
    w = torch.randn((Co, Ci) + k, device=DEV, generator=g) / (Ci * k[0] * k[1] * k[2]) ** 0.5
    xa = Act.empty(N, T, H, W, ldi, C=Ci, zero=True)
    xa.torch_view().copy_(x.permute(0, 2, 3, 4, 1))
    xq, wq = kc.from_act(xa), w.bfloat16().float()
    kf = (k[0], k[1], 1 if k[2] == 1 else 3)
    pf = tuple(kk // 2 for kk in kf)
    wf = ops.fold_weight(w.contiguous(), f, ldo, ldi)
    # ---- fprop + stats
    out = Act.empty(N, T, H, W, ldo, C=Co, zero=True)
    sf = torch.zeros(2 * f * ldo, device=DEV)
    ops.conv_gemm(ops.folded(xa, f), ops.pack_weight(wf, f * ldi), ops.folded(out, f), kf, s1, pf, stats=sf)
    stats = torch.zeros(2 * Co, device=DEV)
    ops.unfold_stats(sf, f, ldo, Co, stats)
    ref = F.conv3d(xq, wq, padding=p)
    got = kc.from_act(out)
    assert kc.relerr(got, ref) < 1e-2
    assert float(out.buf[..., Co:].abs().max()) == 0.0 if ldo > Co else True
    cnt = ref.numel() / Co
    torch.testing.assert_close(stats[:Co] / cnt, got.mean(dim=(0, 2, 3, 4)), rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(stats[Co:] / cnt, (got * got).mean(dim=(0, 2, 3, 4)), rtol=1e-3, atol=1e-4)
    # ---- dgrad (accumulating on top of a base gradient)
    dy = torch.randn((N, Co, T, H, W), device=DEV, generator=g)
    dya = Act.empty(N, T, H, W, ldo, C=Co, zero=True)
    dya.torch_view().copy_(dy.permute(0, 2, 3, 4, 1))
    dyq = kc.from_act(dya)
    xz = torch.zeros_like(xq, requires_grad=True)
    wz = torch.zeros_like(wq, requires_grad=True)
    F.conv3d(xz, wq, padding=p).backward(dyq)
    F.conv3d(xq, wz, padding=p).backward(dyq)
    base = torch.randn((N, Ci, T, H, W), device=DEV, generator=g)
    dxa = Act.empty(N, T, H, W, ldi, C=Ci, zero=True)
    dxa.torch_view().copy_(base.permute(0, 2, 3, 4, 1))
    ref_dx = xz.grad + kc.from_act(dxa)
    dxf = ops.folded(dxa, f)
    ops.conv_gemm(ops.folded(dya, f), ops.pack_weight_dgrad(wf, f * ldo, s1), dxf, kf, s1, pf, transposed=True, residual=dxf)
    torch.cuda.synchronize()
    assert kc.relerr(kc.from_act(dxa), ref_dx) < 1.5e-2
    assert float(dxa.buf[..., Ci:].abs().max()) == 0.0 if ldi > Ci else True
    # ---- wgrad through the folded gradient
    dwf = torch.zeros_like(wf)
    ops.conv_wgrad(ops.folded(xa, f), ops.folded(dya, f), dwf, kf, s1, pf)
    dw = torch.zeros_like(w)
    ops.unfold_wgrad(dwf, f, ldo, ldi, dw)
    torch.cuda.synchronize()
    assert kc.relerr(dw, wz.grad) < 2e-2
    # ---- wgrad with the band map applied in the kernel's epilogue (what the engine uses): same sums, on top of a base
    dw2 = torch.full_like(w, 0.25)
    ops.conv_wgrad_folded(ops.folded(xa, f), ops.folded(dya, f), dw2, kf, pf, f, ldo, ldi)
    torch.cuda.synchronize()
    assert kc.relerr(dw2 - 0.25, wz.grad) < 2e-2
    torch.testing.assert_close(dw2 - 0.25, dw, rtol=1e-3, atol=1e-3 * float(dw.abs().max()))


def test_bad_args_rejected():
    x = Act.empty(1, 2, 4, 4, 12)  # 12 channels: not a multiple of 8
    w = torch.empty(64, 64, dtype=torch.bfloat16, device=DEV)
    out = Act.empty(1, 2, 4, 4, 64)
    with pytest.raises(_ffi.SfbError):
        ops.conv_gemm(x, w, out, (1, 1, 1), (1, 1, 1), (0, 0, 0))


# ---- raw clip staging (sfb_stage_clips_u8) ---------------------------------------------------------
def _stage_expected(frames, crop, pathways, pad):
    from oracle import staging_oracle as st
    windows = None if crop is None else [(int(dy) + pad, int(dx) + pad) for dy, dx in crop.tolist()]
    x = st.stage_clips(frames.cpu(), windows, pad)
    exp = []
    for c0, Cc, ts in pathways:
        xp = torch.permute(x, [0, 2, 1, 3, 4])[:, c0:c0 + Cc, ::ts].contiguous()
        exp.append(st.space_to_depth_bf16(xp, ops.stem_s2d_channels(Cc)))
    return exp


@pytest.mark.parametrize("N,T,H,W,Ct,pathways,crop", [
    (2, 4, 16, 16, 21, ((0, 5, 1), (5, 15, 1)), False),      # reference graph: equal clip lengths
    (2, 8, 20, 36, 21, ((0, 5, 4), (5, 15, 1)), False),      # 8x8 SlowFast: Slow takes every 4th frame
    (3, 2, 20, 20, 21, ((0, 5, 1), (5, 15, 1)), True),       # RandomCrop windows, rows 420 B (not 16-B multiples)
    (2, 2, 18, 22, 20, ((0, 5, 2), (5, 15, 1)), True),
    (1, 2, 64, 250, 21, ((0, 5, 1), (5, 15, 1)), True),      # widest s2d row (W/2 = 125)
    (2, 3, 12, 14, 7, ((0, 3, 1), (3, 4, 3)), True),         # other channel splits
])
def test_stage_clips_u8_bit_identical_to_oracle(N, T, H, W, Ct, pathways, crop):
    g = torch.Generator().manual_seed(N * 1000 + W)
    frames = torch.randint(0, 256, (N, T, H, W, Ct), generator=g, dtype=torch.uint8)
    pad = max(H, W) // 10 + 3
    cyx = None
    if crop:
        cyx = torch.randint(-pad, pad + 1, (N, 2), generator=g, dtype=torch.int32)
        cyx[0, 0], cyx[0, 1] = -pad, pad  # extremes
    exp = _stage_expected(frames, cyx, pathways, pad)
    got = ops.stage_clips_u8(frames.cuda(), pathways, None if cyx is None else cyx.cuda())
    torch.cuda.synchronize()
    for e, o in zip(exp, got):
        assert o.shape == e.shape
        assert torch.equal(o.cpu().view(torch.int16), e.view(torch.int16))


def test_stage_clips_u8_against_torchvision_golden_and_fp32_layout():
    import os
    g = torch.load(os.path.join(os.path.dirname(__file__), "golden", "staging_small.pt"))
    pad = g["padding"]
    cyx = torch.tensor([(i - pad, j - pad) for i, j in g["windows"]], dtype=torch.int32)
    pathways = ((0, 5, 1), (5, 15, 1))
    got = ops.stage_clips_u8(g["frames"].cuda(), pathways, cyx.cuda())
    # the fp32 batch torchvision produced, through the fp32 entry point the strict drop-in path uses
    x = torch.permute(g["aug"].cuda(), [0, 2, 1, 3, 4])
    for (c0, Cc, _), o in zip(pathways, got):
        ref = ops.layout_input_s2d(x[:, c0:c0 + Cc])
        assert torch.equal(o.view(torch.int16), ref.view(torch.int16))


def test_stage_clips_u8_rejects_bad_arguments():
    f = torch.zeros((1, 2, 8, 8, 21), dtype=torch.uint8, device="cuda")
    with pytest.raises(_ffi.SfbError):
        ops.stage_clips_u8(f, ((0, 5, 3), (5, 15, 1)))            # T % t_stride
    with pytest.raises(_ffi.SfbError):
        ops.stage_clips_u8(f, ((0, 5, 1), (10, 15, 1)))           # channel range beyond the frame
    with pytest.raises(_ffi.SfbError):
        ops.stage_clips_u8(torch.zeros((1, 2, 7, 8, 21), dtype=torch.uint8, device="cuda"), ((0, 5, 1),))  # odd H


# ---- evaluation drivers on the device (sfb_video_scores / sfb_sparse_fuse, SURVEY.md 8(f) f4 + f2) ---------------------
@pytest.mark.parametrize("C,sv", [(249, [7, 5, 10, 1, 3]), (10, [1] * 17), (1000, [4, 4]), (249, [64])])
def test_video_scores_match_run_eval_tail(C, sv):
    from oracle import eval_oracle as eo
    from video_classification_b200 import evaluation as ev
    g = torch.Generator().manual_seed(C + len(sv))
    n = sum(sv)
    logits = torch.randn((n, C), generator=g) * 3.0
    labels = torch.cat([torch.full((k,), int(torch.randint(0, C, (1,), generator=g))) for k in sv])
    # make some videos correct on purpose
    for v, (a, k) in enumerate(zip([0] + list(torch.tensor(sv).cumsum(0)[:-1]), sv)):
        if v % 2 == 0:
            logits[a:a + k, labels[a]] += 6.0
    ref = eo.run_eval_tail(logits.numpy(), labels.numpy(), sv)
    acc = ev.EvalAccumulator()
    half = n // 2
    acc.add(logits[:half].cuda(), labels[:half].cuda(), sv[:1])        # batches and videos arrive in pieces
    acc.add(logits[half:].cuda(), labels[half:].cuda(), sv[1:])
    got = acc.result()
    assert got["sv"] == ref["sv"] and (got["t"] == ref["t"]).all()
    assert abs(got["ps"] - ref["ps"]).max() <= 1e-6
    assert abs(got["video_mean"] - ref["video_mean"]).max() <= 1e-6
    assert (got["video_top1"] == ref["video_top1"]).all()
    assert got["acc"] == pytest.approx(float(ref["acc"]))


def test_video_scores_ties_and_errors():
    from video_classification_b200 import evaluation as ev
    s = torch.zeros((3, 300), device="cuda")
    s[:, 17] = 2.0
    s[:, 290] = 2.0   # equal maxima: np.argmax takes the first
    r = ev.video_scores(s, [2, 1])
    assert r["top1"].tolist() == [17, 17]
    with pytest.raises(ValueError):
        ev.video_scores(s, [2, 2])
    with pytest.raises(_ffi.SfbError):
        ev.video_scores(s.cpu(), [2, 1])


@pytest.mark.parametrize("N,P,C", [(33, 5, 249), (4, 6, 10), (300, 1, 7)])
def test_sparse_model_matches_reference_forward_and_gradients(N, P, C):
    from oracle import eval_oracle as eo
    from video_classification_b200 import evaluation as ev
    torch.manual_seed(N + P)
    ref = eo.SparseModelRef(C, P)
    mine = ev.SparseModel(C, P)
    assert list(mine.state_dict().keys()) == list(ref.state_dict().keys())
    mine.load_state_dict(ref.state_dict(), strict=True)
    mine.cuda()
    x = torch.softmax(torch.randn(N, P, C), -1)
    t = torch.randint(0, C, (N,))
    xr = x.clone().requires_grad_(True)
    yr = ref(xr)
    F.cross_entropy(yr, t).backward()
    xm = x.clone().cuda().requires_grad_(True)
    ym = mine(xm)
    F.cross_entropy(ym, t.cuda()).backward()
    assert torch.allclose(ym.detach().cpu(), yr.detach(), rtol=1e-5, atol=1e-6)
    assert torch.allclose(xm.grad.cpu(), xr.grad, rtol=1e-5, atol=1e-7)
    for (k, pm), (_, pr) in zip(mine.named_parameters(), ref.named_parameters()):
        assert torch.allclose(pm.grad.cpu(), pr.grad, rtol=1e-4, atol=1e-7), k


def test_fuse_ensemble_matches_sparse_test_tail():
    from oracle import eval_oracle as eo
    from video_classification_b200 import evaluation as ev
    torch.manual_seed(5)
    P, C, sv = 5, 249, [6, 9, 4, 11]
    n = sum(sv)
    ref_m = eo.SparseModelRef(C, P)
    mine = ev.SparseModel(C, P)
    mine.load_state_dict(ref_m.state_dict())
    mine.cuda()
    ps = torch.softmax(torch.randn(n, P, C) * 2, -1)
    labels = torch.cat([torch.full((k,), i * 7) for i, k in enumerate(sv)])
    with torch.no_grad():
        fused_ref = ref_m(ps).numpy()
    ref = eo.sparse_test_tail(fused_ref, labels.numpy(), sv)
    got = ev.fuse_ensemble(mine, ps.cuda(), labels.cuda(), sv)
    assert abs(got["video_mean"] - ref["video_mean"]).max() <= 1e-5
    assert (got["video_top1"] == ref["video_top1"]).all()
    assert got["acc"] == pytest.approx(float(ref["acc"]))


# ---- input gradient with the BatchNorm(+ReLU) backward reduction fused into its epilogue (sfb_conv_dgrad_bn) ----------
DGRAD_BN_CASES = [
    # name, N, T, H, W, Ci (= channels of dz / y), Co (= gathered dy channels), k, s
    ("c_1x1x1_64_256", 2, 4, 12, 12, 64, 256, (1, 1, 1), (1, 1, 1)),          # dgrad of conv_c -> dz of conv_b (BN tile 64)
    ("b_1x3x3_64_64", 2, 3, 14, 14, 64, 64, (1, 3, 3), (1, 1, 1)),            # dgrad of conv_b -> dz of conv_a
    ("b_1x3x3_128_128_s2", 1, 4, 14, 14, 128, 128, (1, 3, 3), (1, 2, 2)),     # strided conv_b: one launch per parity class
    ("c_1x1x1_256_1024", 1, 8, 7, 7, 256, 1024, (1, 1, 1), (1, 1, 1)),        # 256-wide tile, weights streamed (3 stages)
    ("b_1x3x3_512_512", 1, 4, 7, 7, 512, 512, (1, 3, 3), (1, 1, 1)),          # two N tiles
    ("big_c_1x1x1_128_512", 2, 8, 28, 28, 128, 512, (1, 1, 1), (1, 1, 1)),    # several persistent waves
    ("odd_b_1x3x3_32_32", 2, 5, 13, 11, 32, 32, (1, 3, 3), (1, 1, 1)),        # partial patches, 32-channel boxes
    ("fast_c_1x1x1_8_32", 2, 8, 12, 12, 8, 32, (1, 1, 1), (1, 1, 1)),         # 8 real channels in a 16-channel buffer
    ("fast_b_1x3x3_16_16_s2", 2, 8, 14, 14, 16, 16, (1, 3, 3), (1, 2, 2)),
    ("halo_b_1x3x3_64_64_w7", 2, 3, 28, 7, 64, 64, (1, 3, 3), (1, 1, 1)),     # halo-mode patches (one box per tile)
    ("halo_b_1x3x3_64_64_big", 2, 2, 56, 56, 64, 64, (1, 3, 3), (1, 1, 1)),
]


@pytest.mark.parametrize("case", DGRAD_BN_CASES, ids=[c[0] for c in DGRAD_BN_CASES])
def test_conv_dgrad_bn_fused_reduction(case):
    """out = dgrad * relu-mask(y), dbeta = sum out, dgamma = invstd * sum out*(y - mean), against PyTorch fp32 on the
    same bf16 operands; pad lanes of the output stay zero and the accumulators of neighbouring layers (the arrays sit
    in one arena) are not touched"""
    name, N, T, H, W, Ci, Co, k, s = case
    g = torch.Generator(device=DEV).manual_seed(11)
    p = tuple(kk // 2 for kk in k)
    To, Ho, Wo = ops.conv_out_dims((T, H, W), k, s, p)
    ld = kc.ops.round_up(Ci, 16)
    dy = torch.randn((N, Co, To, Ho, Wo), device=DEV, generator=g)
    w = torch.randn((Co, Ci) + k, device=DEV, generator=g) / (Co * k[0] * k[1] * k[2]) ** 0.5
    y = torch.randn((N, Ci, T, H, W), device=DEV, generator=g) * 1.3 + 0.2
    dya = kc.to_act(dy)
    ya = Act.empty(N, T, H, W, ld, C=Ci, zero=True)
    ya.torch_view().copy_(y.permute(0, 2, 3, 4, 1))
    yq, wq = kc.from_act(ya), w.bfloat16().float()
    x = torch.zeros((N, Ci, T, H, W), device=DEV, requires_grad=True)
    F.conv3d(x, wq, stride=s, padding=p).backward(kc.from_act(dya))
    dz = x.grad.bfloat16().float()                       # the epilogue masks the bf16-rounded gradient
    gamma = torch.rand(Ci, device=DEV, generator=g) + 0.5
    beta = torch.randn(Ci, device=DEV, generator=g) * 0.3
    M = N * T * H * W
    yf = yq.permute(0, 2, 3, 4, 1).reshape(M, Ci)
    mean, var = yf.mean(0), yf.var(0, unbiased=False)
    invstd = (var + 1e-5).rsqrt()
    scale, shift = gamma * invstd, beta - mean * gamma * invstd
    bc = lambda t: t.view(1, -1, 1, 1, 1)
    mask = (yq * bc(scale) + bc(shift)) > 0
    ref = torch.where(mask, dz, torch.zeros_like(dz))
    ref_dbeta = ref.sum(dim=(0, 2, 3, 4))
    ref_dgamma = (ref * (yq - bc(mean))).sum(dim=(0, 2, 3, 4)) * invstd
    # accumulators in the middle of an arena with sentinels around them
    arena = torch.full((Ci + 32,) , 7.0, device=DEV)
    arena2 = torch.full((Ci + 32,), 7.0, device=DEV)
    dgamma, dbeta = arena[16:16 + Ci], arena2[16:16 + Ci]
    dgamma.zero_(); dbeta.zero_()
    wpt = ops.pack_weight_dgrad(w.contiguous(), Co, s)
    out = Act.empty(N, T, H, W, ld, C=Ci, zero=True)
    outp = Act(out.buf, 0, ld)  # a dgrad writes the padded width
    ops.conv_dgrad_bn(dya, wpt, outp, k, s, p, Act(ya.buf, 0, ld), scale, shift, mean, invstd, dgamma, dbeta, C_real=Ci)
    torch.cuda.synchronize()
    got = kc.from_act(out)
    # the kernel evaluates the mask with one fma, PyTorch with two roundings: where y*scale+shift is within rounding of
    # zero either sign is right, those (measure-zero) elements are left out of the elementwise comparison
    amb = (yq * bc(scale) + bc(shift)).abs() < 1e-5
    assert kc.relerr(got.masked_fill(amb, 0.0), ref.masked_fill(amb, 0.0)) < 1e-2, name
    if ld > Ci:
        assert float(out.buf[..., Ci:].abs().max()) == 0.0
    assert kc.relerr(dbeta, ref_dbeta) < 2e-3 + 1e-2 * float(ref.abs().max()) / max(float(ref_dbeta.abs().max()), 1e-6), name
    assert kc.relerr(dgamma, ref_dgamma) < 1e-2, name
    assert bool((arena[:16] == 7.0).all() and (arena[16 + Ci:] == 7.0).all())
    assert bool((arena2[:16] == 7.0).all() and (arena2[16 + Ci:] == 7.0).all())
    # the same result as the two-kernel path: plain dgrad, then sfb_bn_bwd_reduce with the recomputed mask
    out2 = Act.empty(N, T, H, W, ld, C=Ci, zero=True)
    ops.conv_gemm(dya, wpt, Act(out2.buf, 0, ld), k, s, p, transposed=True)
    dg2, db2 = torch.zeros(Ci, device=DEV), torch.zeros(Ci, device=DEV)
    ops.bn_bwd_reduce(out2, ya, None, scale, shift, mean, invstd, True, dg2, db2)
    torch.cuda.synchronize()
    torch.testing.assert_close(dbeta, db2, rtol=2e-3, atol=2e-3 * float(db2.abs().max()))
    torch.testing.assert_close(dgamma, dg2, rtol=2e-3, atol=2e-3 * float(dg2.abs().max()))


@pytest.mark.parametrize("N,T,H,W,Ci,Co,k", [FOLD_CASES[0], FOLD_CASES[2], FOLD_CASES[3]])
def test_conv_dgrad_bn_fused_reduction_pixel_folded(N, T, H, W, Ci, Co, k):
    """the fused reduction on a pixel-folded launch: column j of the folded output is channel j % ld"""
    g = torch.Generator(device=DEV).manual_seed(5)
    ldi, ldo = kc.ops.round_up(Ci, 16), kc.ops.round_up(Co, 16)
    s1, p = (1, 1, 1), tuple(kk // 2 for kk in k)
    f = ops.fold_factor(ldi, ldo, k, s1, W)
    assert f >= 2
    w = torch.randn((Co, Ci) + k, device=DEV, generator=g) / (Co * k[0] * k[1] * k[2]) ** 0.5
    kf = (k[0], k[1], 1 if k[2] == 1 else 3)
    pf = tuple(kk // 2 for kk in kf)
    wf = ops.fold_weight(w.contiguous(), f, ldo, ldi)
    dy = torch.randn((N, Co, T, H, W), device=DEV, generator=g)
    dya = Act.empty(N, T, H, W, ldo, C=Co, zero=True)
    dya.torch_view().copy_(dy.permute(0, 2, 3, 4, 1))
    y = torch.randn((N, Ci, T, H, W), device=DEV, generator=g) * 1.3 + 0.2
    ya = Act.empty(N, T, H, W, ldi, C=Ci, zero=True)
    ya.torch_view().copy_(y.permute(0, 2, 3, 4, 1))
    yq, wq = kc.from_act(ya), w.bfloat16().float()
    x = torch.zeros((N, Ci, T, H, W), device=DEV, requires_grad=True)
    F.conv3d(x, wq, padding=p).backward(kc.from_act(dya))
    dz = x.grad.bfloat16().float()
    gamma = torch.rand(Ci, device=DEV, generator=g) + 0.5
    beta = torch.randn(Ci, device=DEV, generator=g) * 0.3
    yf = yq.permute(0, 2, 3, 4, 1).reshape(-1, Ci)
    mean, var = yf.mean(0), yf.var(0, unbiased=False)
    invstd = (var + 1e-5).rsqrt()
    scale, shift = gamma * invstd, beta - mean * gamma * invstd
    bc = lambda t: t.view(1, -1, 1, 1, 1)
    ref = torch.where((yq * bc(scale) + bc(shift)) > 0, dz, torch.zeros_like(dz))
    dgamma, dbeta = torch.zeros(Ci, device=DEV), torch.zeros(Ci, device=DEV)
    out = Act.empty(N, T, H, W, ldi, C=Ci, zero=True)
    ops.conv_dgrad_bn(ops.folded(dya, f), ops.pack_weight_dgrad(wf, f * ldo, s1), ops.folded(out, f), kf, s1, pf,
                      ops.folded(ya, f), scale, shift, mean, invstd, dgamma, dbeta, C_real=Ci, fold_ld=ldi)
    torch.cuda.synchronize()
    amb = (yq * bc(scale) + bc(shift)).abs() < 1e-5
    assert kc.relerr(kc.from_act(out).masked_fill(amb, 0.0), ref.masked_fill(amb, 0.0)) < 1.5e-2
    if ldi > Ci:
        assert float(out.buf[..., Ci:].abs().max()) == 0.0
    assert kc.relerr(dbeta, ref.sum(dim=(0, 2, 3, 4))) < 2e-2
    assert kc.relerr(dgamma, (ref * (yq - bc(mean))).sum(dim=(0, 2, 3, 4)) * invstd) < 1e-2


@pytest.mark.parametrize("C_", [32, 256, 1024])
def test_bn_bwd_tail2_equals_two_separate_passes(C_):
    """projection-shortcut tail: one reduce + one apply for both BatchNorms == sfb_bn_bwd_reduce / _apply run per layer"""
    N, T, H, W = 2, 3, 6, 7
    M = N * T * H * W
    g = torch.Generator(device=DEV).manual_seed(4)
    dz = kc.to_act(torch.randn((N, C_, T, H, W), device=DEV, generator=g), ld=C_ + 8, coff=8)
    ys = [kc.to_act(torch.randn((N, C_, T, H, W), device=DEV, generator=g) * (1.0 + i) + 0.3 * i) for i in range(2)]
    bits = torch.randint(0, 256, (M, C_ // 8), device=DEV, generator=g, dtype=torch.uint8)
    ref, got = [], []
    sides = []
    for y in ys:
        yf = kc.from_act(y).permute(0, 2, 3, 4, 1).reshape(M, C_)
        mean, invstd = yf.mean(0), (yf.var(0, unbiased=False) + 1e-5).rsqrt()
        gamma = torch.rand(C_, device=DEV, generator=g) + 0.5
        dg, db = torch.zeros(C_, device=DEV), torch.zeros(C_, device=DEV)
        ops.bn_bwd_reduce(dz, y, None, None, None, mean, invstd, True, dg, db, mask_bits=bits)
        dy = Act.empty(N, T, H, W, C_)
        ops.bn_bwd_apply(dz, y, None, None, None, mean, invstd, gamma, dg, db, True, dy, mask_bits=bits)
        ref.append((dg, db, dy))
        dg2, db2 = torch.zeros(C_, device=DEV), torch.zeros(C_, device=DEV)
        dy2 = Act.empty(N, T, H, W, C_)
        got.append((dg2, db2, dy2))
        sides.append((y, mean, invstd, gamma, dg2, db2, dy2))
    ops.bn_bwd_tail2(dz, bits, sides)
    torch.cuda.synchronize()
    for (dg, db, dy), (dg2, db2, dy2) in zip(ref, got):
        torch.testing.assert_close(dg2, dg, rtol=1e-4, atol=1e-4 * float(dg.abs().max()))
        torch.testing.assert_close(db2, db, rtol=1e-4, atol=1e-4 * float(db.abs().max()))
        assert kc.relerr(kc.from_act(dy2), kc.from_act(dy)) < 1e-2


@pytest.mark.parametrize("name", ["1x3x3_64_64", "big_1x3x3_64_64", "halo_1x3x3_64_64_w7", "halo_1x3x3_64_64_w15"])
def test_halo_mode_equals_tap_by_tap_boxes(name, monkeypatch):
    """(1,3,3) 64-channel convs fetch ONE box per tile (patch + border) and read the nine taps through shifted
    descriptors; SFB_HALO=0 fetches one box per tap.  Same MMAs in the same order: the results are identical."""
    case = [c for c in kc.CONV_CASES if c[0] == name][0]
    _, N, T, H, W, Ci, Co, k, s = case
    g = torch.Generator(device=DEV).manual_seed(3)
    x = kc.to_act(torch.randn((N, Ci, T, H, W), device=DEV, generator=g))
    w = torch.randn((Co, Ci) + k, device=DEV, generator=g) / (Ci * 9) ** 0.5
    dy = kc.to_act(torch.randn((N, Co, T, H, W), device=DEV, generator=g))
    wp = ops.pack_weight(w.contiguous(), x.C)
    wpt = ops.pack_weight_dgrad(w.contiguous(), Co, s)
    outs = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("SFB_HALO", flag)
        y = Act.empty(N, T, H, W, Co, zero=True)
        stats = torch.zeros(2 * Co, device=DEV)
        ops.conv_gemm(x, wp, y, k, s, (0, 1, 1), stats=stats)
        dx = Act.empty(N, T, H, W, Ci, zero=True)
        ops.conv_gemm(dy, wpt, dx, k, s, (0, 1, 1), transposed=True)
        torch.cuda.synchronize()
        outs[flag] = (y.buf.clone(), stats.clone(), dx.buf.clone())
    assert torch.equal(outs["0"][0], outs["1"][0])
    assert torch.equal(outs["0"][2], outs["1"][2])
    torch.testing.assert_close(outs["1"][1], outs["0"][1], rtol=1e-5, atol=1e-3)   # fp32 atomics in another order
    assert float(outs["1"][0].float().abs().max()) > 0


@pytest.mark.parametrize("name", ["3x1x1_1024_256", "big_3x1x1_640_256"])
def test_cta_pair_mma_equals_single_cta(name, monkeypatch):
    """The 256-wide streamed-weight convs run as CTA pairs (cta_group::2: one 256-row MMA over two neighbouring patches,
    each CTA staging its own pixels and half of the weight tile); SFB_PAIR=0 runs them one CTA per tile.  Same MMAs on
    the same operands in the same K order: identical results; an odd number of patches leaves the last pair half empty."""
    case = [c for c in kc.CONV_CASES if c[0] == name][0]
    _, N, T, H, W, Ci, Co, k, s = case
    g = torch.Generator(device=DEV).manual_seed(5)
    p = tuple(kk // 2 for kk in k)
    for n_clips in (N, 1):                                      # 1 clip of 8x7x7: 4 patches (3 with T = 5 below)
        Tn = T if n_clips == N else 5
        x = kc.to_act(torch.randn((n_clips, Ci, Tn, H, W), device=DEV, generator=g))
        w = torch.randn((Co, Ci) + k, device=DEV, generator=g) / (Ci * k[0] * k[1] * k[2]) ** 0.5
        dy = kc.to_act(torch.randn((n_clips, Co, Tn, H, W), device=DEV, generator=g))
        wp = ops.pack_weight(w.contiguous(), x.C)
        wpt = ops.pack_weight_dgrad(w.contiguous(), Co, s)
        outs = {}
        for flag in ("0", "1"):
            monkeypatch.setenv("SFB_PAIR", flag)
            y = Act.empty(n_clips, Tn, H, W, Co, zero=True)
            stats = torch.zeros(2 * Co, device=DEV)
            ops.conv_gemm(x, wp, y, k, s, p, stats=stats)
            dx = Act.empty(n_clips, Tn, H, W, Ci, zero=True)
            ops.conv_gemm(dy, wpt, dx, k, s, p, transposed=True)
            torch.cuda.synchronize()
            outs[flag] = (y.buf.clone(), stats.clone(), dx.buf.clone())
        assert torch.equal(outs["0"][0], outs["1"][0])
        assert torch.equal(outs["0"][2], outs["1"][2])
        torch.testing.assert_close(outs["1"][1], outs["0"][1], rtol=1e-5, atol=1e-3)
        ref = F.conv3d(kc.from_act(x), w.bfloat16().float(), stride=s, padding=p)
        assert kc.relerr(kc.from_act(Act(outs["1"][0], 0, Co)), ref) < 2e-2


@pytest.mark.parametrize("name", ["3x1x1_256_128", "big_3x1x1_640_256", "lateral_3x1x1_32_64_s4"])
def test_merged_hw_axis_equals_separate_axes(name, monkeypatch):
    """Temporal convs ((3,1,1), laterals) tile a frame as ONE axis of H*W pixels (patches may cut it anywhere);
    SFB_MERGE_HW=0 keeps H and W as separate patch axes.  Only the tiling differs: identical results."""
    case = [c for c in kc.CONV_CASES if c[0] == name][0]
    _, N, T, H, W, Ci, Co, k, s = case
    g = torch.Generator(device=DEV).manual_seed(6)
    p = tuple(kk // 2 for kk in k)
    x = kc.to_act(torch.randn((N, Ci, T, H, W), device=DEV, generator=g))
    w = torch.randn((Co, Ci) + k, device=DEV, generator=g) / (Ci * k[0]) ** 0.5
    To, Ho, Wo = ops.conv_out_dims((T, H, W), k, s, p)
    dy = kc.to_act(torch.randn((N, Co, To, Ho, Wo), device=DEV, generator=g))
    wp = ops.pack_weight(w.contiguous(), x.C)
    wpt = ops.pack_weight_dgrad(w.contiguous(), dy.C, s)
    outs = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("SFB_MERGE_HW", flag)
        y = Act.empty(N, To, Ho, Wo, Co, zero=True)
        stats = torch.zeros(2 * Co, device=DEV)
        ops.conv_gemm(x, wp, y, k, s, p, stats=stats)
        dx = Act.empty(N, T, H, W, x.C, zero=True)
        ops.conv_gemm(dy, wpt, dx, k, s, p, transposed=True)
        torch.cuda.synchronize()
        outs[flag] = (y.buf.clone(), stats.clone(), dx.buf.clone())
    assert torch.equal(outs["0"][0], outs["1"][0])
    assert torch.equal(outs["0"][2], outs["1"][2])
    torch.testing.assert_close(outs["1"][1], outs["0"][1], rtol=1e-5, atol=1e-3)
    ref = F.conv3d(kc.from_act(x)[:, :Ci], w.bfloat16().float(), stride=s, padding=p)
    assert kc.relerr(kc.from_act(Act(outs["1"][0], 0, Co)), ref) < 2e-2
