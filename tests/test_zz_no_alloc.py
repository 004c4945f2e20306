"""sfb_bn_bwd_noalloc (include/sfb200.h): the BatchNorm(+ReLU) backward whose per-channel sums live in the library's
file-scope `static __device__` array instead of caller memory -- no counterpart in the reference (added for callers that
cannot allocate).  Checked through the C ABI against the existing pair sfb_bn_bwd_reduce + sfb_bn_bwd_apply on the same
inputs, on two calls in a row (the array keeps the first call's sums: the second call, with another channel count and
other data, must not see them), with the tolerances of tests/test_kernels_gpu.py's BatchNorm-backward comparisons.

Collected last on purpose (file name).
"""
import pytest
import torch

import kernel_cases as kc
from video_classification_b200 import _ffi, ops
from video_classification_b200.ops import Act

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _case(C_, relu, seed, dims=(2, 3, 6, 7)):
    N, T, H, W = dims
    M = N * T * H * W
    g = torch.Generator(device=DEV).manual_seed(seed)
    dz = kc.to_act(torch.randn((N, C_, T, H, W), device=DEV, generator=g), ld=C_ + 8, coff=8)
    y = kc.to_act(torch.randn((N, C_, T, H, W), device=DEV, generator=g) * 1.5 + 0.3)
    yf = kc.from_act(y).permute(0, 2, 3, 4, 1).reshape(M, C_)
    mean, invstd = yf.mean(0), (yf.var(0, unbiased=False) + 1e-5).rsqrt()
    gamma = torch.rand(C_, device=DEV, generator=g) + 0.5
    beta = torch.randn(C_, device=DEV, generator=g) * 0.2
    scale = (gamma * invstd).contiguous()
    shift = (beta - mean * scale).contiguous()
    return dict(dims=dims, C=C_, relu=relu, dz=dz, y=y, mean=mean, invstd=invstd, gamma=gamma, scale=scale, shift=shift)


def _existing(c):
    dg, db = torch.zeros(c["C"], device=DEV), torch.zeros(c["C"], device=DEV)
    ops.bn_bwd_reduce(c["dz"], c["y"], None, c["scale"], c["shift"], c["mean"], c["invstd"], c["relu"], dg, db)
    dy = Act.empty(*c["dims"], c["C"])
    ops.bn_bwd_apply(c["dz"], c["y"], None, c["scale"], c["shift"], c["mean"], c["invstd"], c["gamma"], dg, db,
                     c["relu"], dy)
    return dg, db, dy


def _compare(c, dg2, db2, dy2):
    dg, db, dy = _existing(c)
    torch.cuda.synchronize()
    torch.testing.assert_close(dg2, dg, rtol=1e-4, atol=1e-4 * float(dg.abs().max()))
    torch.testing.assert_close(db2, db, rtol=1e-4, atol=1e-4 * float(db.abs().max()))
    assert kc.relerr(kc.from_act(dy2), kc.from_act(dy)) < 1e-2
    assert float(kc.from_act(dy2).abs().max()) > 0


def test_bn_bwd_noalloc_matches_reduce_plus_apply_on_two_calls_in_a_row():
    first = _case(256, True, seed=11)
    second = _case(64, False, seed=12, dims=(1, 2, 5, 9))     # fewer channels, other data: stale sums would show
    outs = []
    for c in (first, second):
        # outputs pre-filled with garbage: dgamma_out / dbeta_out are overwritten, not accumulated
        dg2 = torch.full((c["C"],), 7.0, device=DEV)
        db2 = torch.full((c["C"],), -3.0, device=DEV)
        dy2 = Act.empty(*c["dims"], c["C"])
        outs.append((dg2, db2, dy2))
    torch.cuda.synchronize()
    n0 = _ffi.launch_count()
    free0 = torch.cuda.mem_get_info()[0]
    for c, (dg2, db2, dy2) in zip((first, second), outs):    # back to back on one stream, nothing in between
        ops.bn_bwd_noalloc(c["dz"], c["y"], c["scale"], c["shift"], c["mean"], c["invstd"], c["gamma"], c["relu"],
                           dy2, dg2, db2)
    torch.cuda.synchronize()
    assert _ffi.launch_count() - n0 == 4                      # reduce + apply per call
    for c, o in zip((first, second), outs):
        _compare(c, *o)
    # a third call on the first inputs again (the array now holds the second call's sums), sums not requested
    free1 = torch.cuda.mem_get_info()[0]
    dy3 = Act.empty(*first["dims"], first["C"])
    torch.cuda.synchronize()
    free2 = torch.cuda.mem_get_info()[0]
    ops.bn_bwd_noalloc(first["dz"], first["y"], first["scale"], first["shift"], first["mean"], first["invstd"],
                       first["gamma"], True, dy3)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free2, "device memory shrank across a sfb_bn_bwd_noalloc call"
    assert kc.relerr(kc.from_act(dy3), kc.from_act(outs[0][2])) < 1e-2
    del free0, free1


def test_bn_bwd_noalloc_widest_supported_and_limit():
    c = _case(ops.NOALLOC_MAX_C, True, seed=13, dims=(1, 2, 3, 4))
    dg2, db2 = torch.empty(c["C"], device=DEV), torch.empty(c["C"], device=DEV)
    dy2 = Act.empty(*c["dims"], c["C"])
    ops.bn_bwd_noalloc(c["dz"], c["y"], c["scale"], c["shift"], c["mean"], c["invstd"], c["gamma"], True, dy2, dg2, db2)
    torch.cuda.synchronize()
    _compare(c, dg2, db2, dy2)
    # beyond the static array: refused with the limit named, nothing launched
    Cbig = ops.NOALLOC_MAX_C + 8
    big = Act.empty(1, 1, 2, 2, Cbig, zero=True)
    v = torch.ones(Cbig, device=DEV)
    n0 = _ffi.launch_count()
    with pytest.raises(_ffi.SfbError, match="SFB_NOALLOC_MAX_C = 2048"):
        ops.bn_bwd_noalloc(big, big, v, v, v, v, v, True, Act.empty(1, 1, 2, 2, Cbig))
    assert _ffi.launch_count() == n0
    torch.cuda.synchronize()
