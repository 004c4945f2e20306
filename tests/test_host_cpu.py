"""CPU tests (no GPU, no compute calls): the C-ABI library loads and exports every symbol include/sfb200.h declares,
argument validation rejects bad requests before any launch, the drop-in keeps the reference's module tree /
state_dict, and the engine plans a full training step on the meta device."""
import ctypes as C
import os
import re

import pytest
import torch

from oracle import slowfast_oracle as so
from video_classification_b200 import _ffi
from video_classification_b200 import build as sfb_build
from video_classification_b200.my_slowfast import (FuseFastToSlow, MyFastToSlowFusionBuilder, _MODEL_STAGE_DEPTH,
                                                   init_my_slowfast)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    sfb_build.build()
    return _ffi.lib()


def test_header_symbols_all_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "sfb200.h")).read()
    declared = sorted(set(re.findall(r"\b(sfb_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/sfb200.h but not exported by libsfb200.so"
    assert sorted(declared) == _ffi.ALL_SYMBOLS, "ctypes binding and header disagree"
    assert lib.sfb_version() >= 100
    assert lib.sfb_strerror(0) == b"ok" and b"bad argument" in lib.sfb_strerror(-1)


def test_validation_rejects_before_launch(lib):
    r, k = C.c_int32(), C.c_int32()
    assert lib.sfb_packed_weight_dims(64, 8, 49, C.byref(r), C.byref(k)) == 0
    assert (r.value, k.value) == (64, 448)
    assert lib.sfb_packed_weight_dims(2048, 1280, 3, C.byref(r), C.byref(k)) == 0
    assert r.value == 2048 and k.value == 3840
    assert lib.sfb_packed_weight_dims(64, 5, 49, C.byref(r), C.byref(k)) == -1  # Cg must be a multiple of 8
    n0 = lib.sfb_launch_count()
    g = _ffi.ConvGeom(1, 4, 8, 8, 12, 4, 8, 8, 16, 1, 1, 1, 1, 1, 1, 0, 0, 0, 12, 0, 16, 0, 0)  # Cg=12 not /8
    assert lib.sfb_conv_gemm(C.byref(g), 16, None, 16, 32, 64, 16, None, 0, None) == -1
    assert lib.sfb_conv_gemm(None, 16, None, 16, 32, 64, 16, None, 0, None) == -1
    assert lib.sfb_bn_finalize(None, 8, 10, None, None, 1e-5, 0.1, None, None, None, None, None, None, None) == -1
    assert lib.sfb_fc_fwd(None, None, None, 1, 1, 1, None, None) == -1
    assert lib.sfb_launch_count() == n0, "a rejected call must not launch anything"


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_ffi, "_lib", None)
    monkeypatch.setattr(_ffi, "LIB_PATH", "/nonexistent/libsfb200.so")
    with pytest.raises(_ffi.SfbError, match="no CPU/PyTorch fallback"):
        _ffi.lib()


def test_dropin_module_tree_matches_reference_contract():
    torch.manual_seed(0)
    m = init_my_slowfast(so.make_cfg(249, True), (5, 15), (64, 8))
    torch.manual_seed(0)
    o = so.build_oracle()
    sd_m, sd_o = m.state_dict(), o.state_dict()
    assert list(sd_m.keys()) == list(sd_o.keys()) and len(sd_m) == 734
    for k in sd_m:
        assert sd_m[k].shape == sd_o[k].shape and sd_m[k].dtype == sd_o[k].dtype, k
        assert torch.equal(sd_m[k], sd_o[k]), f"{k}: initialisation differs from the reference (same seed)"
    assert sum(p.numel() for p in m.parameters()) == 38077321
    # the 12 names train.py:94-108 deletes from the canonical checkpoint exist with the reference's shapes
    for k in ("blocks.0.multipathway_blocks.0.conv.weight", "blocks.0.multipathway_blocks.1.conv.weight",
              "blocks.6.proj.weight", "blocks.6.proj.bias", "blocks.0.multipathway_fusion.conv_fast_to_slow.0.weight"):
        assert k in sd_m
    assert tuple(sd_m["blocks.0.multipathway_blocks.0.conv.weight"].shape) == (64, 5, 1, 7, 7)
    assert tuple(sd_m["blocks.0.multipathway_fusion.conv_fast_to_slow.0.weight"].shape) == (16, 8, 3, 1, 1)
    m.load_state_dict(o.state_dict(), strict=True)
    o.load_state_dict(m.state_dict(), strict=True)
    assert _MODEL_STAGE_DEPTH[50] == (3, 4, 6, 3)
    assert isinstance(m.blocks[0].multipathway_fusion, FuseFastToSlow)
    assert isinstance(MyFastToSlowFusionBuilder.build_fusion_builder(8).create_module(64, 4), torch.nn.Identity)
    # MODEL.FUSE = False: nn.Identity fusion, 640 fewer... keys as in the reference
    m2 = init_my_slowfast(so.make_cfg(249, False), (5, 15), (64, 8))
    assert list(m2.state_dict().keys()) == list(so.build_oracle(fuse=False).state_dict().keys())
    with pytest.raises(AssertionError):
        init_my_slowfast(so.make_cfg(), (5,), (64,))


def test_cpu_model_refuses_to_run():
    m = init_my_slowfast(so.make_cfg(), (5, 15), (64, 8))
    with pytest.raises(_ffi.SfbError, match="no CPU fallback"):
        m([torch.randn(1, 5, 4, 64, 64), torch.randn(1, 15, 4, 64, 64)])


@pytest.mark.parametrize("variant,shape", [("REF", (2, 20, 20, 64, 64)), ("HYB", (2, 8, 32, 224, 224))])
def test_engine_plans_a_training_step_on_meta(variant, shape, lib):
    from video_classification_b200.engine import SlowFastEngine
    kw = {} if variant == "REF" else dict(fusion_conv_stride=(4, 1, 1), head_pool_kernel_sizes=((8, 7, 7), (32, 7, 7)))
    m = init_my_slowfast(so.make_cfg(), (5, 15), (64, 8), **kw)
    eng = SlowFastEngine(m, dry_run=True)
    eng.refresh_weights(True)
    plan = eng.build_plan(*shape, True)
    n_conv = len(eng.layers)
    assert n_conv == 2 + 4 + 16 * 2 * 3 + 4 * 2  # stems, laterals, 16 bottlenecks x 2 pathways x 3, 4 shortcuts x 2
    # the two stems run as space-to-depth row-window kernels, everything else as the generic implicit GEMM
    convs = [op for op in plan.fwd if op[0] is lib.sfb_conv_gemm]
    assert len(convs) == n_conv - 2 and len([op for op in plan.fwd if op[0] is lib.sfb_stem_conv_fprop]) == 2
    # (pixel-folded Fast-pathway layers write their gradient through sfb_conv_wgrad_folded)
    wgrads = [op for op in plan.bwd if op[0] in (lib.sfb_conv_wgrad, lib.sfb_conv_wgrad_folded)]
    assert any(op[0] is lib.sfb_conv_wgrad_folded for op in wgrads)
    assert len(wgrads) == n_conv - 2 and len([op for op in plan.bwd if op[0] is lib.sfb_stem_conv_wgrad]) == 2
    dgrads = [op for op in plan.bwd if op[0] in (lib.sfb_conv_gemm, lib.sfb_conv_dgrad_bn)]
    assert len(dgrads) == n_conv - 2  # the stems need no input gradient
    # conv_c's and conv_b's input gradients reduce the BatchNorm backward of conv_b / conv_a in their epilogue:
    # two per bottleneck; only block tails, shortcuts, laterals and stems keep a separate reduction pass
    fused = [op for op in dgrads if op[0] is lib.sfb_conv_dgrad_bn]
    assert len(fused) == 2 * 16 * 2
    reduces = [op for op in plan.bwd if op[0] is lib.sfb_bn_bwd_reduce]
    # ... and the eight blocks with a projection shortcut reduce both BatchNorms of their tail in one pass
    assert len([op for op in plan.bwd if op[0] is lib.sfb_bn_bwd_tail2]) == 8
    assert len(reduces) == n_conv - len(fused) - 16
    assert tuple(plan.logits.shape) == (shape[0], 249)
    # the lateral output lands behind the Slow stage output in ONE buffer (torch.cat never happens)
    bufs = eng.tagged_buffers(plan)
    lat = bufs["blocks.1.multipathway_fusion.conv_fast_to_slow.0.z"]
    slow = bufs["blocks.1.multipathway_blocks.0.res_blocks.2.O"]
    assert lat.buf is slow.buf and slow.coff == 0 and lat.coff == slow.C == 256 and lat.buf.shape[-1] == 320
    with pytest.raises(RuntimeError, match="Sizes of tensors must match"):
        eng.get_plan(torch.empty(1, 5, 8, 64, 64, device="meta"), torch.empty(1, 15, 32 if variant == "REF" else 8, 64, 64, device="meta"), False)


def test_dropout_seed_follows_torch_generator_and_rank(monkeypatch):
    """engine._draw_dropout_seed: reproducible under torch.manual_seed, different from step to step and from rank to
    rank (the reference's nn.Dropout draws from the process generator at every call)"""
    import torch.distributed as dist
    from oracle import slowfast_oracle as so
    from video_classification_b200.engine import SlowFastEngine
    from video_classification_b200.my_slowfast import init_my_slowfast
    eng = SlowFastEngine(init_my_slowfast(so.make_cfg(), (5, 15), (64, 8)), dry_run=True)
    torch.manual_seed(3)
    a1, a2 = eng._draw_dropout_seed(), eng._draw_dropout_seed()
    torch.manual_seed(3)
    b1 = eng._draw_dropout_seed()
    torch.manual_seed(4)
    c1 = eng._draw_dropout_seed()
    assert a1 == b1 and a1 != a2 and a1 != c1
    assert all(-(1 << 63) <= v < (1 << 63) for v in (a1, a2, c1))
    monkeypatch.setattr(dist, "is_initialized", lambda: True)
    monkeypatch.setattr(dist, "get_rank", lambda: 1)
    torch.manual_seed(3)
    r1 = eng._draw_dropout_seed()
    assert r1 != a1


def test_every_plan_owns_its_batchnorm_scratch():
    """the coefficients a backward pass reads are per plan: building / running another shape or an eval plan cannot
    overwrite them (dry-run planning on the meta device: the pointers baked into the op lists are disjoint)"""
    from oracle import slowfast_oracle as so
    from video_classification_b200.engine import SlowFastEngine
    from video_classification_b200.my_slowfast import init_my_slowfast
    eng = SlowFastEngine(init_my_slowfast(so.make_cfg(), (5, 15), (64, 8)), dry_run=True)
    eng.refresh_weights(True)
    p1 = eng.build_plan(2, 8, 8, 64, 64, True)
    p2 = eng.build_plan(2, 8, 8, 64, 64, False)
    p3 = eng.build_plan(1, 8, 8, 64, 64, True)
    assert p1.bn_flat is not p2.bn_flat and p1.bn_flat is not p3.bn_flat
    assert p1.bn_flat.numel() == p2.bn_flat.numel() == p3.bn_flat.numel() > 0
    assert p1.stats_flat.numel() < p1.bn_flat.numel()


# ---- BASELINE configs[0]: Slow-only 3D-ResNet (config/res3d.yaml; reference train.py:79-83) --------------------------
def test_slow_r50_tree_matches_the_create_resnet_restatement():
    from video_classification_b200.res3d import slow_r50
    oracle = so.build_res3d_oracle(head_pool_kernel_size=(4, 4, 4))
    product = slow_r50(5, 400, (4, 4, 4))
    sd_o, sd_p = oracle.state_dict(), product.state_dict()
    assert list(sd_o.keys()) == list(sd_p.keys()) and len(sd_o) == 320
    assert all(sd_o[k].shape == sd_p[k].shape and sd_o[k].dtype == sd_p[k].dtype for k in sd_o)
    assert sum(p.numel() for p in product.parameters()) == 32460368  # slow_r50 (32 454 096) + 2 extra stem channels
    assert tuple(product.blocks[0].conv.weight.shape) == (64, 5, 1, 7, 7)  # train.py:81
    product.load_state_dict(sd_o, strict=True)


def test_slow_r50_plans_on_meta_and_rejects_oversized_head_pool(lib):
    from video_classification_b200.engine import SlowFastEngine
    from video_classification_b200.res3d import slow_r50
    eng = SlowFastEngine(slow_r50(5, 400, (4, 4, 4)), dry_run=True)
    assert eng.P == 1 and len(eng.layers) == 1 + 16 * 3 + 4 and all(l is None for l in eng.laterals)
    eng.refresh_weights(True)
    plan = eng.build_plan(2, 16, 0, 112, 112, True)  # configs[0]: batch 2, 16 x 112^2
    assert len([op for op in plan.fwd if op[0] is lib.sfb_stem_conv_fprop]) == 1
    assert len([op for op in plan.bwd if op[0] in (lib.sfb_conv_wgrad, lib.sfb_conv_wgrad_folded)]) == 16 * 3 + 4
    assert tuple(plan.logits.shape) == (2, 400)
    # the hub model's (4,7,7) head pool does not fit the 4x4 map of a 112^2 clip: same failure as nn.AvgPool3d
    eng7 = SlowFastEngine(slow_r50(5, 400, (4, 7, 7)), dry_run=True)
    eng7.refresh_weights(False)
    with pytest.raises(_ffi.SfbError, match="head pool kernel"):
        eng7.build_plan(2, 16, 0, 112, 112, False)
    eng7.build_plan(1, 8, 0, 224, 224, False)
