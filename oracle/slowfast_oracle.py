"""ORACLE (test infrastructure, not product code; only tests/, __graft_entry__.smoke() and bench.py's CPU
baseline may import it).

Self-contained fp32 PyTorch restatement of the reference's SlowFast hot path, usable where /root/reference
does not exist (the GPU box):

    build_oracle(...)  restates  init_my_slowfast            reference model/my_slowfast.py:44-126
    LateralFuse        restates  FuseFastToSlow(.forward)     reference model/my_slowfast.py:260-283, 334-344
    _lateral_module    restates  MyFastToSlowFusionBuilder.create_module   model/my_slowfast.py:174-244
                                 (+ build_fusion_builder parameters, :246-257)
    everything else    is oracle/pv_restate.py (pytorchvideo 0.1.5 semantics, restated)

Pinned in this container against the reference file imported VERBATIM (oracle/ref_import.py):
tests/test_oracle.py checks identical state_dict keys/shapes and bit-identical logits/gradients on shared
weights, for REF and HYB variants.  The reference itself ships no tests/golden vectors: PARITY UNPINNED by
reference tests; golden fixtures in tests/golden/ are generated from the verbatim import by
oracle/make_golden.py.
"""
from types import SimpleNamespace
from typing import Sequence, Tuple

import torch
import torch.nn as nn

from . import pv_restate as pv

STAGE_DEPTHS_R50 = (3, 4, 6, 3)  # reference model/my_slowfast.py:129-134 (_MODEL_STAGE_DEPTH[50])


class LateralFuse(nn.Module):
    """Fast->Slow lateral connection.  The live forward uses only conv/norm/activation index 0 and a channel
    concat; `residual` and `res_unit` are parameters the reference creates but never reaches from forward()
    (they must exist for state_dict compatibility: train.py:212 loads with strict=True)."""

    def __init__(self, conv_fast_to_slow, residual, norm, activation, res_unit):
        super().__init__()
        # registration order == reference's set_attributes(locals()) order (model/my_slowfast.py:268-283)
        self.conv_fast_to_slow = conv_fast_to_slow
        self.residual = residual
        self.norm = norm
        self.activation = activation
        self.res_unit = res_unit

    def forward(self, x):
        slow, fast = x[0], x[1]
        lat = self.conv_fast_to_slow[0](fast)
        if self.norm is not None:
            lat = self.norm[0](lat)
        if self.activation is not None:
            lat = self.activation[0](lat)
        return [torch.cat([slow, lat], dim=1), fast]


def _lateral_module(slow_dim: int, reduction: int, kernel, stride, max_stage_idx: int, stage_idx: int) -> nn.Module:
    if stage_idx > max_stage_idx:
        return nn.Identity()
    fast_in = slow_dim // reduction
    fast_out = int(fast_in * 2)  # conv_fusion_channel_ratio = 2
    cat_dim = slow_dim + fast_out
    conv = nn.ModuleList([nn.Conv3d(fast_in, fast_out, kernel_size=kernel, stride=stride,
                                    padding=[k // 2 for k in kernel], bias=False)])
    residual = nn.Sequential(nn.Conv3d(slow_dim, cat_dim, kernel_size=1, bias=True), nn.ReLU(inplace=True))
    norm = nn.ModuleList([nn.BatchNorm3d(num_features=fast_in * 2, eps=1e-5, momentum=0.1)])
    act = nn.ModuleList([nn.ReLU()])
    q = cat_dim // 4
    res_unit = nn.Sequential(
        nn.Conv3d(cat_dim, q, (1, 1, 1)), nn.ReLU(inplace=True), nn.BatchNorm3d(q),
        nn.Conv3d(q, q, (1, 3, 3), padding=(0, 1, 1)), nn.ReLU(inplace=True), nn.BatchNorm3d(q),
        nn.Conv3d(q, cat_dim, (1, 1, 1)))
    return LateralFuse(conv, residual, norm, act, res_unit)


def build_oracle(num_class: int = 249, fuse: bool = True, input_channels: Sequence[int] = (5, 15),
                 stem_dim_outs: Sequence[int] = (64, 8), *, lateral_kernel=(3, 1, 1), lateral_stride=(1, 1, 1),
                 head_pool_kernel_sizes=((4, 2, 2), (4, 2, 2)), stem_conv_kernel_sizes=((1, 7, 7), (1, 7, 7)),
                 model_depth: int = 50) -> nn.Module:
    """Defaults == the reference as written (REF).  HYB = lateral_stride (4,1,1), head pools (8,7,7)/(32,7,7)."""
    n_path = len(input_channels)
    assert n_path >= 2 and len(stem_dim_outs) == n_path
    reduction = tuple(int(stem_dim_outs[0]) // int(c) for c in stem_dim_outs[1:])
    depths = pv._MODEL_STAGE_DEPTH[model_depth]
    if fuse:
        def fusion_builder(fusion_dim_in, stage_idx):
            return _lateral_module(fusion_dim_in, reduction[0], tuple(lateral_kernel), tuple(lateral_stride),
                                   len(STAGE_DEPTHS_R50) - 1, stage_idx)
        fusion_ratio = 2 * (n_path - 1)
    else:
        fusion_builder = nn.Identity
        fusion_ratio = 0
    slow_a = ((1, 1, 1), (1, 1, 1), (3, 1, 1), (3, 1, 1))
    fast_a = ((3, 1, 1),) * 4
    return pv.create_slowfast(
        slowfast_channel_reduction_ratio=reduction, slowfast_conv_channel_fusion_ratio=fusion_ratio,
        model_depth=model_depth, model_num_class=num_class, input_channels=tuple(input_channels),
        fusion_builder=fusion_builder,
        stem_function=(pv.create_res_basic_stem,) * n_path, stem_dim_outs=tuple(stem_dim_outs),
        stem_conv_kernel_sizes=tuple(stem_conv_kernel_sizes), stem_conv_strides=((1, 2, 2),) * n_path,
        stem_pool=(nn.MaxPool3d,) * n_path, stem_pool_kernel_sizes=((1, 3, 3),) * n_path,
        stem_pool_strides=((1, 2, 2),) * n_path,
        stage_conv_a_kernel_sizes=[slow_a] + [fast_a] * (n_path - 1),
        stage_conv_b_kernel_sizes=(((1, 3, 3),) * 4,) * n_path, stage_conv_b_num_groups=((1, 1, 1, 1),) * n_path,
        stage_conv_b_dilations=(((1, 1, 1),) * 4,) * n_path, stage_spatial_strides=((1, 2, 2, 2),) * n_path,
        stage_temporal_strides=((1, 1, 1, 1),) * n_path, bottleneck=((pv.create_bottleneck_block,) * 4,) * n_path,
        head_pool_kernel_sizes=tuple(head_pool_kernel_sizes))


def build_res3d_oracle(num_class: int = 400, input_channel: int = 5, head_pool_kernel_size=(4, 7, 7)) -> nn.Module:
    """BASELINE config 1 (config/res3d.yaml): the model the reference's res3d path intends -- pytorchvideo `slow_r50`
    (create_resnet defaults, 400 classes) with its stem convolution replaced by
    Conv3d(5, 64, (1,7,7), stride (1,2,2), padding (0,3,3), bias=False) (reference train.py:79-83, commented out;
    (deprecated)/train_3dresnet.py:47-49).  Building create_resnet with input_channel = 5 gives that module tree
    directly (same state_dict keys as the hub model after the replacement: `blocks.0.conv.weight` is (64,5,1,7,7)).
    The head's own AvgPool3d is (4,7,7) by default, which needs a >= 7x7 final map (224^2 input); for the 112^2 clips of
    the config the survey clips it to (4,4,4) -- pass head_pool_kernel_size."""
    return pv.create_resnet(input_channel=input_channel, model_num_class=num_class,
                            head_pool_kernel_size=tuple(head_pool_kernel_size))


VARIANTS = {
    "REF": dict(),
    "HYB": dict(lateral_stride=(4, 1, 1), head_pool_kernel_sizes=((8, 7, 7), (32, 7, 7))),
}


def make_cfg(num_class=249, fuse=True):
    """Stand-in for the yacs CfgNode the reference passes (only the two knobs the hot path reads:
    config/defaults.py:14 CHALEARN.NUM_CLASS, :44 MODEL.FUSE)."""
    return SimpleNamespace(CHALEARN=SimpleNamespace(NUM_CLASS=num_class), MODEL=SimpleNamespace(FUSE=fuse))


def randomize_bn_(model: nn.Module, seed: int = 123, final_gamma: float = 1.0) -> nn.Module:
    """default init zeroes the last BN gamma of every bottleneck, which makes branch2 and 224
    gradients identically zero; parity runs re-randomise every BatchNorm (affine + running stats), seeded.
    `final_gamma` scales the gamma of each bottleneck's last BatchNorm (`block_final_bn`): 1.0 makes every
    residual branch as strong as its shortcut, which at random init is a numerically chaotic network in
    train mode (batch-stat BN amplifies any perturbation ~1 %/layer, so bf16-vs-fp32 comparisons measure
    chaos, not kernels); ~0.1 is close to the reference's real init (gamma = 0) while keeping every
    gradient non-zero."""
    g = torch.Generator().manual_seed(seed)
    for m in model.modules():
        if isinstance(m, nn.modules.batchnorm._BatchNorm):
            with torch.no_grad():
                m.weight.copy_(torch.rand(m.weight.shape, generator=g) + 0.5)
                if getattr(m, "block_final_bn", False):
                    m.weight.mul_(final_gamma)
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
                m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) + 0.5)
    return model


def synthetic_clips(n: int, t_slow: int, t_fast: int, size: int, input_channels=(5, 15), seed: int = 1,
                    reference_layout: bool = True) -> Tuple[torch.Tensor, torch.Tensor]:
    """Seeded clips.  With reference_layout and t_slow == t_fast the two pathway tensors are the
    NON-CONTIGUOUS views train.py:136-140 produces from one (N,T,21,S,S) batch (channel 20 dropped)."""
    g = torch.Generator().manual_seed(seed)
    if reference_layout and t_slow == t_fast and sum(input_channels) <= 21:
        x = torch.randn((n, t_slow, 21, size, size), generator=g).permute(0, 2, 1, 3, 4)
        c0 = input_channels[0]
        return x[:, 0:c0], x[:, c0:c0 + input_channels[1]]
    xs = torch.randn((n, input_channels[0], t_slow, size, size), generator=g)
    xf = torch.randn((n, input_channels[1], t_fast, size, size), generator=g)
    return xs, xf


def disable_dropout_(model: nn.Module) -> nn.Module:
    for m in model.modules():
        if isinstance(m, nn.Dropout):
            m.p = 0.0
    return model


# ------------------------------------------------------------------------------------------------------------
# bf16-storage twin: same fp32 arithmetic and the same graph, but every tensor the B200 engine keeps in HBM as
# bf16 (conv operands, raw conv outputs, block outputs, conv weights) is rounded to bf16 at that point, with a
# straight-through gradient.  Why: elementwise gradients of a ReLU network are discontinuous in the forward
# activations, so ANY bf16 implementation (PyTorch autocast + cuDNN included, see tools/autocast_yardstick.py:
# median per-parameter rel-L2 0.25 against fp32 on this model) differs from the pure-fp32 oracle through
# ReLU-mask flips, not through kernel bugs.  Against the twin the masks agree, so the 3e-2 gradient bar of
# BASELINE.json's north_star is a meaningful kernel test.  Works on the restated oracle and on the reference's
# verbatim module tree alike (only nn.Conv3d modules and `res_blocks.*` children are hooked).
# ------------------------------------------------------------------------------------------------------------
def _ste_bf16(t: torch.Tensor) -> torch.Tensor:
    return t + (t.to(torch.bfloat16).to(t.dtype) - t).detach()


def round_conv_weights_to_bf16_(model: nn.Module) -> nn.Module:
    with torch.no_grad():
        for m in model.modules():
            if isinstance(m, nn.Conv3d):
                m.weight.copy_(m.weight.to(torch.bfloat16).float())
    return model


def add_bf16_storage_hooks_(model: nn.Module):
    """returns the list of hook handles (call .remove() on each to get the pure-fp32 oracle back)"""
    handles = []
    for name, m in model.named_modules():
        parts = name.split(".")
        if isinstance(m, nn.Conv3d) and ".residual." not in name and ".res_unit." not in name:
            handles.append(m.register_forward_pre_hook(lambda mod, inp: (_ste_bf16(inp[0]),)))
            handles.append(m.register_forward_hook(lambda mod, inp, out: _ste_bf16(out)))
        elif len(parts) >= 2 and parts[-2] == "res_blocks":
            handles.append(m.register_forward_hook(lambda mod, inp, out: _ste_bf16(out)))
    return handles
