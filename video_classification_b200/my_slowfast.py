"""Drop-in for the reference's model/my_slowfast.py (zc402/video-classification).

Same public surface (reference model/my_slowfast.py:44-126, 129-134, 136-257, 260-344):
    init_my_slowfast(cfg, input_channels, stem_dim_outs) -> nn.Module
    _MODEL_STAGE_DEPTH, MyFastToSlowFusionBuilder, FuseFastToSlow
and the same module tree / parameter names / state_dict (734 entries for (5,15),(64,8)), so
train.py:93-123 (delete_mismatch + load_state_dict), :182 (optim.Adam(model.parameters())),
:192/:212 (checkpoint save / strict reload) work unchanged.

What differs is underneath: the torch.nn modules here only HOLD parameters and buffers; forward and
backward of the whole network run in libsfb200.so (hand-written sm_100a kernels: tcgen05 implicit-GEMM
Conv3d whose epilogues produce the BatchNorm batch statistics in training, apply the folded BatchNorm +
residual + ReLU in inference and reduce the BatchNorm backward sums in the input-gradient GEMMs; HBM-bound
normalise / residual-tail passes; zero-copy lateral concat, ...) driven by engine.SlowFastEngine through one
torch.autograd.Function.  There is no PyTorch/cuDNN fallback.

Keyword-only extras (defaults == the reference as written) select the canonical 8/32 alpha=4 graph the
benchmark names: fusion_conv_kernel, fusion_conv_stride, head_pool_kernel_sizes,
stem_conv_kernel_sizes.
"""
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

_MODEL_STAGE_DEPTH = {
    18: (1, 1, 1, 1),
    50: (3, 4, 6, 3),
    101: (3, 4, 23, 3),
    152: (3, 8, 36, 3),
}


def _half(k):
    return tuple(int(v) // 2 for v in k)


class _ParamTree(nn.Module):
    """Base of the parameter-holding tree: sub-modules are never called one by one."""

    def forward(self, *a, **k):  # pragma: no cover
        raise RuntimeError("sub-modules of the B200 SlowFast are parameter holders; call the top-level model")


class _Stem(_ParamTree):
    def __init__(self, cin, cout, kernel):
        super().__init__()
        self.conv = nn.Conv3d(cin, cout, kernel_size=kernel, stride=(1, 2, 2), padding=_half(kernel), bias=False)
        self.norm = nn.BatchNorm3d(cout, eps=1e-5, momentum=0.1)
        self.activation = nn.ReLU()
        self.pool = nn.MaxPool3d(kernel_size=(1, 3, 3), stride=(1, 2, 2), padding=(0, 1, 1))


class _Bottleneck(_ParamTree):
    def __init__(self, cin, cinner, cout, ka, spatial_stride):
        super().__init__()
        self.conv_a = nn.Conv3d(cin, cinner, kernel_size=ka, stride=(1, 1, 1), padding=_half(ka), bias=False)
        self.norm_a = nn.BatchNorm3d(cinner, eps=1e-5, momentum=0.1)
        self.act_a = nn.ReLU()
        self.conv_b = nn.Conv3d(cinner, cinner, kernel_size=(1, 3, 3), stride=(1, spatial_stride, spatial_stride),
                                padding=(0, 1, 1), bias=False)
        self.norm_b = nn.BatchNorm3d(cinner, eps=1e-5, momentum=0.1)
        self.act_b = nn.ReLU()
        self.conv_c = nn.Conv3d(cinner, cout, kernel_size=(1, 1, 1), bias=False)
        self.norm_c = nn.BatchNorm3d(cout, eps=1e-5, momentum=0.1)
        self.norm_c.block_final_bn = True


class _ResBlock(_ParamTree):
    def __init__(self, cin, cinner, cout, ka, spatial_stride):
        super().__init__()
        if cin != cout or spatial_stride != 1:
            self.branch1_conv = nn.Conv3d(cin, cout, kernel_size=(1, 1, 1), stride=(1, spatial_stride, spatial_stride),
                                          bias=False)
            self.branch1_norm = nn.BatchNorm3d(cout, eps=1e-5, momentum=0.1)
        else:
            self.branch1_conv = None
            self.branch1_norm = None
        self.branch2 = _Bottleneck(cin, cinner, cout, ka, spatial_stride)
        self.activation = nn.ReLU()


class _ResStage(_ParamTree):
    def __init__(self, depth, cin, cinner, cout, ka, spatial_stride):
        super().__init__()
        self.res_blocks = nn.ModuleList(
            [_ResBlock(cin if i == 0 else cout, cinner, cout, ka, spatial_stride if i == 0 else 1) for i in range(depth)])


class _MultiPathWayWithFuse(_ParamTree):
    def __init__(self, blocks, fusion):
        super().__init__()
        self.multipathway_blocks = nn.ModuleList(blocks)
        self.multipathway_fusion = fusion


class _PoolConcatPathway(_ParamTree):
    def __init__(self, kernels):
        super().__init__()
        self.pool = nn.ModuleList([nn.AvgPool3d(kernel_size=tuple(k), stride=(1, 1, 1), padding=(0, 0, 0)) for k in kernels])


class _Head(_ParamTree):
    def __init__(self, cin, num_class, dropout_rate=0.5):
        super().__init__()
        self.dropout = nn.Dropout(dropout_rate) if dropout_rate > 0 else None
        self.proj = nn.Linear(cin, num_class)
        self.output_pool = nn.AdaptiveAvgPool3d(1)


class FuseFastToSlow(_ParamTree):
    """Fast->Slow lateral connection (reference model/my_slowfast.py:260-344).  Live path: conv -> BN -> ReLU on
    the Fast tensor, channel-concatenated behind the Slow tensor.  In this implementation the lateral conv
    writes its raw output and, in training, one normalise+ReLU pass (in inference the GEMM epilogue) puts the
    activated values straight behind the Slow channels of the shared concat buffer (no torch.cat copy in either
    direction).  `residual` / `res_unit` are created exactly like the
    reference does and, exactly like there, never used by forward (their .grad stays None)."""

    def __init__(self, conv_fast_to_slow: nn.ModuleList, residual: nn.Module, norm: Optional[nn.ModuleList] = None,
                 activation: Optional[nn.ModuleList] = None, res_unit: nn.Module = None) -> None:
        super().__init__()
        self.conv_fast_to_slow = conv_fast_to_slow
        self.residual = residual
        self.norm = norm
        self.activation = activation
        self.res_unit = res_unit


class MyFastToSlowFusionBuilder:
    """Same constructor / create_module contract as reference model/my_slowfast.py:136-257."""

    def __init__(self, slowfast_channel_reduction_ratio: int, conv_fusion_channel_ratio: float,
                 conv_kernel_size: Tuple[int], conv_stride: Tuple[int], norm: Callable = nn.BatchNorm3d,
                 norm_eps: float = 1e-5, norm_momentum: float = 0.1, activation: Callable = nn.ReLU,
                 max_stage_idx: int = 3) -> None:
        self.slowfast_channel_reduction_ratio = slowfast_channel_reduction_ratio
        self.conv_fusion_channel_ratio = conv_fusion_channel_ratio
        self.conv_kernel_size = tuple(conv_kernel_size)
        self.conv_stride = tuple(conv_stride)
        if norm is not nn.BatchNorm3d or activation is not nn.ReLU:
            raise NotImplementedError("the B200 lateral kernel fuses BatchNorm3d + ReLU; other norm/activation "
                                      "choices are not implemented")
        self.norm, self.norm_eps, self.norm_momentum, self.activation = norm, norm_eps, norm_momentum, activation
        self.max_stage_idx = max_stage_idx

    def create_module(self, fusion_dim_in: int, stage_idx: int) -> nn.Module:
        if stage_idx > self.max_stage_idx:
            return nn.Identity()
        fast_in = fusion_dim_in // self.slowfast_channel_reduction_ratio
        fast_out = int(fast_in * self.conv_fusion_channel_ratio)
        cat_dim = fusion_dim_in + fast_out
        conv = nn.ModuleList([nn.Conv3d(fast_in, fast_out, kernel_size=self.conv_kernel_size, stride=self.conv_stride,
                                        padding=_half(self.conv_kernel_size), bias=False)])
        residual = nn.Sequential(nn.Conv3d(fusion_dim_in, cat_dim, kernel_size=(1, 1, 1), bias=True), nn.ReLU(inplace=True))
        norm = nn.ModuleList([nn.BatchNorm3d(fast_in * self.conv_fusion_channel_ratio, eps=self.norm_eps,
                                             momentum=self.norm_momentum)])
        act = nn.ModuleList([nn.ReLU()])
        q = cat_dim // 4
        res_unit = nn.Sequential(nn.Conv3d(cat_dim, q, (1, 1, 1)), nn.ReLU(inplace=True), nn.BatchNorm3d(q),
                                 nn.Conv3d(q, q, (1, 3, 3), padding=(0, 1, 1)), nn.ReLU(inplace=True), nn.BatchNorm3d(q),
                                 nn.Conv3d(q, cat_dim, (1, 1, 1)))
        return FuseFastToSlow(conv_fast_to_slow=conv, residual=residual, norm=norm, activation=act, res_unit=res_unit)

    @classmethod
    def build_fusion_builder(cls, slowfast_channel_reduction_ratio, conv_kernel_size=(3, 1, 1), conv_stride=(1, 1, 1)):
        return MyFastToSlowFusionBuilder(
            slowfast_channel_reduction_ratio=slowfast_channel_reduction_ratio, conv_fusion_channel_ratio=2,
            conv_kernel_size=conv_kernel_size, conv_stride=conv_stride, norm=nn.BatchNorm3d, activation=nn.ReLU,
            max_stage_idx=len(_MODEL_STAGE_DEPTH[50]) - 1)


def _init_weights(model: nn.Module, fc_std: float = 0.01) -> None:
    """Same initialisation (and same RNG consumption order) as pytorchvideo's init_net_weights, which the
    reference inherits through Net.__init__: Kaiming-normal fan_out convs, BN gamma 1 (0 for the last BN of a
    bottleneck), Linear N(0, 0.01), zero biases."""
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.Conv3d)):
            nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.modules.batchnorm._NormBase):
            m.weight.data.fill_(0.0 if getattr(m, "block_final_bn", False) else 1.0)
            m.bias.data.zero_()
        if isinstance(m, nn.Linear):
            m.weight.data.normal_(mean=0.0, std=fc_std)
            m.bias.data.zero_()


class SlowFastB200(_ParamTree):
    """nn.Module with the reference's `blocks.*` tree; forward([slow, fast]) -> (N, NUM_CLASS) logits."""

    def __init__(self, num_class, fuse, input_channels, stem_dim_outs, fusion_conv_kernel, fusion_conv_stride,
                 head_pool_kernel_sizes, stem_conv_kernel_sizes, model_depth=50):
        super().__init__()
        if len(input_channels) != 2:
            # the reference's FuseFastToSlow.forward (my_slowfast.py:334-344) only consumes x[0], x[1]
            raise NotImplementedError("exactly one Slow and one Fast pathway are supported")
        slow_c, fast_c = int(stem_dim_outs[0]), int(stem_dim_outs[1])
        ratio = slow_c // fast_c
        depths = _MODEL_STAGE_DEPTH[model_depth]
        self.graph_cfg = dict(num_class=num_class, fuse=bool(fuse), input_channels=tuple(input_channels),
                              stem_dim_outs=(slow_c, fast_c), ratio=ratio, depths=depths,
                              fusion_conv_kernel=tuple(fusion_conv_kernel), fusion_conv_stride=tuple(fusion_conv_stride),
                              head_pool_kernel_sizes=tuple(tuple(k) for k in head_pool_kernel_sizes),
                              stem_conv_kernel_sizes=tuple(tuple(k) for k in stem_conv_kernel_sizes))
        if fuse:
            builder = MyFastToSlowFusionBuilder.build_fusion_builder(ratio, fusion_conv_kernel, fusion_conv_stride).create_module
            fusion_ratio = 2
        else:
            builder = lambda fusion_dim_in, stage_idx: nn.Identity()  # noqa: E731  (reference: nn.Identity as builder)
            fusion_ratio = 0
        blocks: List[nn.Module] = []
        stems = [_Stem(input_channels[p], stem_dim_outs[p], stem_conv_kernel_sizes[p]) for p in range(2)]
        blocks.append(_MultiPathWayWithFuse(stems, builder(fusion_dim_in=slow_c, stage_idx=0)))
        slow_ka = ((1, 1, 1), (1, 1, 1), (3, 1, 1), (3, 1, 1))
        dim_in, dim_out = slow_c, slow_c * 4
        for idx, depth in enumerate(depths):
            sstride = (1, 2, 2, 2)[idx]
            slow = _ResStage(depth, dim_in + dim_in * fusion_ratio // ratio, dim_out // 4, dim_out, slow_ka[idx], sstride)
            fast = _ResStage(depth, dim_in // ratio, dim_out // 4 // ratio, dim_out // ratio, (3, 1, 1), sstride)
            blocks.append(_MultiPathWayWithFuse([slow, fast], builder(fusion_dim_in=dim_out, stage_idx=idx + 1)))
            dim_in, dim_out = dim_out, dim_out * 2
        blocks.append(_PoolConcatPathway(head_pool_kernel_sizes))
        blocks.append(_Head(dim_in + dim_in // ratio, num_class))
        self.blocks = nn.ModuleList(blocks)
        _init_weights(self)
        self._engine = None

    # -- nn.Module API ---------------------------------------------------------------------------
    def forward(self, x: List[torch.Tensor]) -> torch.Tensor:
        from .engine import SlowFastEngine, run_model
        from .staging import RawClips
        raw = isinstance(x, RawClips)
        assert raw or (isinstance(x, (list, tuple)) and len(x) == 2), \
            "expects [slow, fast] NCTHW tensors (or staging.RawClips: uint8 frames, normalised on the device)"
        if self._engine is not None and self._engine.params_replaced():
            if self._engine.ddp is not None:
                raise RuntimeError("parameters of the model were replaced (load_state_dict(assign=True)?) after "
                                   "attach_ddp: call parallel.attach_ddp(model) again")
            self._engine = None
        if self._engine is None:
            self._engine = SlowFastEngine(self)
        return run_model(self._engine, x, None) if raw else run_model(self._engine, x[0], x[1])

    def _apply(self, fn, *a, **k):
        self._engine = None  # device / dtype moves invalidate packed weights and buffers
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        r = super().load_state_dict(*a, **k)
        if self._engine is not None:
            self._engine.invalidate_weights()
        return r


def init_my_slowfast(cfg, input_channels: Sequence[int], stem_dim_outs: Sequence[int], *,
                     fusion_conv_kernel=(3, 1, 1), fusion_conv_stride=(1, 1, 1),
                     head_pool_kernel_sizes=((4, 2, 2), (4, 2, 2)),
                     stem_conv_kernel_sizes=((1, 7, 7), (1, 7, 7))) -> nn.Module:
    """Drop-in for reference model/my_slowfast.py:44-126.  Reads cfg.MODEL.FUSE and cfg.CHALEARN.NUM_CLASS."""
    num_c = len(input_channels)
    assert num_c >= 2
    assert len(stem_dim_outs) == num_c
    fuse = cfg.MODEL.FUSE
    if fuse not in (True, False):
        raise ValueError("cfg.MODEL.FUSE must be a bool")
    return SlowFastB200(num_class=int(cfg.CHALEARN.NUM_CLASS), fuse=fuse, input_channels=input_channels,
                        stem_dim_outs=stem_dim_outs, fusion_conv_kernel=fusion_conv_kernel,
                        fusion_conv_stride=fusion_conv_stride, head_pool_kernel_sizes=head_pool_kernel_sizes,
                        stem_conv_kernel_sizes=stem_conv_kernel_sizes)
