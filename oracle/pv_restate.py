"""ORACLE (test infrastructure, not product code).

CPU restatement of the third-party layer library the reference's hot path is built from:
`pytorchvideo` (facebookresearch/pytorchvideo, last PyPI release 0.1.5; NOT pinned by the reference, NOT
vendored under /root/reference, NOT installed here).  Only the symbols the reference imports at
model/my_slowfast.py:14,23-24,34-38 are restated, from the published algorithm of

    pytorchvideo/layers/utils.py   :: set_attributes
    pytorchvideo/models/net.py     :: Net, MultiPathWayWithFuse, DetectionBBoxNetwork (name only)
    pytorchvideo/models/stem.py    :: create_res_basic_stem, ResNetBasicStem
    pytorchvideo/models/resnet.py  :: create_bottleneck_block, create_res_block, create_res_stage,
                                      BottleneckBlock, ResBlock, ResStage
    pytorchvideo/models/head.py    :: create_res_basic_head, ResNetBasicHead
    pytorchvideo/models/slowfast.py:: create_slowfast, PoolConcatPathway, FastToSlowFusionBuilder, FuseFastToSlow
    pytorchvideo/models/weight_init.py :: init_net_weights

Everything is plain torch.nn, fp32.  Pins (see oracle/README.md): parameter count 34.57 M and
50.3 GMAC @ 8/32x224^2 for the canonical create_slowfast() defaults; 38 077 321 parameters / 734
state_dict entries when the reference's own model/my_slowfast.py is imported verbatim on top of it.
PARITY UNPINNED by the reference's own tests (it has none).
"""
import math
from typing import Callable, List, Optional, Tuple

import torch
import torch.nn as nn


# ----------------------------------------------------------------------------- layers/utils.py
def set_attributes(self, params=None):
    if params:
        for k, v in params.items():
            if k != "self":
                setattr(self, k, v)


# ----------------------------------------------------------------------------- models/weight_init.py
def _init_resnet_weights(model: nn.Module, fc_init_std: float = 0.01) -> nn.Module:
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.Conv3d)):
            # c2_msra_fill
            nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.modules.batchnorm._NormBase):
            if m.weight is not None:
                if hasattr(m, "block_final_bn") and m.block_final_bn:
                    m.weight.data.fill_(0.0)
                else:
                    m.weight.data.fill_(1.0)
            if m.bias is not None:
                m.bias.data.zero_()
        if isinstance(m, nn.Linear):
            m.weight.data.normal_(mean=0.0, std=fc_init_std)
            if m.bias is not None:
                m.bias.data.zero_()
    return model


def init_net_weights(model: nn.Module, init_std: float = 0.01, style: str = "resnet") -> None:
    assert style in ("resnet", "vit")
    if style == "resnet":
        _init_resnet_weights(model, init_std)
    else:  # pragma: no cover - not on the SlowFast path
        raise NotImplementedError


# ----------------------------------------------------------------------------- models/net.py
class Net(nn.Module):
    def __init__(self, *, blocks: nn.ModuleList) -> None:
        super().__init__()
        assert blocks is not None
        self.blocks = blocks
        init_net_weights(self)

    def forward(self, x):
        for block in self.blocks:
            x = block(x)
        return x


class DetectionBBoxNetwork(nn.Module):  # name only; never built on this path
    def __init__(self, model: nn.Module, detection_head: nn.Module):
        super().__init__()
        self.model = model
        self.detection_head = detection_head

    def forward(self, x, bboxes):
        feats = self.model(x)
        return self.detection_head(feats, bboxes).view(-1, feats.shape[1] if feats.dim() > 1 else 1)


class MultiPathWayWithFuse(nn.Module):
    def __init__(self, *, multipathway_blocks: nn.ModuleList, multipathway_fusion: Optional[nn.Module],
                 inplace: Optional[bool] = True) -> None:
        super().__init__()
        set_attributes(self, locals())

    def forward(self, x: List[torch.Tensor]) -> torch.Tensor:
        assert isinstance(x, list), "input for MultiPathWayWithFuse needs to be a list of tensors"
        x_out = x if self.inplace else [None] * len(x)
        for idx in range(len(self.multipathway_blocks)):
            if self.multipathway_blocks[idx] is not None:
                x_out[idx] = self.multipathway_blocks[idx](x[idx])
        if self.multipathway_fusion is not None:
            x_out = self.multipathway_fusion(x_out)
        return x_out


# ----------------------------------------------------------------------------- models/stem.py
class ResNetBasicStem(nn.Module):
    def __init__(self, *, conv: nn.Module = None, norm: nn.Module = None, activation: nn.Module = None,
                 pool: nn.Module = None) -> None:
        super().__init__()
        set_attributes(self, locals())
        assert self.conv is not None

    def forward(self, x):
        x = self.conv(x)
        if self.norm is not None:
            x = self.norm(x)
        if self.activation is not None:
            x = self.activation(x)
        if self.pool is not None:
            x = self.pool(x)
        return x


def create_res_basic_stem(*, in_channels: int, out_channels: int, conv_kernel_size=(3, 7, 7), conv_stride=(1, 2, 2),
                          conv_padding=(1, 3, 3), conv_bias: bool = False, conv: Callable = nn.Conv3d,
                          pool: Callable = nn.MaxPool3d, pool_kernel_size=(1, 3, 3), pool_stride=(1, 2, 2),
                          pool_padding=(0, 1, 1), norm: Callable = nn.BatchNorm3d, norm_eps: float = 1e-5,
                          norm_momentum: float = 0.1, activation: Callable = nn.ReLU) -> nn.Module:
    conv_module = conv(in_channels=in_channels, out_channels=out_channels, kernel_size=conv_kernel_size,
                       stride=conv_stride, padding=conv_padding, bias=conv_bias)
    norm_module = None if norm is None else norm(num_features=out_channels, eps=norm_eps, momentum=norm_momentum)
    activation_module = None if activation is None else activation()
    pool_module = None if pool is None else pool(kernel_size=pool_kernel_size, stride=pool_stride, padding=pool_padding)
    return ResNetBasicStem(conv=conv_module, norm=norm_module, activation=activation_module, pool=pool_module)


# ----------------------------------------------------------------------------- models/resnet.py
class BottleneckBlock(nn.Module):
    def __init__(self, *, conv_a=None, norm_a=None, act_a=None, conv_b=None, norm_b=None, act_b=None, conv_c=None,
                 norm_c=None) -> None:
        super().__init__()
        set_attributes(self, locals())
        assert all(op is not None for op in (self.conv_a, self.conv_b, self.conv_c))
        if self.norm_c is not None:
            self.norm_c.block_final_bn = True  # zero-initialised gamma (init_net_weights)

    def forward(self, x):
        x = self.conv_a(x)
        if self.norm_a is not None:
            x = self.norm_a(x)
        if self.act_a is not None:
            x = self.act_a(x)
        x = self.conv_b(x)
        if self.norm_b is not None:
            x = self.norm_b(x)
        if self.act_b is not None:
            x = self.act_b(x)
        x = self.conv_c(x)
        if self.norm_c is not None:
            x = self.norm_c(x)
        return x


def create_bottleneck_block(*, dim_in: int, dim_inner: int, dim_out: int, conv_a_kernel_size=(3, 1, 1),
                            conv_a_stride=(2, 1, 1), conv_a_padding=(1, 0, 0), conv_a: Callable = nn.Conv3d,
                            conv_b_kernel_size=(1, 3, 3), conv_b_stride=(1, 2, 2), conv_b_padding=(0, 1, 1),
                            conv_b_num_groups: int = 1, conv_b_dilation=(1, 1, 1), conv_b: Callable = nn.Conv3d,
                            conv_c: Callable = nn.Conv3d, norm: Callable = nn.BatchNorm3d, norm_eps: float = 1e-5,
                            norm_momentum: float = 0.1, activation: Callable = nn.ReLU) -> nn.Module:
    a = conv_a(in_channels=dim_in, out_channels=dim_inner, kernel_size=conv_a_kernel_size, stride=conv_a_stride,
               padding=conv_a_padding, bias=False)
    na = None if norm is None else norm(num_features=dim_inner, eps=norm_eps, momentum=norm_momentum)
    aa = None if activation is None else activation()
    b = conv_b(in_channels=dim_inner, out_channels=dim_inner, kernel_size=conv_b_kernel_size, stride=conv_b_stride,
               padding=conv_b_padding, bias=False, groups=conv_b_num_groups, dilation=conv_b_dilation)
    nb = None if norm is None else norm(num_features=dim_inner, eps=norm_eps, momentum=norm_momentum)
    ab = None if activation is None else activation()
    c = conv_c(in_channels=dim_inner, out_channels=dim_out, kernel_size=(1, 1, 1), bias=False)
    nc = None if norm is None else norm(num_features=dim_out, eps=norm_eps, momentum=norm_momentum)
    return BottleneckBlock(conv_a=a, norm_a=na, act_a=aa, conv_b=b, norm_b=nb, act_b=ab, conv_c=c, norm_c=nc)


class ResBlock(nn.Module):
    def __init__(self, branch1_conv=None, branch1_norm=None, branch2=None, activation=None,
                 branch_fusion: Callable = None) -> nn.Module:
        super().__init__()
        set_attributes(self, locals())
        assert self.branch2 is not None

    def forward(self, x):
        if self.branch1_conv is None:
            x = self.branch_fusion(x, self.branch2(x))
        else:
            shortcut = self.branch1_conv(x)
            if self.branch1_norm is not None:
                shortcut = self.branch1_norm(shortcut)
            x = self.branch_fusion(shortcut, self.branch2(x))
        if self.activation is not None:
            x = self.activation(x)
        return x


def create_res_block(*, dim_in: int, dim_inner: int, dim_out: int, bottleneck: Callable, use_shortcut: bool = False,
                     branch_fusion: Callable = lambda x, y: x + y, conv_a_kernel_size=(3, 1, 1),
                     conv_a_stride=(2, 1, 1), conv_a_padding=(1, 0, 0), conv_a: Callable = nn.Conv3d,
                     conv_b_kernel_size=(1, 3, 3), conv_b_stride=(1, 2, 2), conv_b_padding=(0, 1, 1),
                     conv_b_num_groups: int = 1, conv_b_dilation=(1, 1, 1), conv_b: Callable = nn.Conv3d,
                     conv_c: Callable = nn.Conv3d, conv_skip: Callable = nn.Conv3d, norm: Callable = nn.BatchNorm3d,
                     norm_eps: float = 1e-5, norm_momentum: float = 0.1, activation_bottleneck: Callable = nn.ReLU,
                     activation_block: Callable = nn.ReLU) -> nn.Module:
    branch1_conv_stride = tuple(x * y for x, y in zip(conv_a_stride, conv_b_stride))
    norm_model = None
    if use_shortcut or (norm is not None and (dim_in != dim_out or math.prod(branch1_conv_stride) != 1)):
        norm_model = norm(num_features=dim_out, eps=norm_eps, momentum=norm_momentum)
    need_skip = (dim_in != dim_out or math.prod(branch1_conv_stride) != 1) or use_shortcut
    return ResBlock(
        branch1_conv=conv_skip(dim_in, dim_out, kernel_size=(1, 1, 1), stride=branch1_conv_stride, bias=False)
        if need_skip else None,
        branch1_norm=norm_model,
        branch2=bottleneck(dim_in=dim_in, dim_inner=dim_inner, dim_out=dim_out, conv_a_kernel_size=conv_a_kernel_size,
                           conv_a_stride=conv_a_stride, conv_a_padding=conv_a_padding, conv_a=conv_a,
                           conv_b_kernel_size=conv_b_kernel_size, conv_b_stride=conv_b_stride,
                           conv_b_padding=conv_b_padding, conv_b_num_groups=conv_b_num_groups,
                           conv_b_dilation=conv_b_dilation, conv_b=conv_b, conv_c=conv_c, norm=norm, norm_eps=norm_eps,
                           norm_momentum=norm_momentum, activation=activation_bottleneck),
        activation=None if activation_block is None else activation_block(),
        branch_fusion=branch_fusion,
    )


class ResStage(nn.Module):
    def __init__(self, res_blocks: nn.ModuleList) -> nn.Module:
        super().__init__()
        self.res_blocks = res_blocks

    def forward(self, x):
        for blk in self.res_blocks:
            x = blk(x)
        return x


def create_res_stage(*, depth: int, dim_in: int, dim_inner: int, dim_out: int, bottleneck: Callable,
                     conv_a_kernel_size=(3, 1, 1), conv_a_stride=(2, 1, 1), conv_a_padding=(1, 0, 0),
                     conv_a: Callable = nn.Conv3d, conv_b_kernel_size=(1, 3, 3), conv_b_stride=(1, 2, 2),
                     conv_b_padding=(0, 1, 1), conv_b_num_groups: int = 1, conv_b_dilation=(1, 1, 1),
                     conv_b: Callable = nn.Conv3d, conv_c: Callable = nn.Conv3d, norm: Callable = nn.BatchNorm3d,
                     norm_eps: float = 1e-5, norm_momentum: float = 0.1, activation: Callable = nn.ReLU) -> nn.Module:
    if isinstance(conv_a_kernel_size[0], int):
        conv_a_kernel_size = [conv_a_kernel_size]
    if isinstance(conv_a_padding[0], int):
        conv_a_padding = [conv_a_padding]
    conv_a_kernel_size = (list(conv_a_kernel_size) * depth)[:depth]
    conv_a_padding = (list(conv_a_padding) * depth)[:depth]
    blocks = []
    for idx in range(depth):
        blocks.append(create_res_block(
            dim_in=dim_in if idx == 0 else dim_out, dim_inner=dim_inner, dim_out=dim_out, bottleneck=bottleneck,
            conv_a_kernel_size=conv_a_kernel_size[idx], conv_a_stride=conv_a_stride if idx == 0 else (1, 1, 1),
            conv_a_padding=conv_a_padding[idx], conv_a=conv_a, conv_b_kernel_size=conv_b_kernel_size,
            conv_b_stride=conv_b_stride if idx == 0 else (1, 1, 1), conv_b_padding=conv_b_padding,
            conv_b_num_groups=conv_b_num_groups, conv_b_dilation=conv_b_dilation, conv_b=conv_b, conv_c=conv_c,
            norm=norm, norm_eps=norm_eps, norm_momentum=norm_momentum, activation_bottleneck=activation,
            activation_block=activation))
    return ResStage(res_blocks=nn.ModuleList(blocks))


# ----------------------------------------------------------------------------- models/head.py
class ResNetBasicHead(nn.Module):
    def __init__(self, pool=None, dropout=None, proj=None, activation=None, output_pool=None) -> None:
        super().__init__()
        set_attributes(self, locals())
        assert self.proj is not None

    def forward(self, x):
        if self.pool is not None:
            x = self.pool(x)
        if self.dropout is not None:
            x = self.dropout(x)
        if self.proj is not None:
            x = x.permute((0, 2, 3, 4, 1))
            x = self.proj(x)
            x = x.permute((0, 4, 1, 2, 3))
        if self.activation is not None:
            x = self.activation(x)
        if self.output_pool is not None:
            x = self.output_pool(x)
            x = x.view(x.shape[0], -1)
        return x


def create_res_basic_head(*, in_features: int, out_features: int, pool: Callable = nn.AvgPool3d,
                          output_size=(1, 1, 1), pool_kernel_size=(1, 7, 7), pool_stride=(1, 1, 1),
                          pool_padding=(0, 0, 0), dropout_rate: float = 0.5, activation: Callable = None,
                          output_with_global_average: bool = True) -> nn.Module:
    if activation is None:
        activation_model = None
    elif activation == nn.Softmax:
        activation_model = activation(dim=1)
    else:
        activation_model = activation()
    if pool is None:
        pool_model = None
    elif pool == nn.AdaptiveAvgPool3d:
        pool_model = pool(output_size)
    else:
        pool_model = pool(kernel_size=pool_kernel_size, stride=pool_stride, padding=pool_padding)
    output_pool = nn.AdaptiveAvgPool3d(1) if output_with_global_average else None
    return ResNetBasicHead(proj=nn.Linear(in_features, out_features), activation=activation_model, pool=pool_model,
                           dropout=nn.Dropout(dropout_rate) if dropout_rate > 0 else None, output_pool=output_pool)


def create_res_roi_pooling_head(**kwargs):  # name only (detection heads are not on this path)
    raise NotImplementedError("create_res_roi_pooling_head is not on the SlowFast classification path")


# ----------------------------------------------------------------------------- models/slowfast.py
_MODEL_STAGE_DEPTH = {18: (1, 1, 1, 1), 50: (3, 4, 6, 3), 101: (3, 4, 23, 3), 152: (3, 8, 36, 3)}


class PoolConcatPathway(nn.Module):
    def __init__(self, retain_list: bool = False, pool: Optional[nn.ModuleList] = None, dim: int = 1) -> None:
        super().__init__()
        set_attributes(self, locals())

    def forward(self, x: List[torch.Tensor]) -> torch.Tensor:
        if self.pool is not None:
            assert len(x) == len(self.pool)
        output = []
        for ind in range(len(x)):
            if x[ind] is not None:
                if self.pool is not None and self.pool[ind] is not None:
                    x[ind] = self.pool[ind](x[ind])
                output.append(x[ind])
        if self.retain_list:
            return output
        return torch.cat(output, 1)


class FuseFastToSlow(nn.Module):
    def __init__(self, conv_fast_to_slow: nn.Module, norm: Optional[nn.Module] = None,
                 activation: Optional[nn.Module] = None) -> None:
        super().__init__()
        set_attributes(self, locals())

    def forward(self, x):
        x_s, x_f = x[0], x[1]
        fuse = self.conv_fast_to_slow(x_f)
        if self.norm is not None:
            fuse = self.norm(fuse)
        if self.activation is not None:
            fuse = self.activation(fuse)
        return [torch.cat([x_s, fuse], 1), x_f]


class FastToSlowFusionBuilder:
    def __init__(self, slowfast_channel_reduction_ratio: int, conv_fusion_channel_ratio: float,
                 conv_kernel_size: Tuple[int], conv_stride: Tuple[int], norm: Callable = nn.BatchNorm3d,
                 norm_eps: float = 1e-5, norm_momentum: float = 0.1, activation: Callable = nn.ReLU,
                 max_stage_idx: int = 3) -> None:
        set_attributes(self, locals())

    def create_module(self, fusion_dim_in: int, stage_idx: int) -> nn.Module:
        if stage_idx > self.max_stage_idx:
            return nn.Identity()
        conv_dim_in = fusion_dim_in // self.slowfast_channel_reduction_ratio
        conv = nn.Conv3d(conv_dim_in, int(conv_dim_in * self.conv_fusion_channel_ratio),
                         kernel_size=self.conv_kernel_size, stride=self.conv_stride,
                         padding=[k // 2 for k in self.conv_kernel_size], bias=False)
        norm_module = None if self.norm is None else self.norm(
            num_features=conv_dim_in * self.conv_fusion_channel_ratio, eps=self.norm_eps, momentum=self.norm_momentum)
        activation_module = None if self.activation is None else self.activation()
        return FuseFastToSlow(conv_fast_to_slow=conv, norm=norm_module, activation=activation_module)


def create_slowfast(
    *,
    slowfast_channel_reduction_ratio=(8,), slowfast_conv_channel_fusion_ratio: int = 2,
    slowfast_fusion_conv_kernel_size=(7, 1, 1), slowfast_fusion_conv_stride=(4, 1, 1), fusion_builder=None,
    input_channels=(3, 3), model_depth: int = 50, model_num_class: int = 400, dropout_rate: float = 0.5,
    norm: Callable = nn.BatchNorm3d, activation: Callable = nn.ReLU,
    stem_function=(create_res_basic_stem, create_res_basic_stem), stem_dim_outs=(64, 8),
    stem_conv_kernel_sizes=((1, 7, 7), (5, 7, 7)), stem_conv_strides=((1, 2, 2), (1, 2, 2)),
    stem_pool=(nn.MaxPool3d, nn.MaxPool3d), stem_pool_kernel_sizes=((1, 3, 3), (1, 3, 3)),
    stem_pool_strides=((1, 2, 2), (1, 2, 2)),
    stage_conv_a_kernel_sizes=(((1, 1, 1), (1, 1, 1), (3, 1, 1), (3, 1, 1)), ((3, 1, 1), (3, 1, 1), (3, 1, 1), (3, 1, 1))),
    stage_conv_b_kernel_sizes=(((1, 3, 3), (1, 3, 3), (1, 3, 3), (1, 3, 3)), ((1, 3, 3), (1, 3, 3), (1, 3, 3), (1, 3, 3))),
    stage_conv_b_num_groups=((1, 1, 1, 1), (1, 1, 1, 1)),
    stage_conv_b_dilations=(((1, 1, 1), (1, 1, 1), (1, 1, 1), (1, 1, 1)), ((1, 1, 1), (1, 1, 1), (1, 1, 1), (1, 1, 1))),
    stage_spatial_strides=((1, 2, 2, 2), (1, 2, 2, 2)), stage_temporal_strides=((1, 1, 1, 1), (1, 1, 1, 1)),
    bottleneck=((create_bottleneck_block,) * 4, (create_bottleneck_block,) * 4),
    head=create_res_basic_head, head_pool: Callable = nn.AvgPool3d, head_pool_kernel_sizes=((8, 7, 7), (32, 7, 7)),
    head_output_size=(1, 1, 1), head_activation: Callable = None, head_output_with_global_average: bool = True,
) -> nn.Module:
    num_pathway = len(input_channels)
    assert model_depth in _MODEL_STAGE_DEPTH
    stage_depths = _MODEL_STAGE_DEPTH[model_depth]
    if isinstance(slowfast_channel_reduction_ratio, int):
        slowfast_channel_reduction_ratio = (slowfast_channel_reduction_ratio,)
    if isinstance(slowfast_fusion_conv_kernel_size[0], int):
        slowfast_fusion_conv_kernel_size = (slowfast_fusion_conv_kernel_size,)
    if isinstance(slowfast_fusion_conv_stride[0], int):
        slowfast_fusion_conv_stride = (slowfast_fusion_conv_stride,)
    if fusion_builder is None:
        fusion_builder = FastToSlowFusionBuilder(
            slowfast_channel_reduction_ratio=slowfast_channel_reduction_ratio[0],
            conv_fusion_channel_ratio=slowfast_conv_channel_fusion_ratio,
            conv_kernel_size=slowfast_fusion_conv_kernel_size[0], conv_stride=slowfast_fusion_conv_stride[0],
            norm=norm, activation=activation, max_stage_idx=len(stage_depths) - 1).create_module

    stems = []
    for p in range(num_pathway):
        stems.append(stem_function[p](
            in_channels=input_channels[p], out_channels=stem_dim_outs[p], conv_kernel_size=stem_conv_kernel_sizes[p],
            conv_stride=stem_conv_strides[p], conv_padding=[s // 2 for s in stem_conv_kernel_sizes[p]],
            pool=stem_pool[p], pool_kernel_size=stem_pool_kernel_sizes[p], pool_stride=stem_pool_strides[p],
            pool_padding=[s // 2 for s in stem_pool_kernel_sizes[p]], norm=norm, activation=activation))
    stages = [MultiPathWayWithFuse(multipathway_blocks=nn.ModuleList(stems),
                                   multipathway_fusion=fusion_builder(fusion_dim_in=stem_dim_outs[0], stage_idx=0))]

    stage_dim_in = stem_dim_outs[0]
    stage_dim_out = stage_dim_in * 4
    for idx in range(len(stage_depths)):
        pathway_stage_dim_in = [stage_dim_in + stage_dim_in * slowfast_conv_channel_fusion_ratio
                                // slowfast_channel_reduction_ratio[0]]
        pathway_stage_dim_inner = [stage_dim_out // 4]
        pathway_stage_dim_out = [stage_dim_out]
        for reduction_ratio in slowfast_channel_reduction_ratio:
            pathway_stage_dim_in.append(stage_dim_in // reduction_ratio)
            pathway_stage_dim_inner.append(stage_dim_out // 4 // reduction_ratio)
            pathway_stage_dim_out.append(stage_dim_out // reduction_ratio)
        stage = []
        for p in range(num_pathway):
            depth = stage_depths[idx]
            stage_conv_a_stride = (stage_temporal_strides[p][idx], 1, 1)
            stage_conv_b_stride = (1, stage_spatial_strides[p][idx], stage_spatial_strides[p][idx])
            ka = stage_conv_a_kernel_sizes[p][idx]
            kb = stage_conv_b_kernel_sizes[p][idx]
            dil = stage_conv_b_dilations[p][idx]
            stage.append(create_res_stage(
                depth=depth, dim_in=pathway_stage_dim_in[p], dim_inner=pathway_stage_dim_inner[p],
                dim_out=pathway_stage_dim_out[p], bottleneck=bottleneck[p][idx], conv_a_kernel_size=ka,
                conv_a_stride=stage_conv_a_stride, conv_a_padding=[s // 2 for s in ka], conv_b_kernel_size=kb,
                conv_b_stride=stage_conv_b_stride,
                conv_b_padding=(kb[0] // 2, dil[1] if dil[1] > 1 else kb[1] // 2, dil[2] if dil[2] > 1 else kb[2] // 2),
                conv_b_num_groups=stage_conv_b_num_groups[p][idx], conv_b_dilation=dil, norm=norm,
                activation=activation))
        stages.append(MultiPathWayWithFuse(multipathway_blocks=nn.ModuleList(stage),
                                           multipathway_fusion=fusion_builder(fusion_dim_in=stage_dim_out,
                                                                              stage_idx=idx + 1)))
        stage_dim_in = stage_dim_out
        stage_dim_out = stage_dim_out * 2

    if head_pool is None:
        pool_model = None
    elif head_pool == nn.AdaptiveAvgPool3d:
        pool_model = [head_pool(head_output_size[p]) for p in range(num_pathway)]
    elif head_pool == nn.AvgPool3d:
        pool_model = [head_pool(kernel_size=head_pool_kernel_sizes[p], stride=(1, 1, 1), padding=(0, 0, 0))
                      for p in range(num_pathway)]
    else:  # pragma: no cover
        raise NotImplementedError(f"Unsupported pool_model type {pool_model}")
    stages.append(PoolConcatPathway(retain_list=False, pool=nn.ModuleList(pool_model)))
    head_in_features = stage_dim_in
    for reduction_ratio in slowfast_channel_reduction_ratio:
        head_in_features = head_in_features + stage_dim_in // reduction_ratio
    if head is not None:
        stages.append(head(in_features=head_in_features, out_features=model_num_class, pool=None,
                           output_size=head_output_size, dropout_rate=dropout_rate, activation=head_activation,
                           output_with_global_average=head_output_with_global_average))
    return Net(blocks=nn.ModuleList(stages))


# ----------------------------------------------------------------------------- models/resnet.py::create_resnet
def create_resnet(
    *,
    input_channel: int = 3, model_depth: int = 50, model_num_class: int = 400, dropout_rate: float = 0.5,
    norm: Callable = nn.BatchNorm3d, activation: Callable = nn.ReLU,
    stem_dim_out: int = 64, stem_conv_kernel_size=(1, 7, 7), stem_conv_stride=(1, 2, 2), stem_pool: Callable = nn.MaxPool3d,
    stem_pool_kernel_size=(1, 3, 3), stem_pool_stride=(1, 2, 2), stem: Callable = create_res_basic_stem,
    stage1_pool: Callable = None, stage1_pool_kernel_size=(2, 1, 1),
    stage_conv_a_kernel_size=((1, 1, 1), (1, 1, 1), (3, 1, 1), (3, 1, 1)),
    stage_conv_b_kernel_size=((1, 3, 3), (1, 3, 3), (1, 3, 3), (1, 3, 3)),
    stage_conv_b_num_groups=(1, 1, 1, 1), stage_conv_b_dilation=((1, 1, 1), (1, 1, 1), (1, 1, 1), (1, 1, 1)),
    stage_spatial_h_stride=(1, 2, 2, 2), stage_spatial_w_stride=(1, 2, 2, 2), stage_temporal_stride=(1, 1, 1, 1),
    bottleneck=(create_bottleneck_block,) * 4,
    head: Callable = create_res_basic_head, head_pool: Callable = nn.AvgPool3d, head_pool_kernel_size=(4, 7, 7),
    head_output_size=(1, 1, 1), head_activation: Callable = None, head_output_with_global_average: bool = True,
) -> nn.Module:
    """pytorchvideo 0.1.5 `create_resnet` (the builder behind the hub entry `slow_r50` that the reference's
    commented-out res3d path loads, reference train.py:79-83, (deprecated)/train_3dresnet.py:47-49): stem, four
    ResStages (dims 64 -> 256 -> 512 -> 1024 -> 2048, inner = out / 4), head with its own AvgPool3d."""
    assert model_depth in _MODEL_STAGE_DEPTH
    stage_depths = _MODEL_STAGE_DEPTH[model_depth]
    blocks = [stem(in_channels=input_channel, out_channels=stem_dim_out, conv_kernel_size=stem_conv_kernel_size,
                   conv_stride=stem_conv_stride, conv_padding=[s // 2 for s in stem_conv_kernel_size], pool=stem_pool,
                   pool_kernel_size=stem_pool_kernel_size, pool_stride=stem_pool_stride,
                   pool_padding=[s // 2 for s in stem_pool_kernel_size], norm=norm, activation=activation)]
    stage_dim_in = stem_dim_out
    stage_dim_out = stage_dim_in * 4
    for idx in range(len(stage_depths)):
        stage_dim_inner = stage_dim_out // 4
        ka, kb, dil = stage_conv_a_kernel_size[idx], stage_conv_b_kernel_size[idx], stage_conv_b_dilation[idx]
        blocks.append(create_res_stage(
            depth=stage_depths[idx], dim_in=stage_dim_in, dim_inner=stage_dim_inner, dim_out=stage_dim_out,
            bottleneck=bottleneck[idx], conv_a_kernel_size=ka, conv_a_stride=(stage_temporal_stride[idx], 1, 1),
            conv_a_padding=[s // 2 for s in ka], conv_b_kernel_size=kb,
            conv_b_stride=(1, stage_spatial_h_stride[idx], stage_spatial_w_stride[idx]),
            conv_b_padding=(kb[0] // 2, dil[1] if dil[1] > 1 else kb[1] // 2, dil[2] if dil[2] > 1 else kb[2] // 2),
            conv_b_num_groups=stage_conv_b_num_groups[idx], conv_b_dilation=dil, norm=norm, activation=activation))
        stage_dim_in = stage_dim_out
        stage_dim_out = stage_dim_out * 2
        if idx == 0 and stage1_pool is not None:
            blocks.append(stage1_pool(kernel_size=stage1_pool_kernel_size, stride=stage1_pool_kernel_size, padding=(0, 0, 0)))
    if head is not None:
        blocks.append(head(in_features=stage_dim_in, out_features=model_num_class, pool=head_pool,
                           output_size=head_output_size, pool_kernel_size=head_pool_kernel_size, dropout_rate=dropout_rate,
                           activation=head_activation, output_with_global_average=head_output_with_global_average))
    return Net(blocks=nn.ModuleList(blocks))
