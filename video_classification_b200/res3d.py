"""Slow-only 3D-ResNet-50 on the B200 kernels: BASELINE.json configs[0] (`config/res3d.yaml`).

The reference's res3d path (train.py:79-89, commented out there; live in (deprecated)/train_3dresnet.py:47-51) is

    model = torch.hub.load('facebookresearch/pytorchvideo', 'slow_r50', pretrained=True)
    model.blocks[0].conv = Conv3d(5, 64, (1, 7, 7), stride=(1, 2, 2), padding=(0, 3, 3), bias=False)

fed NCTHW clips (`_prepare_res3d_data`: permute NTCHW -> NCTHW).  `slow_r50(...)` below builds that module tree -- the
one pytorchvideo's `create_resnet` defaults give: `blocks.0` stem (conv / norm / pool), `blocks.1..4` ResStages of
depth (3, 4, 6, 3) with conv_a kernels (1,1,1),(1,1,1),(3,1,1),(3,1,1), `blocks.5` head (AvgPool3d, Dropout(0.5),
Linear(2048, num_classes), global average) -- with the same 320 state_dict keys, so the hub checkpoint loads with the
reference's own `load_state_dict(strict=False)` once the stem weight is dropped.  It is the Slow pathway of the SlowFast
graph without the lateral inputs; the engine runs it with the same kernels (engine.SlowFastEngine, npath = 1).

forward(x): one fp32 CUDA tensor (N, C, T, H, W), any strides -> (N, num_classes) logits that take part in autograd.
"""
from typing import Sequence

import torch
import torch.nn as nn

from .my_slowfast import _Head, _ParamTree, _ResStage, _Stem, _init_weights, _MODEL_STAGE_DEPTH


class _HeadWithPool(_Head):
    """create_res_basic_head with its own AvgPool3d (registration order of pytorchvideo: pool, dropout, proj, ...)"""

    def __init__(self, cin, num_class, pool_kernel, dropout_rate=0.5):
        super().__init__(cin, num_class, dropout_rate)
        self.pool = nn.AvgPool3d(kernel_size=tuple(pool_kernel), stride=(1, 1, 1), padding=(0, 0, 0))


class SlowOnlyB200(_ParamTree):
    def __init__(self, input_channel=3, model_num_class=400, head_pool_kernel_size=(4, 7, 7), model_depth=50,
                 stem_conv_kernel_size=(1, 7, 7)):
        super().__init__()
        depths = _MODEL_STAGE_DEPTH[model_depth]
        ka = ((1, 1, 1), (1, 1, 1), (3, 1, 1), (3, 1, 1))
        blocks = [_Stem(input_channel, 64, stem_conv_kernel_size)]
        dim_in, dim_out = 64, 256
        for idx, depth in enumerate(depths):
            blocks.append(_ResStage(depth, dim_in, dim_out // 4, dim_out, ka[idx], (1, 2, 2, 2)[idx]))
            dim_in, dim_out = dim_out, dim_out * 2
        blocks.append(_HeadWithPool(dim_in, model_num_class, head_pool_kernel_size))
        self.blocks = nn.ModuleList(blocks)
        self.graph_cfg = dict(npath=1, num_class=model_num_class, fuse=False, input_channels=(int(input_channel),),
                              stem_dim_outs=(64,), depths=depths,
                              head_pool_kernel_sizes=(tuple(head_pool_kernel_size),),
                              stem_conv_kernel_sizes=(tuple(stem_conv_kernel_size),))
        _init_weights(self)
        self._engine = None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        from .engine import SlowFastEngine, run_model
        assert torch.is_tensor(x) and x.dim() == 5, "expects one NCTHW tensor"
        if self._engine is not None and self._engine.params_replaced():
            if self._engine.ddp is not None:
                raise RuntimeError("parameters of the model were replaced (load_state_dict(assign=True)?) after "
                                   "attach_ddp: call parallel.attach_ddp(model) again")
            self._engine = None
        if self._engine is None:
            self._engine = SlowFastEngine(self)
        return run_model(self._engine, x, None)

    def _apply(self, fn, *a, **k):
        self._engine = None  # device / dtype moves invalidate packed weights and buffers
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        r = super().load_state_dict(*a, **k)
        if self._engine is not None:
            self._engine.invalidate_weights()
        return r


def slow_r50(input_channel: int = 5, model_num_class: int = 400, head_pool_kernel_size: Sequence[int] = (4, 7, 7)) -> nn.Module:
    """pytorchvideo `slow_r50` with the reference's 5-channel stem (train.py:81).  head_pool_kernel_size: (4,7,7) is the
    hub model's (224^2 clips, >= 4 frames); the 16 x 112^2 clips of BASELINE configs[0] end in a 4 x 4 map and need
    (4,4,4) or smaller -- nn.AvgPool3d in the reference fails the same way when the window exceeds the map."""
    return SlowOnlyB200(input_channel=input_channel, model_num_class=model_num_class,
                        head_pool_kernel_size=head_pool_kernel_size)
