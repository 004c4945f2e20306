"""ORACLE (test infrastructure).  Import the reference's own model/my_slowfast.py VERBATIM from
/root/reference, on top of the restated pytorchvideo (oracle/pv_restate.py) and the yacs / matplotlib shims.

Only usable where /root/reference exists (the authoring container); the GPU box uses the self-contained
restatement oracle/slowfast_oracle.py plus the committed fixtures under tests/golden/.

What has to be arranged for the import to succeed:
  * sys.path: oracle/shims (pytorchvideo, yacs, matplotlib stand-ins) and /root/reference
  * np.int (removed in numpy >= 1.24; used at model/my_slowfast.py:57)
  * cwd = <tmp>/work with <tmp>/cfg_override.yaml pointing CHALEARN.ROOT at <tmp>/data, and three one-line
    label files, because utils/chalearn.py:15-17 reads them at import time
"""
import os
import sys
import tempfile

REFERENCE_ROOT = "/root/reference"
_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(_HERE)
_cached = None


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "model", "my_slowfast.py"))


def load_reference():
    """Returns (my_slowfast module, get_cfg function) of the unmodified reference."""
    global _cached
    if _cached is not None:
        return _cached
    if not available():
        raise RuntimeError("/root/reference is not present on this machine")
    import numpy as np
    if not hasattr(np, "int"):
        np.int = int  # noqa: NPY001 - reference model/my_slowfast.py:57
    for p in (_REPO, os.path.join(_HERE, "shims"), REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    tmp = tempfile.mkdtemp(prefix="sfb_ref_")
    os.makedirs(os.path.join(tmp, "work"))
    os.makedirs(os.path.join(tmp, "data", "1_Sample"))
    for name in ("train", "test", "valid"):
        with open(os.path.join(tmp, "data", "1_Sample", name + ".txt"), "w") as f:
            f.write("train/001/M_00001.avi train/001/K_00001.avi 1\n")
    with open(os.path.join(tmp, "cfg_override.yaml"), "w") as f:
        f.write("CHALEARN:\n  ROOT: %s\n" % os.path.join(tmp, "data"))
    cwd = os.getcwd()
    os.chdir(os.path.join(tmp, "work"))
    try:
        # the reference has top-level packages named `model`, `config`, `dataset`, `utils`
        for name in ("model", "config", "dataset", "utils"):
            if name in sys.modules and not getattr(sys.modules[name], "__file__", "").startswith(REFERENCE_ROOT):
                del sys.modules[name]
        from config.defaults import get_cfg  # type: ignore
        import model.my_slowfast as ms  # type: ignore
    finally:
        os.chdir(cwd)
    _cached = (ms, get_cfg)
    return _cached


def build_reference_model(num_class=249, fuse=True, input_channels=(5, 15), stem_dim_outs=(64, 8), variant="REF"):
    """variant REF: exactly as written.  HYB: lateral stride (4,1,1) + head pools (8,7,7)/(32,7,7) via test-side
    monkeypatches only (SURVEY.md C.3 vi) -- the reference file itself is untouched."""
    import torch.nn as nn
    ms, get_cfg = load_reference()
    cfg = get_cfg()
    cfg.CHALEARN.NUM_CLASS = num_class
    cfg.MODEL.FUSE = fuse
    if variant == "REF":
        return ms.init_my_slowfast(cfg, input_channels, stem_dim_outs)
    assert variant == "HYB"
    orig_create, orig_builder = ms.create_slowfast, ms.MyFastToSlowFusionBuilder.__dict__["build_fusion_builder"]
    try:
        ms.create_slowfast = lambda **kw: orig_create(**{**kw, "head_pool_kernel_sizes": ((8, 7, 7), (32, 7, 7))})
        ms.MyFastToSlowFusionBuilder.build_fusion_builder = classmethod(
            lambda cls, r: ms.MyFastToSlowFusionBuilder(
                slowfast_channel_reduction_ratio=r, conv_fusion_channel_ratio=2, conv_kernel_size=(3, 1, 1),
                conv_stride=(4, 1, 1), norm=nn.BatchNorm3d, activation=nn.ReLU, max_stage_idx=3))
        return ms.init_my_slowfast(cfg, input_channels, stem_dim_outs)
    finally:
        ms.create_slowfast = orig_create
        ms.MyFastToSlowFusionBuilder.build_fusion_builder = orig_builder
