"""The two configuration knobs of the reference that the SlowFast hot path reads, as a plain namespace.

The reference passes a yacs CfgNode (config/defaults.py) to init_my_slowfast, which reads exactly
`cfg.CHALEARN.NUM_CLASS` (config/defaults.py:14) and `cfg.MODEL.FUSE` (config/defaults.py:44); yacs is not part of
this image, and any object with those two attributes works (a real CfgNode included).  The per-part yaml files of the
reference (config/slowfast-*.yaml) only change batch size, crop folder and checkpoint name -- call-site settings,
listed here so benchmarks and tools can name the reference's own workloads.
"""
from types import SimpleNamespace

# config/crop_cfg.py:18-32 (crop folder -> resize) and the BATCH_SIZE of config/slowfast-<part>.yaml:2
PART_WORKLOADS = {
    "HTAH": dict(size=192, batch=55),
    "LHand": dict(size=64, batch=300),
    "RHand": dict(size=64, batch=300),
    "LHandArm": dict(size=128, batch=80),
    "RHandArm": dict(size=128, batch=80),
    "Torso": dict(size=128, batch=55),
}
CLIP_LEN = 20         # config/defaults.py CHALEARN.CLIP_LEN: both pathways see the same 20 frames (train.py:136-140)
NUM_CLASS = 249       # config/defaults.py:14


def make_cfg(num_class: int = NUM_CLASS, fuse: bool = True) -> SimpleNamespace:
    return SimpleNamespace(CHALEARN=SimpleNamespace(NUM_CLASS=num_class), MODEL=SimpleNamespace(FUSE=fuse))
