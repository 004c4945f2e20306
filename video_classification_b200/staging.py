"""Raw clip staging: the uint8 side of the model's input boundary.

The reference normalises every frame on the host (dataset/chalearn_dataset.py:41-46 `transforms.ToTensor()` +
`transforms.Normalize([0.45]*21, [0.225]*21)`, applied per frame at :113-118), augments the fp32 clip with one
`RandomCrop(size, size // 10)` window (:79-88) and ships fp32 (N, T, 21, H, W) to the device, where
`_prepare_slowfast_data` (train.py:125-145) permutes it to NCTHW and splits channels 0:5 (BGR+UV) / 5:20 (flow).  Here
the loader keeps the frames as cv2 produced them -- uint8 HWC, 21 channels -- and `RawClips` carries them to the
model, whose first kernel (`sfb_stage_clips_u8`) does all of the above on the device: a quarter of the host->device
bytes, no fp32 clip in HBM, and bit-identical stem operands.
"""
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import torch


@dataclass
class RawClips:
    """uint8 clips as the dataset holds them before `self.preprocess`.

    frames  : uint8 CUDA tensor (N, T, H, W, Ct), contiguous (cv2 HWC frames stacked over time and batch)
    crop_yx : optional int32 CUDA tensor (N, 2): RandomCrop window offsets (i - padding, j - padding) per clip
    slow_t_stride : the Slow pathway takes frames 0, s, 2s, ... (1 for the reference graph, alpha for 8x8 SlowFast)
    """
    frames: torch.Tensor
    crop_yx: Optional[torch.Tensor] = None
    mean: float = 0.45
    std: float = 0.225
    slow_channels: Tuple[int, int] = (0, 5)    # train.py:139  x[:, 0:5]
    fast_channels: Tuple[int, int] = (5, 20)   # train.py:140  x[:, 5:20]
    slow_t_stride: int = 1

    def pathways(self) -> Sequence[Tuple[int, int, int]]:
        (s0, s1), (f0, f1) = self.slow_channels, self.fast_channels
        return ((s0, s1 - s0, self.slow_t_stride), (f0, f1 - f0, 1))

    def validate(self):
        f = self.frames
        if f.dtype != torch.uint8 or f.dim() != 5:
            raise TypeError("RawClips.frames must be a uint8 (N, T, H, W, C) tensor")
        if not f.is_contiguous():
            raise ValueError("RawClips.frames must be contiguous")
        if f.shape[1] % self.slow_t_stride:
            raise ValueError("clip length must be a multiple of slow_t_stride")
        if self.crop_yx is not None and (self.crop_yx.dtype != torch.int32 or tuple(self.crop_yx.shape) != (f.shape[0], 2)):
            raise ValueError("crop_yx must be int32 (N, 2)")


def random_crop_offsets(n: int, size: int, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """Per-clip offsets of `transforms.RandomCrop(size, padding=size // 10)` (dataset/chalearn_dataset.py:82-87):
    the window origin (i, j) is uniform in [0, 2*padding]; relative to the unpadded frame that is i - padding."""
    pad = size // 10
    ij = torch.randint(0, 2 * pad + 1, (n, 2), generator=generator, dtype=torch.int32)
    return ij - pad


def prepare_slowfast_data_u8(batch, key: str, *, augment: bool = False, slow_t_stride: int = 1, device="cuda"):
    """uint8 counterpart of `ModelManager._prepare_slowfast_data` (train.py:125-145): batch[key] is the uint8
    (N, T, H, W, 21) clip tensor (pinned host memory recommended), batch['label'] the labels."""
    frames = batch[key].to(device, non_blocking=True)
    crop = None
    if augment:
        crop = random_crop_offsets(frames.shape[0], frames.shape[2]).to(device, non_blocking=True)
    y_true = batch["label"].to(device, non_blocking=True)
    return RawClips(frames, crop, slow_t_stride=slow_t_stride), y_true
