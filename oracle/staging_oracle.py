"""ORACLE (test infrastructure, not product code; only tests/, __graft_entry__.smoke() and bench.py's CPU
baseline may import it).

CPU restatement of the reference's input staging, the uint8 side of the model boundary:

    to_tensor_normalize   restates  self.preprocess = Compose([ToTensor(), Normalize([0.45]*21, [0.225]*21)])
                                    reference dataset/chalearn_dataset.py:41-46, applied per frame at :113-118
                                    (torchvision semantics: HWC uint8 -> CHW, .to(float32).div(255);
                                     then tensor.sub_(mean).div_(std) with fp32 mean / std)
    random_crop           restates  transforms.RandomCrop(size, padding) on the (T, C, H, W) clip,
                                    reference dataset/chalearn_dataset.py:79-88: zero padding of the NORMALISED clip,
                                    one window (i, j) per clip
    prepare_slowfast_data restates  ModelManager._prepare_slowfast_data, reference train.py:136-140:
                                    permute NTCHW -> NCTHW, x[:, 0:5], x[:, 5:20]
    space_to_depth_bf16   restates  nothing in the reference: it is the layout contract of include/sfb200.h
                                    (sfb_layout_input_s2d), used to compare the staged operands bit for bit

Pinned in this container against torchvision's own transforms (the reference's dependency, 0.26 here) by
oracle/make_golden_staging.py -> tests/golden/staging_small.pt; tests/test_oracle.py checks the restatement
against that fixture bit for bit.  The reference ships no tests for this path: parity unpinned by reference tests.
"""
from typing import Optional, Sequence, Tuple

import torch

MEAN, STD = 0.45, 0.225


def to_tensor_normalize(frames_u8: torch.Tensor, mean: float = MEAN, std: float = STD) -> torch.Tensor:
    """uint8 (..., H, W, C) -> fp32 (..., C, H, W)"""
    x = frames_u8.movedim(-1, -3).contiguous().to(torch.float32).div(255)
    m = torch.as_tensor(mean, dtype=torch.float32)
    s = torch.as_tensor(std, dtype=torch.float32)
    return x.sub_(m).div_(s)


def random_crop(clip: torch.Tensor, i: int, j: int, padding: int) -> torch.Tensor:
    """clip (..., H, W) padded by `padding` zeros on every side, window of the original size at (i, j)"""
    H, W = clip.shape[-2:]
    padded = torch.nn.functional.pad(clip, (padding, padding, padding, padding), value=0.0)
    return padded[..., i:i + H, j:j + W]


def stage_clips(frames_u8: torch.Tensor, crop_ij: Optional[Sequence[Tuple[int, int]]] = None, padding: int = 0,
                mean: float = MEAN, std: float = STD) -> torch.Tensor:
    """uint8 (N, T, H, W, C) -> the fp32 (N, T, C, H, W) batch the reference loader yields"""
    x = to_tensor_normalize(frames_u8, mean, std)
    if crop_ij is not None:
        x = torch.stack([random_crop(x[n], int(i), int(j), padding) for n, (i, j) in enumerate(crop_ij)])
    return x


def prepare_slowfast_data(x_ntchw: torch.Tensor, slow=(0, 5), fast=(5, 20), slow_t_stride: int = 1):
    x = torch.permute(x_ntchw, [0, 2, 1, 3, 4])
    return [x[:, slow[0]:slow[1], ::slow_t_stride], x[:, fast[0]:fast[1]]]


def space_to_depth_bf16(x_ncthw: torch.Tensor, cs: int) -> torch.Tensor:
    """fp32 (N, C, T, H, W) -> bf16 (N, T, H/2, W/2, cs), channel (sh*2+sw)*C + c, zero padded"""
    N, C, T, H, W = x_ncthw.shape
    x = x_ncthw.reshape(N, C, T, H // 2, 2, W // 2, 2).permute(0, 2, 3, 5, 4, 6, 1).reshape(N, T, H // 2, W // 2, 4 * C)
    out = torch.zeros((N, T, H // 2, W // 2, cs), dtype=torch.bfloat16)
    out[..., :4 * C] = x.to(torch.bfloat16)
    return out
