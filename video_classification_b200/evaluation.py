"""After the network: the reference's evaluation drivers kept on the device.

    video_scores / EvalAccumulator   Trainer.run_eval's tail, reference train.py:287-370: the reference moves every
                                     batch's logits to the host (`y_pred.cpu().numpy()`, a sync per batch), then does the
                                     softmax, the per-video mean over `samples_per_video` clips and the arg-max in
                                     numpy loops.  Here the logits stay on the device and ONE kernel
                                     (sfb_video_scores) produces clip probabilities, video means and top-1.
    SparseModel                      reference train_sparse.py:89-105: num_class nn.Linear(num_part, 1) modules looped
                                     in Python; same parameters / state_dict keys (`fcs.{c}.weight`, `fcs.{c}.bias`),
                                     forward + backward as one kernel each (sfb_sparse_fuse, sfb_sparse_fuse_bwd).
    fuse_ensemble                    SparseFusionTrainer.test, train_sparse.py:200-238: fuse the part scores, mean per
                                     video, arg-max, accuracy.
"""
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn as nn

from . import _ffi
from ._ffi import check, lib, stream_ptr


def _segments(samples_per_video: Sequence[int], device) -> torch.Tensor:
    sv = torch.as_tensor(list(samples_per_video), dtype=torch.int64)
    if sv.numel() == 0 or int(sv.min()) < 1:
        # a video without clips has no mean / arg-max (the reference's run_eval never produces one: every video
        # yields at least one sliding window, train.py:337-349)
        raise ValueError("samples_per_video must be a non-empty list of positive counts")
    seg = torch.zeros(sv.numel() + 1, dtype=torch.int64)
    seg[1:] = torch.cumsum(sv, 0)
    return seg.to(torch.int32).to(device)


def video_scores(scores: torch.Tensor, samples_per_video: Sequence[int], softmax: bool = True) -> Dict[str, torch.Tensor]:
    """scores: fp32 CUDA (n_clips, C) logits (softmax=True) or fused scores (softmax=False), clips of one video
    contiguous in the order of `samples_per_video`.  Returns clip_probs (n_clips, C), video_mean (V, C), top1 (V,)."""
    if not scores.is_cuda:
        raise _ffi.SfbError("video_scores is CUDA only (no CPU fallback)")
    scores = scores.contiguous().float()
    n, C = scores.shape
    seg = _segments(samples_per_video, scores.device)
    V = seg.numel() - 1
    if int(seg[-1]) != n:
        raise ValueError(f"samples_per_video sums to {int(seg[-1])}, but there are {n} clips")
    probs = torch.empty_like(scores)
    mean = torch.empty((V, C), dtype=torch.float32, device=scores.device)
    top1 = torch.empty((V,), dtype=torch.int32, device=scores.device)
    check(lib().sfb_video_scores(scores.data_ptr(), n, C, seg.data_ptr(), V, 1 if softmax else 0, probs.data_ptr(),
                                 mean.data_ptr(), top1.data_ptr(), stream_ptr()), "sfb_video_scores")
    return dict(clip_probs=probs, video_mean=mean, top1=top1)


class EvalAccumulator:
    """Device-side replacement of the lists `run_eval` keeps (train.py:294-309): add() per batch without a host sync,
    result() once at the end, in the reference's return format {'ps', 't', 'acc', 'sv'} (numpy)."""

    def __init__(self):
        self.logits: List[torch.Tensor] = []
        self.labels: List[torch.Tensor] = []
        self.samples_per_video: List[int] = []

    def add(self, logits: torch.Tensor, labels: torch.Tensor, samples_per_video: Optional[Sequence[int]] = None):
        self.logits.append(logits.detach())
        self.labels.append(labels.detach())
        if samples_per_video is not None:
            self.samples_per_video.extend(int(s) for s in samples_per_video)

    def result(self, softmax: bool = True):
        scores = torch.cat(self.logits, 0)
        labels = torch.cat(self.labels, 0).to(scores.device)
        r = video_scores(scores, self.samples_per_video, softmax)
        seg = _segments(self.samples_per_video, scores.device).long()
        first = labels[seg[:-1]]
        # the reference asserts that all clips of a video carry the same label (train.py:349)
        per_clip_video = torch.repeat_interleave(torch.arange(seg.numel() - 1, device=scores.device), seg[1:] - seg[:-1])
        if not bool((labels == first[per_clip_video]).all()):
            raise AssertionError("clips of one video carry different labels")
        correct = r["top1"].long() == first.long()
        return {"ps": r["clip_probs"].cpu().numpy(), "t": labels.cpu().numpy(),
                "acc": float(correct.sum()) / correct.numel(), "sv": list(self.samples_per_video),
                "video_top1": r["top1"].cpu().numpy(), "video_mean": r["video_mean"].cpu().numpy()}


class _SparseFuseFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        N, P, C = x.shape
        y = torch.empty((N, C), dtype=torch.float32, device=x.device)
        check(lib().sfb_sparse_fuse(x.data_ptr(), w.data_ptr(), b.data_ptr(), N, P, C, y.data_ptr(), stream_ptr()),
              "sfb_sparse_fuse")
        ctx.save_for_backward(x, w)
        ctx.need_dx = x.requires_grad
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        N, P, C = x.shape
        dy = dy.contiguous().float()
        dw = torch.empty_like(w)
        db = torch.empty((C,), dtype=torch.float32, device=x.device)
        dx = torch.empty_like(x) if ctx.need_dx else None
        check(lib().sfb_sparse_fuse_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), N, P, C, dw.data_ptr(), db.data_ptr(),
                                        None if dx is None else dx.data_ptr(), stream_ptr()), "sfb_sparse_fuse_bwd")
        return dx, dw, db


class SparseModel(nn.Module):
    """Drop-in for reference train_sparse.py:89-105 (same constructor, parameters and state_dict keys)."""

    def __init__(self, num_class, num_part) -> None:
        super().__init__()
        self.num_class = num_class
        self.num_part = num_part
        self.fcs = nn.ModuleList([nn.Linear(num_part, 1) for _ in range(num_class)])

    def forward(self, x):
        assert x.size()[1:3] == (self.num_part, self.num_class)
        if not x.is_cuda:
            raise _ffi.SfbError("SparseModel is CUDA only (no CPU fallback)")
        # (C, P) / (C,) views of the 2*C leaf parameters; autograd routes dw / db back to every fcs[c]
        w = torch.cat([fc.weight for fc in self.fcs], 0)
        b = torch.cat([fc.bias for fc in self.fcs], 0)
        return _SparseFuseFn.apply(x.contiguous().float(), w.contiguous(), b.contiguous())


def fuse_ensemble(sparse_model: SparseModel, part_scores: torch.Tensor, labels: torch.Tensor,
                  samples_per_video: Sequence[int]):
    """part_scores: (n_clips, P, C) softmax scores of the P part models (the 'ps' arrays the reference pickles,
    stacked as SparseFusionDataset does).  Mirrors SparseFusionTrainer.test (train_sparse.py:200-238)."""
    with torch.no_grad():
        sparse_model.eval()
        fused = sparse_model(part_scores)
    acc = EvalAccumulator()
    acc.add(fused, labels, samples_per_video)
    return acc.result(softmax=False)
