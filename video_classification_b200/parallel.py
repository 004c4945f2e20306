"""Data-parallel training of the B200 SlowFast: one process per GPU, bucketed gradient all-reduce (NCCL over
NVLink/NVSwitch) overlapped with the rest of the backward pass.

The reference trains on a single GPU (train.py:225-231: forward, CrossEntropy, backward, Adam) and has no
distributed code; BASELINE.json's configs[2] asks for the same step data-parallel over 1/2/4/8 B200.  BatchNorm3d
stays per-replica exactly as in the reference (plain nn.BatchNorm3d, model/my_slowfast.py:143,254), so the only
exchange is the gradient sum.

Design:
  * the engine keeps every live gradient in ONE flat fp32 arena laid out in reverse-forward order (FC head first,
    stems last) -- the order in which backward finishes them -- so a bucket is a contiguous slice: no gather /
    scatter copies, one collective per bucket on the arena itself;
  * bucket boundaries are fixed when the backward plan is built; the op that completes the last parameter of a
    bucket is followed, in the static op list, by a launch entry: record an event on the compute stream, make the
    communication stream wait for it, all-reduce the slice there.  Later dgrad/wgrad kernels keep running;
  * parameters the reference never reaches from forward() (48 of them: `residual` / `res_unit`) are
    not in the arena, get no gradient and are skipped, like torch DDP's find_unused_parameters would;
  * finish: compute stream waits for the communication stream; the 1/world_size scaling is folded into the
    collective (ncclAvg) or applied once on the arena (gloo has no AVG).
"""
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


class GradBucketer:
    """Bucket planner + launcher over a flat gradient arena.  Pure host logic + torch.distributed; works on CPU
    tensors with gloo (tests) and on CUDA tensors with NCCL (product)."""

    def __init__(self, flat: torch.Tensor, spans: Sequence[Tuple[int, int]], bucket_bytes: int = 25 << 20,
                 group=None, tail_bytes: int = 0):
        """spans: (offset, numel) of each parameter's gradient inside `flat`, any order; they are bucketed in
        arena order (ascending offset = order in which backward completes them)."""
        self.flat = flat
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.spans = sorted(spans)
        cap = max(1, bucket_bytes // flat.element_size())
        self.buckets: List[Tuple[int, int]] = []  # (start, end) element offsets
        self.span_bucket = {}
        start, end, members = None, None, []
        for off, n in self.spans:
            if start is not None and (off + n - start) > cap and members:
                self.buckets.append((start, end))
                start, members = None, []
            if start is None:
                start = off
            end = off + n
            members.append(off)
            self.span_bucket[off] = len(self.buckets)
        if members:
            self.buckets.append((start, end))
        # The gradients at the END of the arena (stems, first residual stage) become final last, so the collective of
        # the bucket that holds them cannot start before backward is over and its whole duration is exposed.  Split a
        # small tail off the last bucket: what remains exposed is a latency-sized all-reduce.
        if tail_bytes > 0 and self.buckets:
            s0, e0 = self.buckets[-1]
            cut = None
            for off, n in self.spans:
                if off > s0 and (e0 - off) * flat.element_size() <= tail_bytes:
                    cut = off
                    break
            if cut is not None:
                self.buckets[-1] = (s0, cut)
                self.buckets.append((cut, e0))
                nb = len(self.buckets) - 1
                for off, n in self.spans:
                    if off >= cut:
                        self.span_bucket[off] = nb
        self.is_cuda = flat.is_cuda
        self.comm_stream = torch.cuda.Stream(device=flat.device) if self.is_cuda else None
        backend = dist.get_backend(group) if dist.is_initialized() else "none"
        self.use_avg = backend == "nccl"
        self._pending = None
        self._works = []
        self.launch_order: List[int] = []  # for tests / introspection
        # optional per-bucket timeline (bench.py --ddp-timeline): CUDA events at "gradients of the bucket final" on the
        # compute stream and around the collective on the communication stream
        self.timeline = None
        self.reset_counts()

    def reset_counts(self):
        self._pending = [0] * len(self.buckets)
        for off, _ in self.spans:
            self._pending[self.span_bucket[off]] += 1

    def mark_ready(self, offset: int) -> Optional[int]:
        """plan/record-time: the gradient at `offset` is final.  Returns the bucket index if that completed it."""
        b = self.span_bucket[offset]
        self._pending[b] -= 1
        if self._pending[b] == 0:
            return b
        assert self._pending[b] > 0, "gradient marked ready twice"
        return None

    def launch(self, b: int) -> int:
        """enqueue the all-reduce of bucket b behind everything already on the current stream"""
        self.launch_order.append(b)
        if self.world == 1 or not dist.is_initialized():
            return 0
        s, e = self.buckets[b]
        view = self.flat[s:e]
        op = dist.ReduceOp.AVG if self.use_avg else dist.ReduceOp.SUM
        if self.is_cuda:
            tl = self.timeline
            ev = torch.cuda.Event(enable_timing=tl is not None)
            ev.record(torch.cuda.current_stream(self.flat.device))
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                if tl is not None:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(self.comm_stream)
                dist.all_reduce(view, op=op, group=self.group)
                if tl is not None:
                    e1.record(self.comm_stream)
                    tl.append((b, (e - s) * self.flat.element_size(), ev, e0, e1))
        else:
            self._works.append(dist.all_reduce(view, op=op, group=self.group, async_op=True))
        return 0

    def begin(self):
        self.launch_order = []
        self._works = []

    def finish(self):
        """all buckets reduced and visible to the compute stream; gradients are the mean over ranks"""
        if self.world == 1 or not dist.is_initialized():
            return
        if self.is_cuda:
            torch.cuda.current_stream(self.flat.device).wait_stream(self.comm_stream)
        else:
            for w in self._works:
                w.wait()
        if not self.use_avg:
            s, e = self.buckets[0][0], self.buckets[-1][1]
            self.flat[s:e].mul_(1.0 / self.world)


class EngineDDP:
    """Glue between SlowFastEngine's static backward plan and a GradBucketer (engine.ddp protocol)."""

    def __init__(self, engine, bucket_mb: float = 25.0, group=None, tail_mb: float = 1.0):
        spans = [(engine.grad_offsets[id(p)], p.numel()) for p in engine.live_params]
        self.bucketer = GradBucketer(engine.grad_flat, spans, int(bucket_mb * (1 << 20)), group,
                                     tail_bytes=int(tail_mb * (1 << 20)))
        self.engine = engine

    # called while the backward op list is being built ------------------------------------------
    def begin_plan(self):
        self.bucketer.reset_counts()

    def note_grad_ready(self, plan, lst, param):
        b = self.bucketer.mark_ready(self.engine.grad_offsets[id(param)])
        if b is not None:
            bk = self.bucketer

            def launch(bk=bk, b=b):
                return bk.launch(b)
            launch.is_bucket_launch = True  # the engine cuts its backward CUDA graphs at these entries
            lst.append((launch, (), None))

    # called around every execution of the backward op list ----------------------------------------
    def begin_backward(self, engine):
        self.bucketer.begin()

    def finish_backward(self, engine):
        self.bucketer.finish()


def broadcast_parameters(model: torch.nn.Module, src: int = 0, group=None):
    """same start on every rank (parameters AND BatchNorm buffers), as torch DDP does at construction"""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    with torch.no_grad():
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t.data, src=src, group=group)


def attach_ddp(model, bucket_mb: float = 25.0, group=None, broadcast: bool = True, tail_mb: float = 1.0) -> EngineDDP:
    """model: my_slowfast.SlowFastB200 already on its CUDA device.  After this, loss.backward() leaves the MEAN
    gradient over all ranks in every live parameter's .grad; the training loop (train.py:225-231) is unchanged."""
    from .engine import SlowFastEngine
    if broadcast:
        broadcast_parameters(model, 0, group)
    if model._engine is None:
        model._engine = SlowFastEngine(model)
    eng = model._engine
    eng.plans.clear()  # backward plans must be rebuilt with the bucket launch entries in place
    eng.ddp = EngineDDP(eng, bucket_mb, group, tail_mb=tail_mb)
    if broadcast:
        eng.invalidate_weights()
    return eng.ddp


def gather_part_scores(clip_probs: torch.Tensor, group=None) -> torch.Tensor:
    """Ensemble placement of BASELINE config 4 (one body-part model per GPU; reference train_sparse.py:36-86 runs the
    parts one after another and pickles each part's scores): every rank holds the softmax scores (n_clips, C) of ITS
    part for the same clips; one all-gather returns (n_clips, P, C) with parts in rank order -- the layout
    SparseFusionDataset stacks from the pickles (train_sparse.py:109-131) and SparseModel.forward expects."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return clip_probs.unsqueeze(1).contiguous()
    world = dist.get_world_size(group)
    clip_probs = clip_probs.contiguous()
    out = torch.empty((world,) + tuple(clip_probs.shape), dtype=clip_probs.dtype, device=clip_probs.device)
    dist.all_gather([out[r] for r in range(world)], clip_probs, group=group)  # views of one buffer; gloo and nccl
    return out.permute(1, 0, 2).contiguous()
