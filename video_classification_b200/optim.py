"""Optimizer step of the training loop on the device in ONE launch.

Drop-in for the reference's `optim.Adam(self.model.parameters(), lr=cfg.MODEL.LR)` (ref train.py:182) and its
`self.optim.zero_grad() / loss.backward() / self.optim.step()` use (ref train.py:229-231): same constructor
arguments, same `param_groups` / `state` / `state_dict()` layout as torch.optim.Adam (per parameter `step`,
`exp_avg`, `exp_avg_sq`), so checkpoints move both ways.  `step()` describes every parameter that has a gradient as
one row of a device-resident table (built once, rebuilt only when a pointer changes) and `sfb_adam_step` updates all
of them in a single kernel (`csrc/adam.cuh`), in torch's operation order and roundings (oracle/adam_oracle.py).
There is no fallback: CPU parameters, non-fp32 or non-contiguous tensors raise."""
import ctypes as C
from typing import Dict, List, Tuple

import torch

from . import _ffi


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, *,
                 maximize=False, grad_scale=1.0):
        # argument checks and messages of torch.optim.Adam
        if not 0.0 <= lr:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        if not 0.0 <= weight_decay:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if amsgrad or maximize:
            raise NotImplementedError("the B200 Adam step implements the reference's configuration "
                                      "(amsgrad=False, maximize=False; train.py:182)")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                        foreach=None, capturable=False, differentiable=False, fused=None,
                        decoupled_weight_decay=False)
        super().__init__(params, defaults)
        self.grad_scale = float(grad_scale)
        self._tables: Dict[Tuple[int, int], Tuple[tuple, torch.Tensor, int, int]] = {}
        self.launches = 0

    def _init_state(self, p):
        st = self.state[p]
        if len(st) == 0:
            st["step"] = torch.tensor(0.0, dtype=torch.float32)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @staticmethod
    def _check(p, g):
        if not p.is_cuda:
            raise RuntimeError("video_classification_b200.optim.Adam runs on the GPU only (no CPU fallback): "
                               "move the model to a CUDA device before building the optimizer state")
        if g.is_sparse:
            raise RuntimeError("Adam does not support sparse gradients, please consider SparseAdam instead")
        if p.dtype != torch.float32 or g.dtype != torch.float32:
            raise RuntimeError(f"fp32 master parameters and gradients expected, got {p.dtype} / {g.dtype}")
        if g.device != p.device or g.numel() != p.numel():
            raise RuntimeError("gradient does not match its parameter (device / element count)")
        if not p.is_contiguous() or not g.is_contiguous():
            raise RuntimeError("contiguous parameters and gradients expected")

    def _table(self, key, rows: List[tuple], dev):
        sig = tuple(rows)
        hit = self._tables.get(key)
        if hit is not None and hit[0] == sig:
            return hit
        lib = _ffi.lib()
        arr = (_ffi.AdamJob * len(rows))()
        for a, (pp, gp, mp, vp, n) in zip(arr, rows):
            a.param, a.grad, a.exp_avg, a.exp_avg_sq, a.n = pp, gp, mp, vp, n
        nbytes = int(lib.sfb_adam_table_bytes(len(rows)))
        tab = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        nblk = C.c_int32(0)
        with torch.cuda.device(dev):
            _ffi.check(lib.sfb_adam_plan(arr, len(rows), tab.data_ptr(), nbytes, C.byref(nblk)), "sfb_adam_plan")
        hit = (sig, tab, len(rows), int(nblk.value))
        self._tables[key] = hit
        return hit

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _ffi.lib()
        for gi, group in enumerate(self.param_groups):
            if group.get("amsgrad") or group.get("maximize"):
                raise NotImplementedError("amsgrad / maximize are not part of the reference's configuration")
            beta1, beta2 = group["betas"]
            # one launch per (device, step count); a model trained from scratch has exactly one such class
            classes: Dict[Tuple[torch.device, int], List[tuple]] = {}
            for p in group["params"]:
                g = p.grad
                if g is None:  # the reference model's 48 unused parameters: skipped like torch does
                    continue
                self._check(p, g)
                st = self._init_state(p)
                st["step"] += 1
                t = int(st["step"])
                classes.setdefault((p.device, t), []).append(
                    (p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel()))
            for ci, ((dev, t), rows) in enumerate(classes.items()):
                _, tab, n, nblk = self._table((gi, ci), rows, dev)
                with torch.cuda.device(dev):
                    _ffi.check(lib.sfb_adam_step(tab.data_ptr(), n, nblk, t, float(group["lr"]), float(beta1),
                                                 float(beta2), float(group["eps"]), float(group["weight_decay"]),
                                                 self.grad_scale, torch.cuda.current_stream(dev).cuda_stream),
                               "sfb_adam_step")
                self.launches += 1
        return loss
