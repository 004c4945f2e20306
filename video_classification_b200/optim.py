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
        if betas[0] < 0.5:
            # torch's lerp switches formula at weight >= 0.5 (weight = 1 - beta1); the kernel reproduces the
            # weight < 0.5 branch only, which every usual beta1 (the reference keeps the default 0.9) takes
            raise NotImplementedError("beta1 < 0.5 is not implemented by sfb_adam_step (bit-exactness with torch's "
                                      "lerp is only established for beta1 >= 0.5)")
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
        self._plans: Dict[int, dict] = {}
        self._epoch = 0  # bumped whenever state tensors may have been replaced (load_state_dict, add_param_group)
        self.launches = 0

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._epoch += 1

    def add_param_group(self, param_group):
        super().add_param_group(param_group)
        self._epoch = getattr(self, "_epoch", 0) + 1

    def _init_state(self, p):
        st = self.state[p]
        if len(st) == 0:
            st["step"] = torch.tensor(0.0, dtype=torch.float32)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @staticmethod
    def _check(p, g, st):
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
        for k in ("exp_avg", "exp_avg_sq"):
            m = st[k]
            if m.device != p.device or m.dtype != torch.float32 or m.numel() != p.numel() or not m.is_contiguous():
                raise RuntimeError(f"optimizer state {k} does not match its parameter")

    def _plan_group(self, live: List[torch.nn.Parameter], sig: tuple) -> dict:
        """One table per (device, step count) class: a model trained from scratch has exactly one."""
        lib = _ffi.lib()
        classes: Dict[Tuple[torch.device, int], dict] = {}
        for p in live:
            if not p.is_cuda:
                self._check(p, p.grad, None)
            st = self._init_state(p)
            self._check(p, p.grad, st)
            if st["step"].is_cuda:  # a checkpoint of a capturable / fused torch optimizer: keep torch's CPU layout
                st["step"] = st["step"].detach().to("cpu", torch.float32)
            c = classes.setdefault((p.device, int(st["step"])), dict(rows=[], steps=[], touched=[]))
            c["rows"].append((p.data_ptr(), p.grad.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(),
                              p.numel()))
            c["steps"].append(st["step"])
            c["touched"] += [p, st["exp_avg"], st["exp_avg_sq"]]
        out = []
        for (dev, t), c in classes.items():
            rows = c["rows"]
            arr = (_ffi.AdamJob * len(rows))()
            for a, (pp, gp, mp, vp, n) in zip(arr, rows):
                a.param, a.grad, a.exp_avg, a.exp_avg_sq, a.n = pp, gp, mp, vp, n
            nbytes = int(lib.sfb_adam_table_bytes(len(rows)))
            tab = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            nblk = C.c_int32(0)
            with torch.cuda.device(dev):
                _ffi.check(lib.sfb_adam_plan(arr, len(rows), tab.data_ptr(), nbytes, C.byref(nblk)), "sfb_adam_plan")
            out.append(dict(dev=dev, t=t, tab=tab, n=len(rows), nblk=int(nblk.value), steps=c["steps"],
                            touched=c["touched"]))
        return dict(sig=sig, epoch=self._epoch, classes=out)

    def _launch(self, group, cl):
        beta1, beta2 = group["betas"]
        with torch.cuda.device(cl["dev"]):
            _ffi.check(_ffi.lib().sfb_adam_step(cl["tab"].data_ptr(), cl["n"], cl["nblk"], cl["t"], float(group["lr"]),
                                                float(beta1), float(beta2), float(group["eps"]),
                                                float(group["weight_decay"]), self.grad_scale,
                                                torch.cuda.current_stream(cl["dev"]).cuda_stream), "sfb_adam_step")
        # the kernel wrote through raw pointers: tell autograd these tensors changed in place, as torch's own in-place
        # update does (the engine re-packs its bf16 conv operands when a weight's version counter moves)
        torch.autograd.graph.increment_version(cl["touched"])
        self.launches += 1

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            if group.get("amsgrad") or group.get("maximize"):
                raise NotImplementedError("amsgrad / maximize are not part of the reference's configuration")
            live, ptrs = [], []
            for p in group["params"]:
                g = p.grad
                if g is None:  # the reference model's 48 unused parameters: skipped like torch does
                    continue
                live.append(p)
                ptrs.append(p.data_ptr())
                ptrs.append(g.data_ptr())
                st = self.state.get(p)
                if st:  # replaced / cleared state tensors must rebuild the table, not update orphans
                    ptrs.append(st["exp_avg"].data_ptr())
                    ptrs.append(st["exp_avg_sq"].data_ptr())
                else:
                    ptrs += [0, 0]
            if not live:
                continue
            sig = tuple(ptrs)
            plan = self._plans.get(gi)
            if plan is None or plan["epoch"] != self._epoch or plan["sig"] != sig:
                plan = self._plans[gi] = self._plan_group(live, sig)
                # state tensors created by the planning pass are part of the signature from now on
                plan["sig"] = tuple(v for p in live for v in (p.data_ptr(), p.grad.data_ptr(),
                                                              self.state[p]["exp_avg"].data_ptr(),
                                                              self.state[p]["exp_avg_sq"].data_ptr()))
            for cl in plan["classes"]:
                cl["t"] += 1
                torch._foreach_add_(cl["steps"], 1)
                self._launch(group, cl)
        return loss

    def replay_last(self):
        """Launch the tables of the last step() again with the same step count (measurement only: bench.py times the
        kernel alone with it; the parameters move as in a real step)."""
        for gi, group in enumerate(self.param_groups):
            for cl in self._plans.get(gi, {"classes": []})["classes"]:
                self._launch(group, cl)
