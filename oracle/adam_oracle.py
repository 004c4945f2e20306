"""TEST INFRASTRUCTURE ONLY (imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg; never by the
product path).

CPU restatement of the optimizer step on the hot path's tail: the reference builds
`optim.Adam(self.model.parameters(), lr=cfg.MODEL.LR)` (ref train.py:182) and calls `zero_grad / backward / step`
per batch (ref train.py:229-231).  The algorithm lives in torch (a dependency of the reference, torch.optim.Adam,
single-tensor form `_single_tensor_adam`, amsgrad=False, maximize=False, capturable=False); this file restates it in
numpy float32, one correctly rounded operation per line, in the order torch applies them:

    g      = fma(param, weight_decay, grad)                 # grad.add(param, alpha=weight_decay), only if != 0
    m      = fma(g - m, 1 - beta1, m)                       # exp_avg.lerp_(grad, 1 - beta1), weight < 0.5 branch
    v      = fma((1 - beta2) * g, g, v * beta2)             # exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
    denom  = sqrt(v) / sqrt(1 - beta2^t) + eps
    param  = param + ((-lr / (1 - beta1^t)) * m) / denom    # param.addcdiv_(exp_avg, denom, value=-step_size)

fma = fused multiply-add with ONE rounding: that is what torch's vectorised CPU kernels (and its CUDA kernels) issue for
these three updates; every other operation is rounded on its own.  Found by comparing the variants against
torch.optim.Adam on 5x10^5 elements: this one is bit-identical on the vectorised body (the last numel % 8 elements of
a CPU tensor take torch's scalar tail, which differs by at most 1 ulp on `param`; the pin test states both).

The scalars (1 - beta1, step size, sqrt of the bias correction) are computed in Python doubles and rounded to fp32
once, as torch does when it hands a Python scalar to an fp32 kernel.  Pinned by tests/test_adam_cpu.py against
torch.optim.Adam itself (foreach=False, CPU): the reference's own call."""
import math

import numpy as np

F = np.float32


def fma(a, b, c):
    """fp32 fused multiply-add: the fp32 x fp32 product is exact in fp64, so one fp64 add + one rounding to fp32
    reproduces the single rounding (up to a 2^-29 double-rounding corner)."""
    return (np.asarray(a, F).astype(np.float64) * np.float64(F(b)) + np.asarray(c, F).astype(np.float64)).astype(F)


def fma_vv(a, b, c):
    return (np.asarray(a, F).astype(np.float64) * np.asarray(b, F).astype(np.float64)
            + np.asarray(c, F).astype(np.float64)).astype(F)


def adam_scalars(step: int, lr: float, beta1: float, beta2: float):
    """step = 1-based count of this update.  Returns (step_size, bias_correction2_sqrt) as Python doubles."""
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    return lr / bc1, math.sqrt(bc2)


def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
              grad_scale=1.0):
    """In-place update of the float32 numpy arrays param / exp_avg / exp_avg_sq from grad; `step` is the 1-based
    count of this update.  grad_scale multiplies the gradient first (1/world_size for a summed all-reduce)."""
    beta1, beta2 = betas
    step_size, bc2_sqrt = adam_scalars(step, lr, beta1, beta2)
    g = grad.astype(F, copy=True)
    if grad_scale != 1.0:
        g = g * F(grad_scale)
    if weight_decay != 0.0:
        g = fma(param, weight_decay, g)
    exp_avg[...] = fma(g - exp_avg, 1.0 - beta1, exp_avg)
    exp_avg_sq *= F(beta2)
    exp_avg_sq[...] = fma_vv(F(1.0 - beta2) * g, g, exp_avg_sq)
    denom = np.sqrt(exp_avg_sq) / F(bc2_sqrt) + F(eps)
    param += (F(-step_size) * exp_avg) / denom
    return param, exp_avg, exp_avg_sq
