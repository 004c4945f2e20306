"""CPU tests of the optimizer-step row: the numpy restatement is pinned against torch.optim.Adam
itself -- the call the reference makes (train.py:182, 231) -- and the C ABI validates its arguments without a GPU."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import adam_oracle as ao
from video_classification_b200 import _ffi
from video_classification_b200 import build as sfb_build


def _ulps(a, b):
    return np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64))


@pytest.mark.parametrize("wd", [0.0, 1e-2])
@pytest.mark.parametrize("lr", [1e-4, 1e-3])
def test_oracle_pinned_on_torch_adam(wd, lr):
    torch.manual_seed(3)
    n = 65536 + 5
    p = torch.nn.Parameter(torch.randn(n))
    opt = torch.optim.Adam([p], lr=lr, weight_decay=wd, foreach=False)
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    for t in range(1, 7):
        g = torch.randn(n) * (10.0 ** torch.randint(-6, 2, (n,)).float())
        if t == 3:
            g[::7] = 0.0  # dead gradients: sqrt(v) + eps path
        pn = p.detach().numpy().copy()  # teacher-forced: every step starts from torch's own parameters
        p.grad = g.clone()
        opt.step()
        ao.adam_step(pn, g.numpy(), m, v, t, lr=lr, weight_decay=wd)
        st = opt.state[p]
        # moments: bit-identical, all steps.  param: bit-identical on torch's vectorised body; its scalar tail (a few
        # elements per parallel chunk) rounds `value * m / denom` differently: <= 2 ulp on <= 0.1 % of the elements
        assert (st["exp_avg"].numpy() == m).all()
        assert (st["exp_avg_sq"].numpy() == v).all()
        d = _ulps(p.detach().numpy(), pn)
        assert d.max() <= 2 and (d != 0).mean() <= 1e-3, (t, d.max(), (d != 0).sum())


def test_oracle_grad_scale_is_a_prescale():
    rng = np.random.default_rng(0)
    p0 = rng.standard_normal(1000).astype(np.float32)
    g = rng.standard_normal(1000).astype(np.float32)
    a = [p0.copy(), np.zeros_like(p0), np.zeros_like(p0)]
    b = [p0.copy(), np.zeros_like(p0), np.zeros_like(p0)]
    ao.adam_step(a[0], g, a[1], a[2], 1, grad_scale=0.25)
    ao.adam_step(b[0], g * np.float32(0.25), b[1], b[2], 1)
    for x, y in zip(a, b):
        assert (x == y).all()


def test_adam_abi_validates_without_gpu():
    sfb_build.build()
    lib = _ffi.lib()
    assert lib.sfb_adam_table_bytes(0) == 0 and lib.sfb_adam_table_bytes(10) >= 10 * 40
    n0 = lib.sfb_launch_count()
    nblk = C.c_int32()
    jobs = (_ffi.AdamJob * 1)()
    assert lib.sfb_adam_plan(jobs, 1, None, 0, C.byref(nblk)) == -1          # no table
    assert lib.sfb_adam_plan(None, 1, 4096, 4096, C.byref(nblk)) == -1       # no jobs
    jobs[0].param, jobs[0].grad, jobs[0].exp_avg, jobs[0].exp_avg_sq, jobs[0].n = 4096, 4096, 4096, 0, 16
    assert lib.sfb_adam_plan(jobs, 1, 4096, 4096, C.byref(nblk)) == -1       # NULL moment
    assert lib.sfb_adam_step(None, 1, 1, 1, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1.0, None) == -1
    assert lib.sfb_adam_step(4096, 1, 1, 0, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1.0, None) == -1   # step counts from 1
    assert lib.sfb_adam_step(4096, 1, 1, 1, 1e-3, 1.0, 0.999, 1e-8, 0.0, 1.0, None) == -1   # beta1 < 1
    assert lib.sfb_adam_step(4096, 1, 1, 1, -1.0, 0.9, 0.999, 1e-8, 0.0, 1.0, None) == -1   # lr >= 0
    assert lib.sfb_launch_count() == n0


def test_adam_constructor_mirrors_torch_errors_and_has_no_cpu_path():
    from video_classification_b200.optim import Adam
    w = torch.nn.Parameter(torch.zeros(4))
    for kw, msg in ((dict(lr=-1.0), "Invalid learning rate"), (dict(eps=-1.0), "Invalid epsilon"),
                    (dict(betas=(1.0, 0.9)), "Invalid beta parameter at index 0"),
                    (dict(betas=(0.9, 1.0)), "Invalid beta parameter at index 1"),
                    (dict(weight_decay=-1.0), "Invalid weight_decay")):
        with pytest.raises(ValueError, match=msg):
            Adam([w], **kw)
        with pytest.raises(ValueError, match=msg):
            torch.optim.Adam([w], **kw)
    with pytest.raises(NotImplementedError):
        Adam([w], amsgrad=True)
    opt = Adam([w], lr=1e-4)
    ref = torch.optim.Adam([w], lr=1e-4)
    assert set(opt.param_groups[0]) == set(ref.param_groups[0])
    opt.step()  # no gradient anywhere: nothing to do, like torch
    w.grad = torch.ones(4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        opt.step()
