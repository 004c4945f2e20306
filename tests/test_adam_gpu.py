"""GPU parity of the optimizer-step row: sfb_adam_step through the drop-in optimizer against the
oracle (bit-exact: same operations, same roundings) and against torch.optim.Adam on the same device."""
import numpy as np
import pytest
import torch

from oracle import adam_oracle as ao
from video_classification_b200 import _ffi
from video_classification_b200.optim import Adam

pytestmark = pytest.mark.gpu

# ragged sizes: empty tail, one element, below / at / above one 4096-element block, odd counts (scalar path)
SIZES = [1, 7, 64, 249, 4095, 4096, 4097, 3 * 4096 + 5, 64 * 249, 8 * 3 * 3 * 64 * 64 + 3]


def _params(seed, sizes, dev="cuda"):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(n, generator=g).to(dev)) for n in sizes]


def _grad(seed, n):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, generator=g) * (10.0 ** torch.randint(-6, 2, (n,), generator=g).float())
    x[::11] = 0.0
    return x


@pytest.mark.parametrize("wd,scale", [(0.0, 1.0), (1e-2, 1.0), (0.0, 0.125)])
def test_adam_step_bit_exact_against_oracle(wd, scale):
    ps = _params(1, SIZES)
    ps.append(torch.nn.Parameter(torch.randn(1000, device="cuda")))  # never gets a gradient: skipped like torch
    host = [(p.detach().cpu().numpy().copy(), np.zeros(p.numel(), np.float32), np.zeros(p.numel(), np.float32))
            for p in ps]
    opt = Adam(ps, lr=1e-4, weight_decay=wd, grad_scale=scale)
    n0 = _ffi.launch_count()
    for t in range(1, 6):
        for i, p in enumerate(ps[:-1]):
            g = _grad(100 * t + i, p.numel())
            # odd element offset every other tensor: the unaligned (scalar) path of the kernel
            if i % 2:
                buf = torch.empty(p.numel() + 1, device="cuda")
                buf[1:].copy_(g)
                p.grad = buf[1:]
            else:
                p.grad = g.cuda()
            ao.adam_step(host[i][0], g.numpy(), host[i][1], host[i][2], t, lr=1e-4, weight_decay=wd, grad_scale=scale)
        opt.step()
        for i, p in enumerate(ps[:-1]):
            st = opt.state[p]
            assert int(st["step"]) == t
            assert (st["exp_avg"].cpu().numpy() == host[i][1]).all(), (t, i)
            assert (st["exp_avg_sq"].cpu().numpy() == host[i][2]).all(), (t, i)
            assert (p.detach().cpu().numpy() == host[i][0]).all(), (t, i)
    assert (ps[-1].detach().cpu().numpy() == host[-1][0]).all() and len(opt.state[ps[-1]]) == 0
    assert _ffi.launch_count() - n0 == 5 and opt.launches == 5, "one launch per optimizer step"


@pytest.mark.parametrize("wd", [0.0, 1e-2])
def test_adam_matches_torch_adam_on_device_and_swaps_state_dicts(wd):
    sizes = [64 * 5 * 49, 64, 64, 2048 * 512, 249 * 2304, 249]
    a, b = _params(2, sizes), _params(2, sizes)
    ours, ref = Adam(a, lr=1e-3, weight_decay=wd), torch.optim.Adam(b, lr=1e-3, weight_decay=wd, foreach=False)
    for t in range(1, 5):
        for i, (p, q) in enumerate(zip(a, b)):
            g = _grad(7 * t + i, p.numel()).cuda()
            p.grad, q.grad = g.clone(), g.clone()
        ours.step()
        ref.step()
        if t == 2:  # checkpoints move both ways (same state_dict layout)
            sd_o, sd_r = ours.state_dict(), ref.state_dict()
            assert sd_o["state"].keys() == sd_r["state"].keys()
            assert all(set(sd_o["state"][k]) == set(sd_r["state"][k]) for k in sd_o["state"])
            ours2 = Adam(a, lr=1e-3, weight_decay=wd)
            ours2.load_state_dict(sd_r)
            ref2 = torch.optim.Adam(b, lr=1e-3, weight_decay=wd, foreach=False)
            ref2.load_state_dict(sd_o)
            ours, ref = ours2, ref2
    for p, q in zip(a, b):
        # torch's CUDA kernels divide by the bias-correction scalar through a reciprocal multiply: 1-2 ulp on the
        # update, i.e. <= 1e-6 relative to the step size lr on the parameter.  Tolerance: 2e-6 * lr-scale absolute.
        assert (p - q).abs().max().item() <= 2e-6 * 1e-3 + 2.4e-7 * q.abs().max().item()
    for p, q in zip(a, b):
        so, sr = ours.state[p], ref.state[q]
        assert int(so["step"]) == int(sr["step"]) == 4
        # moments: identical operations on both sides; with weight decay the 1-2 ulp parameter difference feeds back
        # through g = grad + wd * param, so the bound is relative to the tensor's scale, not per element
        for k in ("exp_avg", "exp_avg_sq"):
            assert (so[k] - sr[k]).abs().max().item() <= 1e-6 * sr[k].abs().max().item()


def test_adam_full_model_size_properties():
    """34 M parameters in 332 tensors (the live parameters of the benchmark model): first step moves every element
    with a non-zero gradient by exactly -lr*sign(g) up to rounding, zero gradients leave parameters untouched, and a
    second optimizer over the same tensors replays bit-identically (idempotent planning, no cross-tensor writes)."""
    g = torch.Generator().manual_seed(5)
    sizes = [int(x) for x in torch.randint(16, 200000, (331,), generator=g)] + [2048 * 512 * 3]
    ps = [torch.nn.Parameter(torch.randn(n, device="cuda")) for n in sizes]
    qs = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    before = [p.detach().clone() for p in ps]
    for p, q in zip(ps, qs):
        gr = torch.randn(p.numel(), device="cuda")
        gr[::3] = 0.0
        p.grad, q.grad = gr, gr.clone()
    o1, o2 = Adam(ps, lr=1e-4), Adam(qs, lr=1e-4)
    n0 = _ffi.launch_count()
    o1.step()
    o2.step()
    assert _ffi.launch_count() - n0 == 2
    for p, q, p0 in zip(ps, qs, before):
        assert torch.equal(p, q)
        d = (p.detach() - p0)
        z = p.grad == 0
        assert (d[z] == 0).all()
        # |g| >> eps: m/(sqrt(v)) = sign(g) after bias correction
        big = p.grad.abs() > 1e-3
        assert torch.allclose(d[big], -1e-4 * torch.sign(p.grad[big]), rtol=1e-3, atol=2.5e-7 * float(p0.abs().max()))


def test_training_loop_with_own_adam_tracks_torch_adam_on_the_oracle():
    """The reference loop (train.py:225-231) on both sides: oracle + torch.optim.Adam against product + our Adam, same
    weights, same clips, 3 steps.  After every step the engine must have re-packed its bf16 conv operands from the
    updated fp32 weights (the kernel writes through raw pointers and bumps the version counters itself)."""
    import torch.nn.functional as F
    import parity_utils as pu
    from oracle import slowfast_oracle as so
    pu.fp32_reference_mode()
    oracle, product = pu.build_pair("REF", True, device="cuda")
    oracle.train()
    product.train()
    xs, xf = so.synthetic_clips(2, 8, 8, 64)
    xs, xf = xs.cuda(), xf.cuda()
    y = torch.tensor([3, 200], device="cuda")
    o_ref, o_own = torch.optim.Adam(oracle.parameters(), lr=1e-3, foreach=False), Adam(product.parameters(), lr=1e-3)
    with torch.no_grad():
        product([xs, xf])  # the engine is built by the first forward
    eng = product._engine
    w = eng.layers[0].conv.weight
    losses = []
    for t in range(3):
        v0, w0 = w._version, w.detach().clone()
        for model, opt in ((oracle, o_ref), (product, o_own)):
            out = model([xs, xf])
            loss = F.cross_entropy(out, y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(float(loss.detach()))
        assert w._version > v0 and not torch.equal(w, w0), "the optimizer step must be visible to autograd"
        assert abs(losses[-1] - losses[-2]) <= 2e-2 * abs(losses[-2]), (t, losses)
    # next forward re-packs: every layer's operand version follows its weight, and a pointwise layer's packed rows are
    # the bf16 rounding of the CURRENT fp32 weight
    with torch.no_grad():
        product([xs, xf])
    assert all(l.packed_version == l.conv.weight._version for l in eng.layers)
    pw = next(l for l in eng.layers if l.taps == 1 and l.fold_f == 1 and l.wp is not None)
    Co, Ci = pw.conv.weight.shape[:2]
    assert torch.equal(pw.wp[:Co, :Ci].float(), pw.conv.weight.detach().view(Co, Ci).bfloat16().float())
    # parameters stay on the oracle's trajectory: Adam's first steps move every element by ~lr, so compare in units of lr
    n_bad = 0
    for (k, p), q in zip(product.named_parameters(), oracle.parameters()):
        if p.grad is None:
            continue
        d = (p - q).abs().max().item()
        n_bad += d > 3 * 1e-3 * 3  # never further apart than the 3 steps' total travel
    assert n_bad == 0
    assert losses[-1] < losses[1], "three steps on one batch must reduce its loss"


@pytest.mark.parametrize("kind", ["fused", "foreach", "raw"])
def test_engine_repacks_operands_after_any_optimizer(kind):
    """Whatever updates the fp32 weights between two training steps -- torch's fused or foreach Adam, or a write that
    never touches a version counter -- the next forward runs on operands packed from the CURRENT weights."""
    import torch.nn.functional as F
    import parity_utils as pu
    from oracle import slowfast_oracle as so
    _, product = pu.build_pair("REF", True, device="cuda")
    product.train()
    xs, xf = so.synthetic_clips(2, 8, 8, 64)
    xs, xf = xs.cuda(), xf.cuda()
    y = torch.tensor([3, 200], device="cuda")
    opt = None if kind == "raw" else torch.optim.Adam(product.parameters(), lr=1e-3, fused=kind == "fused",
                                                      foreach=kind == "foreach")
    F.cross_entropy(product([xs, xf]), y).backward()
    eng = product._engine
    pw = next(l for l in eng.layers if l.taps == 1 and l.fold_f == 1 and l.wp is not None)
    Co, Ci = pw.conv.weight.shape[:2]
    old = pw.wp[:Co, :Ci].clone()
    if opt is not None:
        opt.step()
    else:  # through a detached alias and an out= kernel on its storage view: the parameter's own counter may not move
        alias = torch.as_strided(pw.conv.weight.detach().view(-1).view(torch.int32).view(torch.float32), (Co * Ci,), (1,))
        torch.mul(alias.clone(), 1.5, out=alias)
    n0 = _ffi.launch_count()
    product([xs, xf])
    assert not torch.equal(pw.wp[:Co, :Ci], old)
    assert torch.equal(pw.wp[:Co, :Ci].float(), pw.conv.weight.detach().view(Co, Ci).bfloat16().float())
    assert _ffi.launch_count() - n0 > 0
