"""ORACLE (test infrastructure).  Generates tests/golden/*.pt from the reference's OWN model/my_slowfast.py
imported verbatim (oracle/ref_import.py; only pytorchvideo underneath is restated).  Run in the authoring
container where /root/reference exists:   python -m oracle.make_golden
The fixtures travel to the GPU box; /root/reference does not.

Each fixture: seeded weights (torch.manual_seed(seed) before init_my_slowfast, then BatchNorm
re-randomisation with bn_seed / final_gamma), seeded clips and labels, and the reference's fp32 results:
eval logits, train-mode logits, loss, and for every parameter with a gradient its L2 norm plus 64 sampled
elements (indices stored)."""
import os
import sys

import torch
import torch.nn.functional as F

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(_HERE)
if _REPO not in sys.path:
    sys.path.insert(0, _REPO)

from oracle import ref_import, slowfast_oracle as so  # noqa: E402

FINAL_GAMMA = 0.1  # keep equal to tests/parity_utils.py
CASES = {
    "ref_small": dict(variant="REF", N=2, S=64, Ts=8, Tf=8),
    "hyb_small": dict(variant="HYB", N=2, S=224, Ts=8, Tf=32),
}


def weight_checksum(model):
    return sum(float(v.double().abs().sum()) for v in model.state_dict().values() if v.dtype.is_floating_point)


def make(name, variant, N, S, Ts, Tf, seed=0, bn_seed=123, clip_seed=1):
    torch.manual_seed(seed)
    model = ref_import.build_reference_model(249, True, (5, 15), (64, 8), variant=variant)
    so.randomize_bn_(model, bn_seed, FINAL_GAMMA)
    so.disable_dropout_(model)
    xs, xf = so.synthetic_clips(N, Ts, Tf, S, seed=clip_seed)
    labels = torch.randint(0, 249, (N,), generator=torch.Generator().manual_seed(7))
    chk = weight_checksum(model)
    model.eval()
    with torch.no_grad():
        eval_logits = model([xs, xf]).clone()
    model.train()
    out = model([xs, xf])
    loss = F.cross_entropy(out, labels)
    loss.backward()
    g = torch.Generator().manual_seed(99)
    grad_norm, grad_idx, grad_samples = {}, {}, {}
    for n, p in model.named_parameters():
        if p.grad is None:
            continue
        flat = p.grad.flatten()
        idx = torch.randperm(flat.numel(), generator=g)[:64].clone()
        grad_norm[n] = float(flat.norm())
        grad_idx[n] = idx
        grad_samples[n] = flat[idx].clone()
    fx = dict(meta=dict(variant=variant, N=N, S=S, Ts=Ts, Tf=Tf, seed=seed, bn_seed=bn_seed, clip_seed=clip_seed,
                        final_gamma=FINAL_GAMMA, labels=labels.tolist(), weight_checksum=chk,
                        n_params=sum(p.numel() for p in model.parameters()), n_keys=len(model.state_dict()),
                        source="reference model/my_slowfast.py imported verbatim (commit 0d76ea9) + oracle/pv_restate.py"),
              eval_logits=eval_logits, train_logits=out.detach().clone(), loss=float(loss.detach()), grad_norm=grad_norm,
              grad_idx=grad_idx, grad_samples=grad_samples, state_dict_keys=list(model.state_dict().keys()))
    os.makedirs(os.path.join(_REPO, "tests", "golden"), exist_ok=True)
    path = os.path.join(_REPO, "tests", "golden", name + ".pt")
    torch.save(fx, path)
    print(f"{name}: loss {float(loss.detach()):.6f} params {fx['meta']['n_params']} keys {fx['meta']['n_keys']} "
          f"grads {len(grad_norm)} -> {path} ({os.path.getsize(path) / 1e3:.0f} kB)")


if __name__ == "__main__":
    if not ref_import.available():
        raise SystemExit("/root/reference not present: fixtures can only be generated in the authoring container")
    for name, kw in CASES.items():
        make(name, **kw)
