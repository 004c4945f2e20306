"""Helpers shared by the parity tests: build (oracle, product) model pairs with identical seeded weights and
compare logits / gradients with the tolerances BASELINE.json's north_star states:
    logits  : max|a-b| / max|b| <= 2e-2 (bf16 kernels vs fp32 oracle), identical top-1
    grads   : per-parameter relative L2 error <= 3e-2
"""
import os

import torch

from oracle import slowfast_oracle as so
from video_classification_b200.my_slowfast import init_my_slowfast

LOGIT_TOL = 2e-2
GRAD_TOL = 3e-2
# gamma scale of each bottleneck's last BatchNorm in parity runs (see oracle.slowfast_oracle.randomize_bn_)
FINAL_GAMMA = 0.1

PRODUCT_KW = {
    "REF": {},
    "HYB": dict(fusion_conv_stride=(4, 1, 1), head_pool_kernel_sizes=((8, 7, 7), (32, 7, 7))),
}


def build_pair(variant="REF", fuse=True, num_class=249, seed=0, bn_seed=123, device="cuda", final_gamma=None, twin=False,
               input_channels=(5, 15)):
    """(oracle, product) on identical seeded weights.  twin=True: the oracle becomes the bf16-storage twin
    (oracle.slowfast_oracle.add_bf16_storage_hooks_; conv weights pre-rounded to bf16 in BOTH models) -- the
    comparison partner for gradients; twin=False: the pure fp32 reference path -- the partner for logits/top-1."""
    torch.manual_seed(seed)
    oracle = so.build_oracle(num_class=num_class, fuse=fuse, input_channels=tuple(input_channels), **so.VARIANTS[variant])
    if final_gamma is None:
        final_gamma = float(os.environ.get("SFB_FINAL_GAMMA", FINAL_GAMMA))
    so.randomize_bn_(oracle, bn_seed, final_gamma)
    so.disable_dropout_(oracle)
    if twin:
        so.round_conv_weights_to_bf16_(oracle)
    product = init_my_slowfast(so.make_cfg(num_class=num_class, fuse=fuse), tuple(input_channels), (64, 8),
                               **PRODUCT_KW[variant])
    product.load_state_dict(oracle.state_dict(), strict=True)
    product.blocks[6].dropout.p = 0.0
    if twin:
        so.add_bf16_storage_hooks_(oracle)
    return oracle.to(device), product.to(device)


def rel_max(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-12)).item()


def rel_l2(a, b):
    return ((a - b).norm() / b.norm().clamp_min(1e-20)).item()


def fp32_reference_mode():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


def grad_report(product, ref_grads: dict):
    """returns list of (name, rel_l2) for every parameter that has a reference gradient; checks None-ness matches"""
    out = []
    for name, p in product.named_parameters():
        rg = ref_grads.get(name)
        if rg is None:
            assert p.grad is None, f"{name}: reference has no gradient (dead parameter) but product produced one"
            continue
        assert p.grad is not None, f"{name}: missing gradient"
        out.append((name, rel_l2(p.grad.float().cpu(), rg.float().cpu())))
    return out


def capture_forward_state(model):
    """hooks on an oracle/twin model that record, in the engine's tag vocabulary (NDHWC):
    `<conv>.y` raw conv outputs, `<res block>.O` block outputs, `<stem>.O` stem outputs, and the pooled feature
    vector entering the FC head.  Returns (state dict, feat holder list, handles)."""
    import torch.nn as nn
    state, feat, handles = {}, [], []

    def ndhwc(t):
        return t.detach().permute(0, 2, 3, 4, 1).contiguous()

    def rec(key):
        return lambda mod, inp, out: state.__setitem__(key, ndhwc(out))

    for name, m in model.named_modules():
        parts = name.split(".")
        if isinstance(m, nn.Conv3d) and ".residual." not in name and ".res_unit." not in name:
            handles.append(m.register_forward_hook(rec(name + ".y")))
        elif len(parts) >= 2 and parts[-2] == "res_blocks":
            handles.append(m.register_forward_hook(rec(name + ".O")))
        elif name in ("blocks.0.multipathway_blocks.0", "blocks.0.multipathway_blocks.1"):
            handles.append(m.register_forward_hook(rec(name + ".O")))
        elif name == "blocks.6.proj":
            handles.append(m.register_forward_pre_hook(
                lambda mod, inp: feat.append(inp[0].detach().reshape(inp[0].shape[0], -1, inp[0].shape[-1]).mean(1))))
    return state, feat, handles


def assert_same_top1(got, ref, tol=LOGIT_TOL):
    """identical top-1, where the oracle itself decides: a clip whose two best oracle logits are closer than the
    logit tolerance (tol * max|ref|, the error BASELINE.json allows a bf16 path) is a numerical tie that either
    implementation may break either way; there the product's choice must be one of the tied classes."""
    slack = tol * float(ref.abs().max())
    g1, r1 = got.argmax(1), ref.argmax(1)
    for n in range(ref.shape[0]):
        if int(g1[n]) == int(r1[n]):
            continue
        gap = float(ref[n, r1[n]] - ref[n, g1[n]])
        assert gap <= slack, f"clip {n}: top-1 {int(g1[n])} vs oracle {int(r1[n])}, oracle margin {gap:.3e} > slack {slack:.3e}"
