"""Helpers shared by the parity tests: build (oracle, product) model pairs with identical seeded weights and
compare logits / gradients with the tolerances BASELINE.json's north_star states:
    logits  : max|a-b| / max|b| <= 2e-2 (bf16 kernels vs fp32 oracle), identical top-1
    grads   : per-parameter relative L2 error <= 3e-2
"""
import torch

from oracle import slowfast_oracle as so
from video_classification_b200.my_slowfast import init_my_slowfast

LOGIT_TOL = 2e-2
GRAD_TOL = 3e-2

PRODUCT_KW = {
    "REF": {},
    "HYB": dict(fusion_conv_stride=(4, 1, 1), head_pool_kernel_sizes=((8, 7, 7), (32, 7, 7))),
}


def build_pair(variant="REF", fuse=True, num_class=249, seed=0, bn_seed=123, device="cuda"):
    torch.manual_seed(seed)
    oracle = so.build_oracle(num_class=num_class, fuse=fuse, **so.VARIANTS[variant])
    so.randomize_bn_(oracle, bn_seed)
    so.disable_dropout_(oracle)
    product = init_my_slowfast(so.make_cfg(num_class=num_class, fuse=fuse), (5, 15), (64, 8), **PRODUCT_KW[variant])
    product.load_state_dict(oracle.state_dict(), strict=True)
    product.blocks[6].dropout.p = 0.0
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
