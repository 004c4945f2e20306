"""Layer-by-layer diagnostic on the B200: snapshots every tagged intermediate of the engine (raw conv outputs,
block outputs, conv-output gradients, block-output gradients) and compares it with the oracle's hooks.
Prints forward entries in forward order and backward entries in backward order, so the first line that goes
bad localises a bug.   Usage: python tools/layer_diag.py [REF|HYB] [N] [S] [Ts] [Tf] [train|eval]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.nn as nn
import torch.nn.functional as F

import parity_utils as pu
from oracle import slowfast_oracle as so

variant = sys.argv[1] if len(sys.argv) > 1 else "REF"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2
S = int(sys.argv[3]) if len(sys.argv) > 3 else 64
Ts = int(sys.argv[4]) if len(sys.argv) > 4 else 8
Tf = int(sys.argv[5]) if len(sys.argv) > 5 else Ts
mode = sys.argv[6] if len(sys.argv) > 6 else "train"
twin = (sys.argv[7] == "twin") if len(sys.argv) > 7 else True
pu.fp32_reference_mode()
oracle, product = pu.build_pair(variant, True, twin=twin)
xs, xf = so.synthetic_clips(N, Ts, Tf, S)
xs, xf = xs.cuda(), xf.cuda()
y = torch.randint(0, 249, (N,), generator=torch.Generator().manual_seed(7)).cuda()

ref_f, ref_b, order_f, order_b = {}, {}, [], []


def ndhwc(t):
    return t.permute(0, 2, 3, 4, 1).float()


def hook(name, kind):
    def fwd(mod, inp, out):
        ref_f[name + kind] = ndhwc(out.detach())
        order_f.append(name + kind)

    def bwd(mod, gin, gout):
        key = name + (".dy" if kind == ".y" else ".dO")
        ref_b[key] = ndhwc(gout[0].detach())
        order_b.append(key)
    return fwd, bwd


for name, mod in oracle.named_modules():
    kind = None
    if isinstance(mod, nn.Conv3d) and ".residual." not in name and ".res_unit." not in name:
        kind = ".y"
    elif name.split(".")[-2:-1] == ["res_blocks"]:
        kind = ".O"
    if kind:
        f, b = hook(name, kind)
        mod.register_forward_hook(f)
        mod.register_full_backward_hook(b)

oracle.train(mode == "train")
product.train(mode == "train")
cap = {}
if mode == "train":
    ref = oracle([xs, xf])
    F.cross_entropy(ref, y).backward()
    product([xs, xf])  # builds the engine
    product._engine.capture = cap
    got = product([xs, xf])
    F.cross_entropy(got, y).backward()
else:
    with torch.no_grad():
        ref = oracle([xs, xf])
        product([xs, xf])
        product._engine.capture = cap
        got = product([xs, xf])
torch.cuda.synchronize()
print(f"== {variant} N={N} S={S} T={Ts}/{Tf} {mode}; logits rel-max {pu.rel_max(got.detach(), ref.detach()):.3e}")


def show(keys, refd):
    for k in keys:
        if k not in cap:
            print(f"  ---      {k} (not captured)")
            continue
        a, b = cap[k], refd[k]
        if a.shape != b.shape:
            print(f"  SHAPE    {k} {tuple(a.shape)} vs {tuple(b.shape)}")
            continue
        e = pu.rel_l2(a, b)
        print(f"  {'OK ' if e < 3e-2 else 'BAD'} {e:.3e} {k}  |ref|={b.norm().item():.3e}")


print("-- forward (forward order)")
show(order_f, ref_f)
if mode == "train":
    print("-- backward (backward order)")
    show(order_b, ref_b)
    rep = pu.grad_report(product, {n: p.grad for n, p in oracle.named_parameters()})
    print("-- parameter gradients (module order)")
    for n, e in rep:
        print(f"  {'OK ' if e < 3e-2 else 'BAD'} {e:.3e} {n}")

if mode == "train" and os.environ.get("SFB_EMU"):
    # re-derive one BatchNorm backward in torch from the engine's OWN captured inputs
    eng = product._engine
    blk = os.environ["SFB_EMU"]  # e.g. blocks.4.multipathway_blocks.1.res_blocks.2
    L = [l for l in eng.layers if l.conv_name == blk + ".branch2.conv_c"][0]
    dO, yc, O = cap[blk + ".dO"], cap[blk + ".branch2.conv_c.y"], cap[blk + ".O"]
    M = yc.numel() // yc.shape[-1]
    g = dO * (O > 0).float()
    xh = (yc - L.mean) * L.invstd
    db = g.reshape(M, -1).sum(0)
    dg = (g * xh).reshape(M, -1).sum(0)
    dy = L.bn.weight.detach() * L.invstd * (g - db / M - xh * dg / M)
    k = blk + ".branch2.conv_c.dy"
    print("EMU: emu-vs-captured %.3e  emu-vs-ref %.3e  captured-vs-ref %.3e" % (pu.rel_l2(dy, cap[k]), pu.rel_l2(dy, ref_b[k]), pu.rel_l2(cap[k], ref_b[k])))
    print("EMU: dgamma emu-vs-engine %.3e  dbeta %.3e" % (pu.rel_l2(dg, L.dgamma), pu.rel_l2(db, L.dbeta)))
    print("EMU: ref dO vs cap %.3e ; mask mismatch frac %.3e" % (pu.rel_l2(dO, ref_b[blk + ".dO"]), ((O > 0) != (ref_f[blk + ".O"] > 0)).float().mean().item()))
    bnm = dict(oracle.named_modules())[blk + ".branch2.norm_c"]
    print("EMU: engine mean vs torch batch mean %.3e ; invstd %.3e" % (
        pu.rel_l2(L.mean, ref_f[blk + ".branch2.conv_c.y"].reshape(M, -1).mean(0)),
        pu.rel_l2(L.invstd, 1.0 / torch.sqrt(ref_f[blk + ".branch2.conv_c.y"].reshape(M, -1).var(0, unbiased=False) + 1e-5))))
