"""Yardstick for bf16-vs-fp32 parity on the B200: how far is PyTorch's own bf16 path (autocast, cuDNN) from the
fp32 oracle on the same weights/clips, next to how far the libsfb200 product is.  ReLU-mask flips make
elementwise gradient comparisons across precisions noisy; this puts a number on it.
Usage: python tools/autocast_yardstick.py [REF|HYB] [N] [S] [Ts] [Tf]"""
import copy
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.nn.functional as F

import parity_utils as pu
from oracle import slowfast_oracle as so

variant = sys.argv[1] if len(sys.argv) > 1 else "REF"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 4
S = int(sys.argv[3]) if len(sys.argv) > 3 else 64
Ts = int(sys.argv[4]) if len(sys.argv) > 4 else 8
Tf = int(sys.argv[5]) if len(sys.argv) > 5 else Ts
pu.fp32_reference_mode()
oracle, product = pu.build_pair(variant, True)
ac = copy.deepcopy(oracle)
xs, xf = so.synthetic_clips(N, Ts, Tf, S)
xs, xf = xs.cuda(), xf.cuda()
y = torch.randint(0, 249, (N,), generator=torch.Generator().manual_seed(7)).cuda()
for m in (oracle, product, ac):
    m.train()
ref = oracle([xs, xf]); F.cross_entropy(ref, y).backward()
got = product([xs, xf]); F.cross_entropy(got, y).backward()
with torch.autocast("cuda", dtype=torch.bfloat16):
    a = ac([xs, xf])
F.cross_entropy(a.float(), y).backward()
torch.cuda.synchronize()
print(f"== {variant} N={N} S={S} T={Ts}/{Tf} final_gamma={os.environ.get('SFB_FINAL_GAMMA', pu.FINAL_GAMMA)}")
print("logits rel-max: product %.3e   autocast %.3e" % (pu.rel_max(got.detach(), ref.detach()), pu.rel_max(a.detach().float(), ref.detach())))
gr = {n: p.grad for n, p in oracle.named_parameters() if p.grad is not None}
gp = dict(product.named_parameters())
ga = dict(ac.named_parameters())
ep = sorted(pu.rel_l2(gp[n].grad, g) for n, g in gr.items())
ea = sorted(pu.rel_l2(ga[n].grad.float(), g) for n, g in gr.items())
epa = sorted(pu.rel_l2(gp[n].grad, ga[n].grad.float()) for n in gr)
q = lambda v, f: v[min(len(v) - 1, int(f * len(v)))]
for nm, v in (("product vs fp32", ep), ("autocast vs fp32", ea), ("product vs autocast", epa)):
    print(f"grad rel-L2 {nm:20s}: median {q(v, .5):.3e}  p90 {q(v, .9):.3e}  max {v[-1]:.3e}  frac<=3e-2 {sum(e <= 3e-2 for e in v) / len(v):.2f}")
