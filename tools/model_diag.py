"""End-to-end diagnostic on the B200: product vs oracle (logits, gradients), prints a compact report.
Usage: python tools/model_diag.py [REF|HYB] [N] [S] [Ts] [Tf] [fuse]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.nn.functional as F

import parity_utils as pu
from oracle import slowfast_oracle as so

variant = sys.argv[1] if len(sys.argv) > 1 else "REF"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2
S = int(sys.argv[3]) if len(sys.argv) > 3 else 64
Ts = int(sys.argv[4]) if len(sys.argv) > 4 else 8
Tf = int(sys.argv[5]) if len(sys.argv) > 5 else Ts
fuse = (sys.argv[6] != "0") if len(sys.argv) > 6 else True
twin = (sys.argv[7] == "twin") if len(sys.argv) > 7 else False
pu.fp32_reference_mode()
oracle, product = pu.build_pair(variant, fuse, twin=twin)
xs, xf = so.synthetic_clips(N, Ts, Tf, S)
xs, xf = xs.cuda(), xf.cuda()
y = torch.randint(0, 249, (N,), generator=torch.Generator().manual_seed(7)).cuda()
print(f"== {variant} fuse={fuse} N={N} S={S} T={Ts}/{Tf}", flush=True)
oracle.eval(); product.eval()
with torch.no_grad():
    ref = oracle([xs, xf]); got = product([xs, xf])
torch.cuda.synchronize()
print("eval  logits rel-err %.3e  top1 equal %s  (ref max %.3e)" % (pu.rel_max(got, ref), torch.equal(got.argmax(1), ref.argmax(1)), ref.abs().max().item()), flush=True)
oracle.train(); product.train()
ref = oracle([xs, xf]); F.cross_entropy(ref, y).backward()
got = product([xs, xf]); loss = F.cross_entropy(got, y); loss.backward()
torch.cuda.synchronize()
print("train logits rel-err %.3e  top1 equal %s" % (pu.rel_max(got.detach(), ref.detach()), torch.equal(got.argmax(1), ref.argmax(1))), flush=True)
rep = pu.grad_report(product, {n: p.grad for n, p in oracle.named_parameters()})
bad = [(n, e) for n, e in rep if e > pu.GRAD_TOL]
print(f"grads: {len(rep)} compared, {len(bad)} above {pu.GRAD_TOL}; median {sorted(e for _, e in rep)[len(rep)//2]:.3e}")
for n, e in sorted(rep, key=lambda t: -t[1])[:12]:
    print(f"   {e:.3e}  {n}")
# timing of the product (train step fwd+bwd), a few iterations
for mode in ("train", "eval"):
    product.train(mode == "train")
    ts = []
    for it in range(5):
        torch.cuda.synchronize(); t0 = time.time()
        if mode == "train":
            out = product([xs, xf]); F.cross_entropy(out, y).backward()
        else:
            with torch.no_grad():
                product([xs, xf])
        torch.cuda.synchronize(); ts.append(time.time() - t0)
    print(f"{mode}: ms/iter {[round(t*1e3,2) for t in ts]}", flush=True)
eng = product._engine
print("plans:", {k: round(v.act_bytes / 1e9, 3) for k, v in eng.plans.items()})
