"""Run-to-run determinism of the batch-N training step: same inputs, (A) two forwards + backwards, (B) one forward,
two backwards.  Prints the worst per-parameter relative L2 difference of the weight gradients."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.nn.functional as F
import parity_utils as pu
from oracle import slowfast_oracle as so

N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
_, product = pu.build_pair("HYB", True)
xs, xf = so.synthetic_clips(N, 8, 32, 224, seed=5)
xs, xf = xs.cuda(), xf.cuda()
y = torch.randint(0, 249, (N,), generator=torch.Generator().manual_seed(7)).cuda()
product.train()

def grads():
    return {n: p.grad.clone() for n, p in product.named_parameters() if p.grad is not None}

def worst(a, b, k=6):
    r = sorted(((pu.rel_l2(a[n], b[n]), n) for n in a if a[n].dim() >= 2), reverse=True)
    return r[:k]

out = product([xs, xf]); logits1 = out.detach().clone()
dl, = torch.autograd.grad(F.cross_entropy(out, y), out, retain_graph=True)
out.backward(dl, retain_graph=True); g1 = grads()
product.zero_grad(set_to_none=True)
out.backward(dl, retain_graph=True); g1b = grads()
product.zero_grad(set_to_none=True)
print("same forward, backward twice:", worst(g1, g1b))
out2 = product([xs, xf]); print("logits run-to-run rel max:", pu.rel_max(out2.detach(), logits1))
out2.backward(dl); g2 = grads()
print("two forwards:", worst(g1, g2))
num = sum(float((g1[n] - g2[n]).pow(2).sum()) for n in g1); den = sum(float(g1[n].pow(2).sum()) for n in g1)
print("two forwards, whole gradient rel L2:", (num / den) ** 0.5)
tagged = eng_bufs = None
eng = product._engine
