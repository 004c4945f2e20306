"""Data-parallel training of the drop-in model checked ON THE GPU (reference step: train.py:225-231; the gradient
exchange is video_classification_b200/parallel.py).

Two variants of the same worker:
  * `nccl`  : 2 ranks on 2 GPUs over NCCL -- the product path; skipped on a box with fewer than 2 GPUs;
  * `gloo`  : 2 ranks sharing GPU 0, the all-reduce going through gloo on CUDA tensors -- exercises the engine's
              bucket plumbing (launch-on-last-ready, side-stream fences, arena views) on every single-GPU box.

Checked per rank:
  (a) every bucket was launched only when its gradients were final: the slice as it was at launch time equals the
      gradient a purely local backward of the SAME forward state produces (up to the fp32 atomics-order noise of two
      backward runs), and the reduced arena equals the mean over ranks of those launch-time slices to fp32 rounding;
  (b) after 3 Adam steps the parameters are bit-identical across the ranks (BatchNorm running statistics stay
      per-replica, as with the reference's plain nn.BatchNorm3d);
  (c) a tiny bucket size puts bucket boundaries inside a layer's (weight, gamma, beta) triple.
"""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, backend, ret):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    import torch.nn.functional as F
    import parity_utils as pu
    from oracle import slowfast_oracle as so
    from video_classification_b200.optim import Adam
    from video_classification_b200.parallel import attach_ddp

    import datetime
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank if backend == "nccl" else 0)
    torch.cuda.set_device(dev)
    # a rank that fails must not leave its peer waiting in a collective for NCCL's default 10 minutes
    tmo = datetime.timedelta(seconds=90)
    if backend == "nccl":
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev, timeout=tmo)
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world, timeout=tmo)
    log_dir = os.path.join(os.path.dirname(here), "gpurun_out")
    log = open(os.path.join(log_dir, f"ddp_{backend}_rank{rank}.log"), "w") if os.path.isdir(log_dir) else None

    def mark(msg):
        if log is not None:
            log.write(msg + "\n")
            log.flush()

    try:
        mark("process group up")
        # rank 1 starts from DIFFERENT weights: attach_ddp's broadcast must make the replicas equal
        _, product = pu.build_pair("REF", True, seed=rank, device=dev)
        product.blocks[6].dropout.p = 0.5  # as the reference trains; per-rank masks (engine._draw_dropout_seed)
        product.train()
        ddp = attach_ddp(product, bucket_mb=0.02)
        eng = product._engine
        bk = ddp.bucketer
        sd = [p.detach().flatten() for p in product.parameters()]
        flat0 = torch.cat(sd)
        both = [torch.empty_like(flat0) for _ in range(world)]
        dist.all_gather(both, flat0)
        assert torch.equal(both[0], both[1]), "broadcast_parameters did not equalise the replicas"
        mark("broadcast ok")

        # (c) bucket boundaries inside a (weight, gamma, beta) triple
        offs = eng.grad_offsets
        split = 0
        for l in eng.layers:
            bs = {bk.span_bucket[offs[id(p)]] for p in (l.conv.weight, l.bn.weight, l.bn.bias)}
            split += len(bs) > 1
        assert split >= 5 and len(bk.buckets) > 50, (split, len(bk.buckets))

        xs, xf = so.synthetic_clips(2, 8, 8, 64, seed=10 + rank)
        xs, xf = xs.to(dev), xf.to(dev)
        y = torch.randint(0, 249, (2,), generator=torch.Generator().manual_seed(20 + rank)).to(dev)

        # ---- (a) ----
        snaps = {}
        orig_launch = bk.launch

        def launch(b):
            s, e = bk.buckets[b]
            snaps[b] = bk.flat[s:e].clone()  # compute stream, ordered before the bucket's event record
            return orig_launch(b)

        bk.launch = launch
        torch.manual_seed(5)
        out = product([xs, xf])
        loss = F.cross_entropy(out, y)
        dl, = torch.autograd.grad(loss, out, retain_graph=True)
        out.backward(dl, retain_graph=True)       # data-parallel backward: arena = mean over ranks
        torch.cuda.synchronize(dev)
        mark("data-parallel backward done")
        reduced = eng.grad_flat.clone()
        assert sorted(snaps) == list(range(len(bk.buckets))), "a bucket was never launched"
        product.zero_grad(set_to_none=True)           # .grad aliases the arena: a second backward must not accumulate
        bk.launch = orig_launch                       # (the snapshots stay those of the data-parallel run)
        w_saved, bk.world = bk.world, 1               # purely local backward of the SAME forward state
        out.backward(dl)
        bk.world = w_saved
        torch.cuda.synchronize(dev)
        local = eng.grad_flat.clone()
        lo, hi = bk.buckets[0][0], bk.buckets[-1][1]
        at_launch = torch.zeros_like(local)
        for b, (s, e) in enumerate(bk.buckets):
            at_launch[s:e] = snaps[b]
        # launch-time slices vs the finished local gradient: a bucket launched before a contribution landed would be
        # off by that whole contribution; two runs of the same backward differ by atomics order only
        num = float((at_launch[lo:hi] - local[lo:hi]).norm())
        den = float(local[lo:hi].norm())
        assert num <= 5e-3 * den, f"bucket launched before its gradients were final: rel {num / den:.3e}"
        # per tensor the same holds up to the run-to-run noise of small tensors (atomics order -> bf16 rounding of the
        # activation gradients: up to ~0.2 on 8-element BatchNorm vectors); a missing contribution is O(1)
        for p in eng.live_params:
            o, n = offs[id(p)], p.numel()
            a, b_ = at_launch[o:o + n], local[o:o + n]
            assert float((a - b_).norm()) <= 0.35 * float(b_.norm()) + 1e-12, "per-tensor launch-time gradient differs"
        # the reduced arena is the mean over ranks of what each rank handed over (fp32 rounding of a 2-term mean)
        gathered = [torch.empty_like(at_launch) for _ in range(world)]
        dist.all_gather(gathered, at_launch)
        mean = (gathered[0] + gathered[1]) * 0.5
        err = float((reduced[lo:hi] - mean[lo:hi]).abs().max())
        scale = float(mean[lo:hi].abs().max())
        assert err <= 1e-5 * scale, f"all-reduce result is not the mean of the ranks' gradients: {err:.3e} vs {scale:.3e}"

        mark("(a) checks passed")
        # ---- (b) three optimizer steps, replicas stay bit-identical ----
        opt = Adam(product.parameters(), lr=1e-3)
        for it in range(3):
            out = product([xs, xf])
            loss = F.cross_entropy(out, y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            torch.cuda.synchronize(dev)
            mark(f"step {it} done, graphs: {sum(len(pl.graphs) for pl in eng.plans.values())} errors: {eng.graph_errors[:1]}")
        torch.cuda.synchronize(dev)
        # in the plain training loop .grad of every live parameter aliases the arena the all-reduce ran on (autograd
        # adopts the returned views; with retain_graph above it clones them, which is why (a) reads the arena itself)
        for p in eng.live_params:
            assert p.grad is not None and p.grad.data_ptr() == eng.grad_views[id(p)].data_ptr()
        flat1 = torch.cat([p.detach().flatten() for p in product.parameters()])
        both = [torch.empty_like(flat1) for _ in range(world)]
        dist.all_gather(both, flat1)
        assert torch.equal(both[0], both[1]), "replicas diverged after 3 Adam steps"
        assert not torch.equal(flat1, flat0), "parameters did not move"
        assert bool(torch.isfinite(flat1).all())
        ret[rank] = "ok"
        mark("ok")
    except BaseException as e:  # noqa: BLE001 -- reported to the parent, which fails the test
        import traceback
        ret[rank] = "".join(traceback.format_exception(type(e), e, e.__traceback__))[-3000:]
        mark("FAILED: " + ret[rank])
        os._exit(1)  # do not join a process group whose peer may be stuck in a collective
    finally:
        if log is not None:
            log.close()
    try:
        dist.destroy_process_group()
    except Exception:  # pragma: no cover
        pass


def _run(backend):
    import torch.multiprocessing as mp
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    try:
        mp.spawn(_worker, args=(world, port, backend, ret), nprocs=world, join=True)
    except Exception as e:  # a worker exited non-zero: its message is in `ret`
        if not any(ret.get(r) not in (None, "ok") for r in range(world)):
            raise e
    bad = {r: ret.get(r) for r in range(world) if ret.get(r) != "ok"}
    assert not bad, "\n".join(f"rank {r}: {m}" for r, m in bad.items())


@pytest.mark.timeout(240)
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
def test_ddp_nccl_two_gpus():
    _run("nccl")


@pytest.mark.timeout(240)
def test_ddp_gloo_two_ranks_one_gpu():
    _run("gloo")
