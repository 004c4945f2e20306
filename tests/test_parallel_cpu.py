"""CPU tests (gloo, world_size 2) of the data-parallel gradient exchange: bucket planning over the flat gradient
arena, launch-on-last-ready ordering, skipping of parameters without gradient, mean over ranks."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from video_classification_b200.parallel import GradBucketer


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_bucket_planning_is_contiguous_and_in_arena_order():
    flat = torch.zeros(1000)
    spans = [(0, 100), (104, 300), (408, 50), (464, 500)]
    b = GradBucketer(flat, spans, bucket_bytes=4 * 420)
    assert b.buckets == [(0, 404), (408, 458), (464, 964)]
    assert [b.span_bucket[o] for o, _ in spans] == [0, 0, 1, 2]
    # a bucket launches exactly when its LAST member is marked ready, whatever the order
    assert b.mark_ready(104) is None
    assert b.mark_ready(464) == 2
    assert b.mark_ready(0) == 0
    assert b.mark_ready(408) == 1
    b.reset_counts()
    assert b.mark_ready(0) is None


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(100 + rank)
        flat = torch.zeros(4096)
        spans = [(0, 1000), (1000, 24), (1024, 2000), (3024, 8), (3032, 1000)]  # [4032, 4096) = padding, untouched
        bk = GradBucketer(flat, spans, bucket_bytes=4 * 1500)
        assert len(bk.buckets) >= 3
        grads = torch.randn(4096)
        grads[4032:] = 7.0
        bk.begin()
        # "backward": gradients become final out of arena order; launch when a bucket completes
        for off, n in [(1000, 24), (0, 1000), (3024, 8), (1024, 2000), (3032, 1000)]:
            flat[off:off + n] = grads[off:off + n]
            b = bk.mark_ready(off)
            if b is not None:
                bk.launch(b)
        flat[4032:] = grads[4032:]
        bk.finish()
        allg = [torch.zeros(4096) for _ in range(world)]
        dist.all_gather(allg, grads)
        want = torch.stack(allg).mean(0)
        ok = torch.allclose(flat[:4032], want[:4032], atol=1e-6) and bool((flat[4032:] == 7.0).all())
        ok = ok and sorted(bk.launch_order) == list(range(len(bk.buckets)))
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_bucketed_allreduce_gloo_world2():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    assert all(ret.get(r) for r in range(world)), dict(ret)


def test_engine_plan_inserts_one_launch_per_bucket():
    """dry-run plan (meta device): every bucket gets exactly one launch entry, placed after the op that finalises
    its last gradient; the FC head's bucket launches first (reverse-forward order)."""
    from oracle import slowfast_oracle as so
    from video_classification_b200.engine import SlowFastEngine
    from video_classification_b200.my_slowfast import init_my_slowfast
    from video_classification_b200.parallel import EngineDDP
    m = init_my_slowfast(so.make_cfg(), (5, 15), (64, 8))
    eng = SlowFastEngine(m, dry_run=True)
    eng.ddp = EngineDDP(eng, bucket_mb=8.0)
    eng.refresh_weights(True)
    plan = eng.build_plan(2, 8, 8, 64, 64, True)
    launches = [i for i, (fn, args, tag) in enumerate(plan.bwd) if getattr(fn, "is_bucket_launch", False)]
    nb = len(eng.ddp.bucketer.buckets)
    assert nb >= 10 and len(launches) == nb
    assert launches[-1] == len(plan.bwd) - 1 or launches[-1] > len(plan.bwd) - 8  # stems finish last
    eng.ddp.begin_backward(eng)
    for i in launches:
        plan.bwd[i][0]()
    assert eng.ddp.bucketer.launch_order[0] == 0  # head-first
    assert sorted(eng.ddp.bucketer.launch_order) == list(range(nb))
    # all live gradients are covered, dead parameters are not in the arena
    covered = sum(n for _, n in eng.ddp.bucketer.spans)
    dead = sum(p.numel() for n, p in m.named_parameters() if ".residual." in n or ".res_unit." in n)
    assert covered == sum(p.numel() for p in eng.live_params) == 38077321 - dead


def _gather_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from video_classification_b200.parallel import gather_part_scores
        n, C = 6, 11
        mine = torch.full((n, C), float(rank)) + torch.arange(n).view(n, 1) * 0.01
        got = gather_part_scores(mine)
        ok = tuple(got.shape) == (n, world, C)
        for r in range(world):
            ok = ok and torch.equal(got[:, r], torch.full((n, C), float(r)) + torch.arange(n).view(n, 1) * 0.01)
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_ensemble_part_scores_gather_gloo_world2():
    """one part model per rank -> (n_clips, P, C) in rank order on every rank (BASELINE config 4 placement)"""
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gather_worker, args=(world, port, ret), nprocs=world, join=True)
    assert all(ret.get(r) for r in range(world)), dict(ret)


def test_ensemble_gather_single_process_is_a_view_with_one_part():
    from video_classification_b200.parallel import gather_part_scores
    x = torch.rand(4, 9)
    assert torch.equal(gather_part_scores(x), x.unsqueeze(1))


def test_tail_bucket_is_split_off_the_last_bucket():
    """the gradients at the end of the arena (stems, first residual stage) are final last: they get a small bucket of
    their own so that only a latency-sized all-reduce is exposed after backward"""
    flat = torch.zeros(4000)
    spans = [(0, 1000), (1000, 1000), (2000, 1000), (3000, 600), (3600, 200), (3800, 100), (3900, 100)]
    b = GradBucketer(flat, spans, bucket_bytes=4 * 2500, tail_bytes=4 * 450)
    # without the tail: [(0,2000), (2000,4000)]; the last 400 elements (3 spans <= 450) are cut off
    assert b.buckets == [(0, 2000), (2000, 3600), (3600, 4000)]
    assert [b.span_bucket[o] for o, _ in spans] == [0, 0, 1, 1, 2, 2, 2]
    b.reset_counts()
    assert b.mark_ready(3900) is None and b.mark_ready(3800) is None and b.mark_ready(3600) == 2
    # the tail never swallows the whole last bucket: at least the bucket's first gradient stays in it
    b2 = GradBucketer(flat, spans, bucket_bytes=4 * 2500, tail_bytes=4 * 5000)
    assert b2.buckets == [(0, 2000), (2000, 3000), (3000, 4000)]
    b3 = GradBucketer(flat, spans, bucket_bytes=4 * 2500)
    assert b3.buckets == [(0, 2000), (2000, 4000)]
