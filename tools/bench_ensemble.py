"""BASELINE config 4 (config/slowfast-HTAH.yaml): multi-part ensemble, ONE body-part SlowFast per GPU, fused logits.
The reference runs the part models one after another, pickles each part's clip scores and fuses them in a second
program (train_sparse.py:36-86, 109-131, 200-238).  Here rank r holds part r's model; per batch of clips every rank
runs its eval forward, sfb_video_scores (softmax), ONE all-gather of the (n_clips, C) scores over NCCL
(parallel.gather_part_scores) and the SparseModel fusion + per-video mean / arg-max on the device (fuse_ensemble).

launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node P --master-addr 127.0.0.1 --master-port 29517 \
            tools/bench_ensemble.py [--steps K] [--warmup W] [--batch B]
Prints one JSON line on rank 0; time = CUDA events, max over ranks; a step = one batch of B clips through all P parts."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from video_classification_b200 import _ffi  # noqa: E402
from video_classification_b200.config import CLIP_LEN, PART_WORKLOADS, make_cfg  # noqa: E402
from video_classification_b200.evaluation import SparseModel, fuse_ensemble, video_scores  # noqa: E402
from video_classification_b200.my_slowfast import init_my_slowfast  # noqa: E402
from video_classification_b200.parallel import broadcast_parameters, gather_part_scores  # noqa: E402

# the ensemble's parts in the order reference train_sparse.py:36 loads them (+ the optional Torso), each at its own crop
# size (config/crop_cfg.py:18-32); ranks beyond the sixth repeat the list
PARTS = ["HTAH", "LHandArm", "LHand", "RHandArm", "RHand", "Torso", "HTAH", "LHandArm"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--clips-per-video", type=int, default=4)
    ap.add_argument("--shapes", default="ref", choices=["ref", "hyb"],
                    help="ref: the reference's per-part crops (64/128/192^2, 20 frames on both pathways, reference graph); "
                         "hyb: every part at the headline 8/32 x 224^2 shape")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    B, C = a.batch, bench.NUM_CLASS
    torch.manual_seed(100 + rank)  # every part has its own weights
    part = PARTS[rank % len(PARTS)]
    if a.shapes == "ref":
        size, ts, tf, kw = PART_WORKLOADS[part]["size"], CLIP_LEN, CLIP_LEN, {}
    else:
        size, ts, tf, kw = bench.SIZE, bench.T_SLOW, bench.T_FAST, bench.HYB_KW
    model = init_my_slowfast(make_cfg(C, True), (5, 15), (64, 8), **kw).to(dev).eval()
    torch.manual_seed(7)
    fuser = SparseModel(C, world).to(dev).eval()
    if world > 1:
        broadcast_parameters(fuser, 0)
    g = torch.Generator().manual_seed(1000 + rank)  # each part sees its own crop of the same clips
    xs = torch.randn((B, 5, ts, size, size), generator=g).to(dev)
    xf = torch.randn((B, 15, tf, size, size), generator=g).to(dev)
    labels = torch.randint(0, C, (B // a.clips_per_video,), generator=torch.Generator().manual_seed(5))
    labels = labels.repeat_interleave(a.clips_per_video).to(dev)
    sv = [a.clips_per_video] * (B // a.clips_per_video)

    def step():
        with torch.no_grad():
            logits = model([xs, xf])
            own = video_scores(logits, sv, softmax=True)["clip_probs"]
            ps = gather_part_scores(own)
        return own, ps, fuse_ensemble(fuser, ps, labels, sv)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(a.warmup, 3)):
        own, ps, res = step()
    # the gathered block of this rank is its own score matrix, bit for bit; every rank fuses the same tensor
    assert ps.shape == (B, world, C) and torch.equal(ps[:, rank, :], own)
    if world > 1:
        chk = ps.double().sum().reshape(1)
        all_chk = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(all_chk, chk)
        assert all(float(c) == float(all_chk[0]) for c in all_chk), "ranks fused different score tensors"
    barrier()
    l0 = _ffi.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        own, ps, res = step()  # fuse_ensemble's result() reads accuracy back: one sync per batch, as run_eval has
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = float(ms)
        print(json.dumps({
            "metric": "ensemble clips/sec, one body-part SlowFast-R50 per GPU, fused logits",
            "shapes": a.shapes, "part_sizes": {p_: (PART_WORKLOADS[p_]["size"] if a.shapes == "ref" else bench.SIZE)
                                               for p_ in PARTS[:world]},
            "value": B / (ms * 1e-3), "unit": "clips/s", "n_gpus": world, "parts": PARTS[:world], "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": ms, "part_forwards_per_s": world * B / (ms * 1e-3),
            "config": {"workload": "configs[3]: slowfast-HTAH ensemble, eval forward per part + softmax + all-gather "
                                   "(n_clips, C) + SparseModel fusion + per-video mean / arg-max",
                       "batch": B, "clips_per_video": a.clips_per_video, "num_class": C,
                       "allgather_bytes_per_rank": B * C * 4},
            "dtype": "bf16", "data": "synthetic", "gpu_launches": _ffi.launch_count() - l0,
            "acc_random_weights": float(res["acc"])}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
