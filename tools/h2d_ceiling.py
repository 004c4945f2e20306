"""Plain host->device copy ceiling of the box, all ranks at once (VERDICT r1 item 4 / reference train.py:125-145: the
`.cuda()` of every batch).  One process per GPU (launch with torch.distributed.run), each copies a pinned host buffer to
its GPU with cudaMemcpyAsync (torch `.copy_(non_blocking=True)`), all ranks started together behind a barrier; reports
GB/s per rank and the aggregate, for each rank alone as well (rank r copies while the others idle) so that a shared
host-side limit (PCIe switch uplinks, memory channels of one socket) shows up as aggregate < sum of solo rates.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tools/h2d_ceiling.py [--mb 1024] [--reps 10] [--bind]
--bind pins the rank's threads to the CPUs of its GPU's NUMA node before the buffer is allocated (bench.py does the same).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--bind", action="store_true")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    numa = bench.bind_to_gpu_numa_node(local) if a.bind else {"numa_node": None, "note": "not requested"}
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n = a.mb << 20
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    host.fill_(rank + 1)  # first touch on this rank's (bound) CPUs
    devb = torch.empty(n, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def run():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        devb.copy_(host, non_blocking=True)
        barrier()
        e0.record()
        for _ in range(a.reps):
            devb.copy_(host, non_blocking=True)
        e1.record()
        torch.cuda.synchronize(dev)
        return n * a.reps / (e0.elapsed_time(e1) * 1e-3) / 1e9

    together = run()
    barrier()
    solo = 0.0
    for r in range(world):  # one rank at a time
        barrier()
        if r == rank:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.reps):
                devb.copy_(host, non_blocking=True)
            e1.record()
            torch.cuda.synchronize(dev)
            solo = n * a.reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
        barrier()
    vals = torch.tensor([together, solo], device=dev, dtype=torch.float64)
    allv = [torch.empty_like(vals) for _ in range(world)]
    if world > 1:
        dist.all_gather(allv, vals)
    else:
        allv = [vals]
    if rank == 0:
        tg = [float(v[0]) for v in allv]
        so = [float(v[1]) for v in allv]
        print(json.dumps({"what": "pinned host -> device cudaMemcpyAsync, GB/s", "n_gpus": world, "mb_per_copy": a.mb,
                          "reps": a.reps, "bind": bool(a.bind), "host_affinity_rank0": numa,
                          "all_ranks_together_gbs": [round(v, 2) for v in tg], "aggregate_together_gbs": round(sum(tg), 2),
                          "each_rank_alone_gbs": [round(v, 2) for v in so], "sum_of_solo_gbs": round(sum(so), 2),
                          "host_cpus": len(os.sched_getaffinity(0)), "cpu": bench.cpu_model_string()}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
