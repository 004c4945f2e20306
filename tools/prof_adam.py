"""sfb_adam_step alone on the live-parameter shapes of the benchmark model (332 tensors, 34.05 M elements): CUDA-event
time of the kernel (replayed tables, no host planning) and, under ncu, the capture target for profiles/.
usage: python tools/prof_adam.py [reps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import slowfast_oracle as so  # noqa: E402  (only make_cfg)
from video_classification_b200.my_slowfast import init_my_slowfast  # noqa: E402
from video_classification_b200.optim import Adam  # noqa: E402
import bench  # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    torch.manual_seed(0)
    model = init_my_slowfast(so.make_cfg(bench.NUM_CLASS, True), (5, 15), (64, 8), **bench.HYB_KW).cuda()
    live = [p for n, p in model.named_parameters()]
    arena = torch.randn(sum((p.numel() + 7) // 8 * 8 for p in live), device="cuda") * 1e-3
    o = 0
    for p in live:  # gradients as views of one flat arena, as the engine hands them over
        p.grad = arena[o:o + p.numel()].view(p.shape)
        o += (p.numel() + 7) // 8 * 8
    opt = Adam(model.parameters(), lr=1e-4)
    opt.step()
    n = sum(p.numel() for p in live)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        opt.replay_last()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        opt.replay_last()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"adam_step_kernel: {len(live)} tensors, {n} elements, {ms:.4f} ms per launch, "
          f"{n * 28 / ms / 1e6:.0f} GB/s algorithmic (16 B read + 12 B written per element)")
    # the drop-in call including its host side
    import time
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        opt.step()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"optim.Adam.step() host time: {(t1 - t0) / reps * 1e3:.3f} ms per call")


if __name__ == "__main__":
    main()
