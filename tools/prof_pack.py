"""Times the batched operand re-pack of one training step (sfb_pack_run: engine.refresh_weights after an optimizer
update) on the headline model, with CUDA events (launches back to back, so the events see kernel time).  Measured r3w: 0.54 ms
for 152 MB of fp32 weights; a variant that packed through 32 x 32 transposing shared-memory tiles was slower (0.63 ms:
45 KB of shared memory per block and mostly idle threads on the many narrow layers) and was dropped.

    python tools/prof_pack.py [--reps 20]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from video_classification_b200 import _ffi, my_slowfast  # noqa: E402
from video_classification_b200.config import make_cfg  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    model = my_slowfast.init_my_slowfast(make_cfg(bench.NUM_CLASS, True), (5, 15), (64, 8), **bench.HYB_KW).to(dev)
    model.train()
    xs = torch.randn(2, 5, 8, 224, 224, device=dev)
    xf = torch.randn(2, 15, 32, 224, 224, device=dev)
    model([xs, xf]).sum().backward()          # per-layer path allocates every operand
    eng = model._engine
    params = list(model.parameters())
    with torch.no_grad():
        torch._foreach_add_(params, 0.0)      # new version counters, as after an optimizer step
    eng._bwd_since_pack = True
    eng.refresh_weights(True)                 # builds the job tables
    tabs = eng._pack_tables(True)
    assert tabs is not None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    times = []
    for _ in range(3):
        for tab, n, nblk in tabs["levels"]:
            _ffi.check(eng.lib.sfb_pack_run(tab.data_ptr(), n, nblk, eng.stream), "sfb_pack_run")
    torch.cuda.synchronize(dev)
    for _ in range(3):
        e0.record()
        for _ in range(a.reps):               # back to back: the launches queue up, the events see kernel time
            for tab, n, nblk in tabs["levels"]:
                _ffi.check(eng.lib.sfb_pack_run(tab.data_ptr(), n, nblk, eng.stream), "sfb_pack_run")
        e1.record()
        torch.cuda.synchronize(dev)
        times += [e0.elapsed_time(e1) / a.reps] * 2
    times = [0, 0, 0] + times
    times = sorted(times[3:])
    nbytes = sum(p.numel() for p in params) * 4
    print(json.dumps({"what": "sfb_pack_run over the engine's job tables (all conv operands of a training step), launches back to back", "reps": a.reps,
                      "ms_median": times[len(times) // 2], "ms_min": times[0], "fp32_weight_bytes": nbytes}))


if __name__ == "__main__":
    main()
