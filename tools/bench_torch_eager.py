"""The library bar on the same B200: the reference's module tree run by PyTorch
eager on the GPU, (a) as the reference runs it -- fp32, NCDHW, cuDNN / cuBLAS / ATen (train.py:122,225-231) -- and
(b) tuned -- bf16 autocast + channels_last_3d -- on the headline workload (HYB graph, 8/32 frames, 224^2) with the same
step definition as bench.py (forward, CrossEntropy, zero_grad, backward, Adam; or the eval forward), CUDA events,
clocks sampled.  NOT the reference arm and not product code: the model is the oracle restatement of the reference
path (this tool lives beside bench.py's CPU leg as a measured baseline).

    python tools/bench_torch_eager.py [--mode train|infer] [--batch 64] [--steps 10] [--warmup 5] [--out FILE]

A batch that does not fit (fp32 eager keeps every intermediate) is halved until it does; clips/s is per the batch
that ran and says so.  Prints / writes one JSON object: {"fp32_ncdhw": {...}, "bf16_channels_last": {...}}.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import bench  # noqa: E402
from oracle import slowfast_oracle as so  # noqa: E402


def run(kind, mode, batch, steps, warmup, dev):
    w = bench.WORKLOADS["hyb"]
    torch.manual_seed(0)
    model = so.build_oracle(num_class=bench.NUM_CLASS, fuse=True, **so.VARIANTS["HYB"])
    so.randomize_bn_(model, 123, 0.1)
    model = model.to(dev)
    bf16 = kind == "bf16_channels_last"
    if bf16:
        model = model.to(memory_format=torch.channels_last_3d)
    train = mode == "train"
    model.train(train)
    opt = torch.optim.Adam(model.parameters(), lr=1e-4, fused=True) if train else None
    B = batch
    while True:
        try:
            g = torch.Generator().manual_seed(1000)
            xs = torch.randn((B, 5, w["ts"], w["size"], w["size"]), generator=g).to(dev)
            xf = torch.randn((B, 15, w["tf"], w["size"], w["size"]), generator=g).to(dev)
            if bf16:
                xs = xs.contiguous(memory_format=torch.channels_last_3d)
                xf = xf.contiguous(memory_format=torch.channels_last_3d)
            y = torch.randint(0, bench.NUM_CLASS, (B,), generator=g).to(dev)

            def step():
                if train:
                    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
                        out = model([xs, xf])
                    loss = F.cross_entropy(out.float(), y)
                    opt.zero_grad()
                    loss.backward()
                    opt.step()
                    return loss
                with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
                    return model([xs, xf])

            for _ in range(warmup):
                step()
            torch.cuda.synchronize(dev)
            break
        except torch.cuda.OutOfMemoryError:
            del xs, xf
            if opt is not None:
                opt.zero_grad(set_to_none=True)
            torch.cuda.empty_cache()
            if B == 1:
                raise
            B //= 2
    sampler = bench.ClockSampler(dev.index or 0)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize(dev)
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1) / steps
    flops = w["f_train"] if train else w["f_fwd"]
    out = {"clips_per_s": B / (ms * 1e-3), "ms_per_step": ms, "batch": B, "requested_batch": batch, "steps": steps,
           "warmup": warmup, "tflops": B / (ms * 1e-3) * flops / 1e3, "peak_mem_gb": torch.cuda.max_memory_allocated(dev) / 1e9,
           "clocks": clocks, "cudnn_benchmark": torch.backends.cudnn.benchmark,
           "what": ("fp32 NCDHW eager, TF32 off, as reference train.py:122,225-231 runs the model" if not bf16 else
                    "torch.autocast(bf16) + channels_last_3d eager, cuDNN bf16 convolutions, fp32 BatchNorm") +
                   ("; step = fwd + CE + zero_grad + bwd + torch.optim.Adam(fused=True)" if train else "; eval forward under no_grad")}
    del model, opt, xs, xf
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats(dev)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="train", choices=["train", "infer"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.backends.cudnn.benchmark = True  # let the library pick its best algorithm per layer
    res = {"workload": bench.workload_name(a.mode, "hyb"), "mode": a.mode, "gpu": torch.cuda.get_device_name(dev),
           "torch": torch.__version__, "cudnn": torch.backends.cudnn.version()}
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    res["fp32_ncdhw"] = run("fp32_ncdhw", a.mode, a.batch, a.steps, a.warmup, dev)
    res["bf16_channels_last"] = run("bf16_channels_last", a.mode, a.batch, a.steps, a.warmup, dev)
    txt = json.dumps(res)
    print(txt, flush=True)
    if a.out:
        with open(a.out, "w") as f:
            f.write(txt + "\n")


if __name__ == "__main__":
    main()
