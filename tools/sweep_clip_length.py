"""BASELINE.json configs[4]: `config/res2d.yaml` 2-D ResNet-50 per-frame baseline vs SlowFast, throughput sweep over clip
length T = 8 ... 64 at 224^2.

  * SlowFast side: the drop-in model on the libsfb200 kernels, HYB rule of the headline graph (Fast pathway sees all T
    frames, Slow pathway T/4, lateral stride 4, head pools (T/4,7,7) / (T,7,7)); training step (fwd + CE + bwd + Adam,
    reference train.py:225-231) and eval forward, clips resident in HBM, CUDA events.
  * res2d side (baseline only -- no new kernels for the 2-D model): the reference's
    `_init_res2d_model` (train.py:64-76): torchvision ResNet-50 whose conv1 is replaced by
    Conv2d(T*5, 64, 7, stride 2, padding 3), fed the clip's frames stacked on the channel axis (N, T*5, H, W), run by
    PyTorch on the same GPU -- fp32 as the reference runs it and bf16 autocast + channels_last.  1000-way head as in the
    reference (it keeps torchvision's fc).

    python tools/sweep_clip_length.py [--T 8 16 32 64] [--batch 64] [--steps 5] [--warmup 3] [--out FILE]
One JSON object per T on stdout (and in --out as JSON lines).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import bench  # noqa: E402
from video_classification_b200.config import make_cfg  # noqa: E402
from video_classification_b200.my_slowfast import init_my_slowfast  # noqa: E402
from video_classification_b200.optim import Adam  # noqa: E402

DEV = torch.device("cuda", 0)
S = 224


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(DEV)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(DEV)
    return e0.elapsed_time(e1) / steps


def slowfast(T, B, steps, warmup):
    ts = max(T // 4, 1)
    kw = dict(fusion_conv_stride=(4, 1, 1), head_pool_kernel_sizes=((ts, 7, 7), (T, 7, 7)))
    torch.manual_seed(0)
    model = init_my_slowfast(make_cfg(bench.NUM_CLASS, True), (5, 15), (64, 8), **kw).to(DEV)
    g = torch.Generator().manual_seed(1)
    xs = torch.randn((B, 5, ts, S, S), generator=g).to(DEV)
    xf = torch.randn((B, 15, T, S, S), generator=g).to(DEV)
    y = torch.randint(0, bench.NUM_CLASS, (B,), generator=g).to(DEV)
    opt = Adam(model.parameters(), lr=1e-4)
    model.train()

    def train_step():
        loss = F.cross_entropy(model([xs, xf]), y)
        opt.zero_grad()
        loss.backward()
        opt.step()

    ms_t = timed(train_step, steps, warmup)
    model.eval()

    def infer_step():
        with torch.no_grad():
            model([xs, xf])

    ms_i = timed(infer_step, steps, warmup)
    mem = torch.cuda.max_memory_allocated(DEV) / 1e9
    del model, opt, xs, xf
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats(DEV)
    return {"batch": B, "train_clips_per_s": B / ms_t * 1e3, "train_ms": ms_t, "infer_clips_per_s": B / ms_i * 1e3,
            "infer_ms": ms_i, "peak_mem_gb": mem, "slow_frames": ts, "fast_frames": T}


def res2d(T, B, steps, warmup, bf16):
    import torchvision
    torch.manual_seed(0)
    model = torchvision.models.resnet50(weights=None)
    model.conv1 = nn.Conv2d(T * 5, 64, kernel_size=(7, 7), stride=(2, 2), padding=(3, 3), bias=False)  # train.py:66
    model = model.to(DEV)
    x = torch.randn((B, T * 5, S, S), generator=torch.Generator().manual_seed(1)).to(DEV)
    if bf16:
        model = model.to(memory_format=torch.channels_last)
        x = x.contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, bench.NUM_CLASS, (B,), generator=torch.Generator().manual_seed(2)).to(DEV)
    opt = torch.optim.Adam(model.parameters(), lr=1e-4, fused=True)
    model.train()

    def train_step():
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            out = model(x)
        loss = F.cross_entropy(out.float(), y)
        opt.zero_grad()
        loss.backward()
        opt.step()

    ms_t = timed(train_step, steps, warmup)
    model.eval()

    def infer_step():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            model(x)

    ms_i = timed(infer_step, steps, warmup)
    del model, opt, x
    torch.cuda.empty_cache()
    return {"batch": B, "train_clips_per_s": B / ms_t * 1e3, "train_ms": ms_t, "infer_clips_per_s": B / ms_i * 1e3,
            "infer_ms": ms_i}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--T", type=int, nargs="+", default=[8, 16, 32, 64])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    torch.cuda.set_device(DEV)
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    lines = []
    for T in a.T:
        B = a.batch if T <= 32 else max(a.batch // 2, 1)  # T = 64 doubles every Fast-pathway activation
        row = {"config": "configs[4]: res2d per-frame baseline vs SlowFast, 224^2", "T": T, "gpu": torch.cuda.get_device_name(DEV),
               "slowfast_b200": slowfast(T, B, a.steps, a.warmup),
               "res2d_torch_fp32": res2d(T, B, a.steps, a.warmup, False),
               "res2d_torch_bf16_channels_last": res2d(T, B, a.steps, a.warmup, True)}
        txt = json.dumps(row)
        print(txt, flush=True)
        lines.append(txt)
    if a.out:
        with open(a.out, "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
