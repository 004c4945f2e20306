#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on its config: train clips/s of SlowFast-R50 8x8 (slow 8 / fast 32 frames, alpha=4,
beta=1/8, 224^2, 5+15 input channels, 249 classes), bf16 storage / fp32 accumulate, batch 64 per GPU, data parallel.

    python bench.py --gpus N --steps K --warmup W          our arm (libsfb200 kernels through the drop-in model)
    python bench.py --impl reference ...                    the reference's own CPU path (oracle restatement) on host cores
    python bench.py --mode infer [--batch 1|8|64]           BASELINE configs[1]: single-part inference (eval forward)
    python bench.py --workload ref_lhand                    the reference's own shapes (slowfast-LHand.yaml: 300 x 20 x 64^2)

A training step is what reference train.py:225-231 does per batch: forward, CrossEntropy, zero_grad, backward (+ bucketed
gradient all-reduce when N > 1), Adam step; an inference step is train.py:304-306 (eval forward under no_grad).
`value` keeps the clips resident in HBM; `e2e` copies every step's fp32 clips from pinned host memory (double-buffered
on a copy stream) and reads the result back.  One JSON line on stdout.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NUM_CLASS = 249
HYB_KW = dict(fusion_conv_stride=(4, 1, 1), head_pool_kernel_sizes=((8, 7, 7), (32, 7, 7)))
# workload constants (SURVEY.md 8d / BASELINE.md 2): algorithmic conv+FC GFLOP per clip, 1 MAC = 2 FLOP
WORKLOADS = {
    "hyb": dict(graph_kw=HYB_KW, oracle="HYB", ts=8, tf=32, size=224, batch=64, f_fwd=100.616, f_train=293.98,
                label="slowfast-LHand SlowFast-R50, slow 5x8x224^2 + fast 15x32x224^2 (alpha=4 lateral), 249 classes"),
    # the reference exactly as written: both pathways see the same 20 frames (train.py:136-140), config/slowfast-*.yaml
    "ref_lhand": dict(graph_kw={}, oracle="REF", ts=20, tf=20, size=64, batch=300, f_fwd=18.936, f_train=55.93,
                      label="reference graph as written, slowfast-LHand.yaml: slow 5x20x64^2 + fast 15x20x64^2, batch 300"),
    "ref_htah": dict(graph_kw={}, oracle="REF", ts=20, tf=20, size=192, batch=55, f_fwd=170.737, f_train=504.3,
                     label="reference graph as written, slowfast-HTAH.yaml: slow 5x20x192^2 + fast 15x20x192^2, batch 55"),
    # BASELINE configs[0] (config/res3d.yaml): Slow-only 3D-ResNet-50 (pytorchvideo slow_r50 with the reference's 5-channel
    # stem, train.py:79-83) on 16 x 112^2 clips, batch 2 -- the reference's own CPU-runnable case; head pool (4,4,4)
    "res3d": dict(graph_kw={}, oracle="RES3D", ts=16, tf=0, size=112, batch=2, f_fwd=44.6045, f_train=132.240,
                  label="config/res3d.yaml Slow-only 3D-ResNet-50 (slow_r50, 5-channel stem), 5x16x112^2, batch 2, 400 classes"),
}
# kept for tools/ that import them
T_SLOW, T_FAST, SIZE = 8, 32, 224
F_FWD_GF, F_TRAIN_GF = 100.616, 293.98


def metric_name(mode, wl):
    if wl == "res3d":
        return f"{'train' if mode == 'train' else 'inference'} clips/sec 3D-ResNet-50 (slow_r50) 16x112^2"
    shape = "8x8 224^2" if wl == "hyb" else f"{WORKLOADS[wl]['ts']}x{WORKLOADS[wl]['size']}^2 (reference shapes)"
    return f"{'train' if mode == 'train' else 'inference'} clips/sec SlowFast-R50 {shape}"


def workload_name(mode, wl):
    w = WORKLOADS[wl]
    if wl == "hyb":
        return ("configs[2]: " + w["label"].replace("SlowFast-R50,", "SlowFast-R50 training,")) if mode == "train" else \
            ("configs[1]: " + w["label"].replace("SlowFast-R50,", "SlowFast-R50 single-part inference,"))
    if wl == "res3d":
        return "configs[0]: " + w["label"] + (", fwd+bwd+Adam" if mode == "train" else ", eval forward")
    return ("SURVEY 8(d) config 3 (REF shapes), training: " if mode == "train" else "SURVEY 8(d) config 2 (REF row), inference: ") + w["label"]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], burst=d["bf16_tflops"], sustained=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, burst=1590.0, sustained=1400.0, src="fallback")


def cpu_model_string():
    try:
        with open("/proc/cpuinfo") as f:
            for ln in f:
                if ln.lower().startswith("model name"):
                    return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


class ClockSampler(threading.Thread):
    """samples SM clock / throttle reasons of one GPU with NVML every 200 ms while the timed region runs"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            self._halt.wait(0.2)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own path (fp32 eager PyTorch on host cores).  The reference is pure Python and cannot
# travel to the GPU box; what is timed is the oracle restatement (oracle/slowfast_oracle.py), which
# tests/test_oracle.py pins bit-identical to the verbatim import of the reference's model/my_slowfast.py.
# ------------------------------------------------------------------------------------------------------------------
def cpu_steps(mode, wl, n_clips, steps, warmup, threads=None):
    """-> (seconds per step, threads used).  One step = train.py:225-231 (train) or :304-306 (infer) on n_clips clips."""
    import torch
    import torch.nn.functional as F
    from oracle import slowfast_oracle as so
    # explicit: under torch.distributed.run the launcher exports OMP_NUM_THREADS=1, which would time ONE thread
    threads = int(threads or len(os.sched_getaffinity(0)) or os.cpu_count() or 1)
    torch.set_num_threads(threads)
    threads = torch.get_num_threads()
    w = WORKLOADS[wl]
    torch.manual_seed(0)
    if w["oracle"] == "RES3D":
        model = so.build_res3d_oracle(num_class=400, input_channel=5, head_pool_kernel_size=(4, 4, 4))
        so.randomize_bn_(model, 123, 0.1)
        x1 = torch.randn((n_clips, 5, w["ts"], w["size"], w["size"]), generator=torch.Generator().manual_seed(1))
        inp = x1
        ncls = 400
    else:
        model = so.build_oracle(num_class=NUM_CLASS, fuse=True, **so.VARIANTS[w["oracle"]])
        so.randomize_bn_(model, 123, 0.1)
        xs, xf = so.synthetic_clips(n_clips, w["ts"], w["tf"], w["size"], seed=1)
        inp = None
        ncls = NUM_CLASS
    y = torch.randint(0, ncls, (n_clips,), generator=torch.Generator().manual_seed(7))
    times = []
    if mode == "train":
        model.train()
        opt = torch.optim.Adam(model.parameters(), lr=1e-4)
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            out = model(inp if inp is not None else [xs, xf])
            loss = F.cross_entropy(out, y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            float(loss.detach())
            if it >= warmup:
                times.append(time.perf_counter() - t0)
    else:
        model.eval()
        with torch.no_grad():
            for it in range(warmup + steps):
                t0 = time.perf_counter()
                out = model(inp if inp is not None else [xs, xf])
                int(out.argmax(1)[0])
                if it >= warmup:
                    times.append(time.perf_counter() - t0)
    return sum(times) / len(times), threads


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # one CPU measurement per box; other ranks exit 0 without work
    n = args.cpu_clips
    sec, threads = cpu_steps(args.mode, args.workload, n, args.steps, args.warmup)
    v = n / sec
    what = "fwd+CE+bwd+Adam" if args.mode == "train" else "eval forward"
    line = {"impl": "reference", "metric": metric_name(args.mode, args.workload), "value": v, "unit": "clips/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": workload_name(args.mode, args.workload), "batch_per_step": n,
                       "note": f"bounded sample: each step runs {n} clips instead of the GPU arm's batch (CPU throughput is "
                               f"batch-insensitive); {what}; {threads} threads on '{cpu_model_string()}'"},
            "cpu_baseline": {"value": v, "unit": "clips/s", "cores": threads, "kind": "port", "cpu": cpu_model_string(),
                             "sample": f"{args.steps} steps ({what}) of {n} clips after {args.warmup} warm-up, fp32 eager "
                                       "PyTorch CPU, oracle restatement pinned bit-identical to the reference's verbatim "
                                       "my_slowfast.py"},
            "e2e": {"value": v, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------------------
_ORIG_AFFINITY = None


def bind_to_gpu_numa_node(local):
    """Pin this rank's threads (and, through first-touch, its pinned host buffers) to the CPUs of the NUMA node its
    GPU hangs off, BEFORE anything is allocated: with 8 ranks copying 6.7 GB per step each, buffers on the far socket
    halve the aggregate host->device rate.  Best effort; returns a short description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        path = f"/sys/bus/pci/devices/{bus.lower()[-12:]}/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return {"numa_node": None, "note": "no NUMA affinity reported for the GPU"}
        cpus = []
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        global _ORIG_AFFINITY
        allowed = os.sched_getaffinity(0)
        _ORIG_AFFINITY = set(allowed)
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "cpus": len(cpus)}
    except Exception as e:  # noqa: BLE001
        return {"numa_node": None, "note": f"not bound ({type(e).__name__})"}


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from video_classification_b200 import _ffi
    from video_classification_b200.config import make_cfg
    from video_classification_b200.my_slowfast import init_my_slowfast
    from video_classification_b200.parallel import attach_ddp
    from video_classification_b200.optim import Adam

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N")
    numa = bind_to_gpu_numa_node(local) if not args.no_numa_bind else {"numa_node": None, "note": "disabled"}
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    w = WORKLOADS[args.workload]
    train = args.mode == "train"
    B = args.batch or w["batch"]
    TS, TF, S = w["ts"], w["tf"], w["size"]
    torch.manual_seed(0)
    single = args.workload == "res3d"
    ncls = 400 if single else NUM_CLASS
    if single:
        from video_classification_b200.res3d import slow_r50
        model = slow_r50(5, ncls, (4, 4, 4)).to(dev)
    else:
        model = init_my_slowfast(make_cfg(NUM_CLASS, True), (5, 15), (64, 8), **w["graph_kw"]).to(dev)
    model.train(train)
    if world > 1 and train:
        attach_ddp(model, bucket_mb=args.bucket_mb, tail_mb=args.tail_mb)
    # same update rule as the reference's optim.Adam (train.py:182), same constructor: sfb_adam_step updates all 332
    # live parameters in ONE launch (csrc/adam.cuh) instead of torch's 10 + 5 fused multi-tensor launches
    opt = Adam(model.parameters(), lr=1e-4) if train else None
    g = torch.Generator().manual_seed(1000 + rank)
    # synthetic clips, per-rank seed; resident copies for `value`, a pinned host copy for `e2e`
    host = [torch.randn((B, 5, TS, S, S), generator=g).pin_memory()]
    if not single:
        host.append(torch.randn((B, 15, TF, S, S), generator=g).pin_memory())
    y = torch.randint(0, ncls, (B,), generator=g).to(dev)
    resident = [h.to(dev) for h in host]
    h2d_bytes = sum(h.numel() * 4 for h in host)

    if train:
        def step(*xs):
            out = model(xs[0] if single else list(xs))
            loss = F.cross_entropy(out, y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            return loss
        d2h_bytes = 4
    else:
        def step(*xs):
            with torch.no_grad():
                return model(xs[0] if single else list(xs))
        d2h_bytes = B * ncls * 4

    def read_back(res):
        """device -> host read of the step's result: the loss (train.py:232) or the logits (train.py:306)"""
        return float(res.detach()) if train else res.cpu()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = _ffi.launch_count()
        # `ncu --nvtx --nvtx-include timed_all` lists exactly the timed launches (a start/end range also covers the
        # backward kernels autograd launches from its own thread; the push/pop range `timed/` only sees this thread)
        rid = torch.cuda.nvtx.range_start("timed_all")
        torch.cuda.nvtx.range_push("timed")
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.nvtx.range_pop()
        torch.cuda.nvtx.range_end(rid)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / steps, _ffi.launch_count() - l0

    # ---- value: inputs resident in HBM ----
    for _ in range(args.warmup):
        step(*resident)
    sampler = ClockSampler(local)
    sampler.start()
    ms_step, launches = timed(lambda: step(*resident), args.steps)
    clocks = sampler.stop()
    last = step(*resident)
    loss_val = float(last.detach()) if train else None
    r = dict(rank=rank, world=world, dev=dev, model=model, ms_step=ms_step, launches=launches, clocks=clocks,
             e2e_ms=None, e2e_u8=None, h2d_bytes=h2d_bytes, d2h_bytes=d2h_bytes, loss=loss_val, step=step,
             resident=resident, B=B, opt=opt, numa=numa, replica_check=None, train=train)

    # ---- replicas must still hold identical parameters after the timed steps (N > 1) ----
    if world > 1 and train:
        with torch.no_grad():
            flat = torch.cat([p.detach().flatten() for p in model.parameters()])
            chk = torch.stack([flat.view(torch.int32).to(torch.int64).sum(), flat.double().sum().view(torch.int64)])
        allc = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        same = all(bool((c == allc[0]).all()) for c in allc)
        r["replica_check"] = {"ranks": world, "param_checksum_equal": same,
                              "what": "int64 sum of the fp32 bit patterns + bit pattern of the double sum of all parameters, "
                                      "all-gathered after the timed steps"}
        if not same:
            raise SystemExit(f"rank {rank}: data-parallel replicas diverged (parameter checksums differ across ranks)")

    # ---- optional: where the collectives sit inside one backward pass (per-bucket CUDA events, eager launches) ----
    if world > 1 and train and args.ddp_timeline:
        eng = model._engine
        bk = eng.ddp.bucketer
        saved = eng.use_graphs
        eng.use_graphs = False
        step(*resident)
        bk.timeline = []
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        barrier()
        out = model(list(resident))
        loss = F.cross_entropy(out, y)
        opt.zero_grad()
        t0.record()
        loss.backward()
        t1.record()
        torch.cuda.synchronize(dev)
        rows = [{"bucket": b, "bytes": nbytes, "ready_ms": t0.elapsed_time(ev), "allreduce_start_ms": t0.elapsed_time(e0),
                 "allreduce_end_ms": t0.elapsed_time(e1)} for b, nbytes, ev, e0, e1 in bk.timeline]
        bk.timeline = None
        eng.use_graphs = saved
        r["ddp_timeline"] = {"backward_ms": t0.elapsed_time(t1), "buckets": rows,
                             "note": "rank 0, one eager backward pass: ready = the bucket's last gradient is final (compute "
                                     "stream), all-reduce start / end on the communication stream; backward_ms ends when the "
                                     "compute stream has waited for the last collective"}
        opt.step()

    if args.skip_e2e:  # profiling runs only (ncu launch lists); the JSON line then carries no e2e number
        return r
    # ---- e2e: pinned host -> device every step (double-buffered prefetch on a copy stream) + result read-back ----
    copy_stream = torch.cuda.Stream(dev)

    def run_e2e(host_tensors, dbuf, run_step):
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        state = {"i": 0}

        def prefetch(slot):
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[slot])
                for d, h in zip(dbuf[slot], host_tensors):
                    d.copy_(h, non_blocking=True)
                ready[slot].record(copy_stream)

        for s_ in range(2):
            consumed[s_].record()

        def e2e_step():
            i = state["i"]
            slot = i & 1
            if i == 0:
                prefetch(0)
            prefetch(slot ^ 1)  # next step's clips travel while this step computes
            torch.cuda.current_stream().wait_event(ready[slot])
            res = run_step(dbuf[slot])
            consumed[slot].record()
            read_back(res)
            state["i"] = i + 1

        for _ in range(2):
            e2e_step()
        ms, _ = timed(e2e_step, args.steps)
        torch.cuda.synchronize(dev)
        return ms

    dbuf = [[torch.empty_like(t) for t in resident] for _ in range(2)]
    r["e2e_ms"] = run_e2e(host, dbuf, lambda bufs: step(*bufs))
    del dbuf

    # ---- e2e_u8: the same step fed uint8 frames as the dataset holds them before ToTensor/Normalize:
    # (B, T, S, S, 21) uint8 from pinned host memory, normalised + split + laid out by sfb_stage_clips_u8 ----
    if single:  # the uint8 staging kernel feeds the two SlowFast stems; the res3d loader path has no such split
        return r
    from video_classification_b200.staging import RawClips
    host_u8 = [torch.randint(0, 256, (B, TF, S, S, 21), generator=g, dtype=torch.uint8).pin_memory()]
    dbuf_u8 = [[torch.empty_like(h, device=dev) for h in host_u8] for _ in range(2)]

    def step_u8(bufs):
        raw = RawClips(bufs[0], slow_t_stride=TF // TS)
        if not train:
            with torch.no_grad():
                return model(raw)
        out = model(raw)
        loss = F.cross_entropy(out, y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        return loss

    r["e2e_u8"] = dict(ms=run_e2e(host_u8, dbuf_u8, step_u8), h2d_bytes=host_u8[0].numel())
    del dbuf_u8, host_u8
    return r


def profile_adam(r, peaks, reps=5):
    """sfb_adam_step alone, CUDA events on the launching stream; its 28 B/element working set (0.95 GB) exceeds L2."""
    import torch
    opt = r["opt"]
    n = sum(p.numel() for g in opt.param_groups for p in g["params"] if p.grad is not None)
    saved = [[p.detach().clone() for p in g["params"]] for g in opt.param_groups]  # timing must not train further
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    opt.replay_last()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        opt.replay_last()  # the kernel alone: the same tables, no host-side planning
    e1.record()
    torch.cuda.synchronize()
    with torch.no_grad():
        for g, sv in zip(opt.param_groups, saved):
            torch._foreach_copy_(list(g["params"]), sv)
    ms = e0.elapsed_time(e1) / reps
    gbs = n * 28 / ms / 1e6
    return {"kernel": "adam_step_kernel (sfb_adam_step)", "launches_per_step": 1, "ms": ms, "elements": n,
            "algorithmic_bytes_per_launch": n * 28, "bound": "hbm", "achieved": gbs, "peak": peaks.get("hbm"),
            "unit": "GB/s", "frac": gbs / peaks["hbm"] if peaks.get("hbm") else None}


def profile_kernels(r, peaks, steps=2):
    """per-plan-entry CUDA-event timing over `steps` extra steps (events on the launching stream)."""
    import torch
    eng = r["model"]._engine
    eng.prof = []
    for _ in range(steps):
        r["step"](*r["resident"])
    table, flops = eng.profile_summary()
    if os.environ.get("SFB_DUMP_OPS"):
        with open(os.environ["SFB_DUMP_OPS"], "w") as f:
            for name, desc, ms, gf in eng.profile_ops()[: len(eng.prof) // steps]:
                f.write(f"{ms:9.4f} ms {gf / ms if ms > 0 else 0:9.1f} TF/s {name:20s} {desc}\n")
    eng.prof = None
    torch.cuda.synchronize()
    per_step = {k: {"launches": n // steps, "ms": t / steps} for k, (n, t) in table.items()}
    for k, f in flops.items():
        if k.endswith("#bytes"):
            per_step[k[:-6]]["algorithmic_bytes"] = f / steps
            continue
        if k.endswith("#kernels"):
            per_step[k[:-8]]["kernel_launches"] = f / steps
            continue
        per_step[k]["algorithmic_tflop"] = f / steps / 1e12
        per_step[k]["tflops"] = f / steps / 1e9 / per_step[k]["ms"]
    return per_step


def conv_traffic_from_ncu():
    """DRAM bytes (read + write) per conv_tma_kernel LAUNCH, averaged over the launches of one training step, from the
    newest committed ncu pass `profiles/r*_traffic.json` (tools/summarize_traffic.py; measured under ncu with cold
    caches, so a static input of this line, never a timing)"""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")))
    if not files:
        return {}
    try:
        with open(files[-1]) as f:
            tab = json.load(f)
    except (OSError, ValueError):
        return {}
    rows = [v for k, v in tab.items() if k.startswith("conv_tma_kernel")]
    n = sum(v["launches"] for v in rows)
    if not n:
        return {}
    byt = sum(v["dram_read_bytes"] + v["dram_write_bytes"] for v in rows)
    return {"traffic": byt / n, "traffic_unit": "DRAM bytes per kernel launch (mean over the training step's launches)",
            "traffic_launches": n, "traffic_source": os.path.basename(files[-1])}


def library_baseline_from_profiles():
    """PyTorch eager on the same B200 (the 'library kernel to beat'): numbers of the newest committed
    run of tools/bench_torch_eager.py (profiles/r*_torch_eager.json); a static input of this line, measured on another
    box of the same pool -- re-measure with `python tools/bench_torch_eager.py`."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_torch_eager.json")))
    if not files:
        return None
    try:
        with open(files[-1]) as f:
            d = json.load(f)
    except (OSError, ValueError):
        return None
    d["source"] = os.path.basename(files[-1])
    return d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="train", choices=["train", "infer"])
    ap.add_argument("--workload", default="hyb", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="clips per GPU (default: the workload's, 64 for the headline)")
    ap.add_argument("--bucket-mb", type=float, default=25.0)
    ap.add_argument("--cpu-clips", type=int, default=2, help="clips per CPU step (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-numa-bind", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs only: skip the host-buffer leg")
    ap.add_argument("--ddp-timeline", action="store_true", help="N > 1: add per-bucket all-reduce timing of one backward pass")
    ap.add_argument("--tail-mb", type=float, default=1.0, help="N > 1: size of the split-off last gradient bucket")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    args.warmup = max(args.warmup, 3)
    peaks = load_peaks()
    r = run_gpu_arm(args)
    world, B, train = r["world"], r["B"], r["train"]
    w = WORKLOADS[args.workload]
    flops_clip = w["f_train"] if train else w["f_fwd"]
    kernels = None
    adam = None
    if not args.no_profile:
        kernels = profile_kernels(r, peaks)
        adam = profile_adam(r, peaks) if train else None
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if r["rank"] != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    clips_s = world * B / (r["ms_step"] * 1e-3)
    e2e_clips_s = world * B / (r["e2e_ms"] * 1e-3) if r["e2e_ms"] else None
    tf_gpu = clips_s / world * flops_clip / 1e3
    step_desc = "fwd + CrossEntropy + bwd + bucketed grad all-reduce + Adam (reference train.py:225-231)" if train else \
        "eval forward under no_grad (reference train.py:304-306), BatchNorm folded, no dropout"
    line = {"metric": metric_name(args.mode, args.workload), "value": clips_s, "unit": "clips/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args.mode, args.workload), "batch_per_gpu": B, "global_batch": world * B,
                       "step": step_desc,
                       "parallelism": f"dp{world}", "l2": "working set (GBs of activations per step) exceeds the 126 MB L2"
                       if B * w["size"] ** 2 * w["tf"] > 5e7 else "small batch: activations may stay L2-resident between kernels",
                       "weights": "random init (seed 0)", "loss_after_bench": r["loss"], "host_affinity": r["numa"]},
            "e2e": {"value": e2e_clips_s, "unit": "clips/s", "ms_per_step": r["e2e_ms"],
                    "h2d_bytes_per_step": r["h2d_bytes"], "d2h_bytes_per_step": r["d2h_bytes"],
                    "note": "fp32 NCTHW clips from pinned host memory each step (as train.py:136 .cuda()), "
                            "prefetched on a copy stream; " + ("loss.item() each step" if train else "logits .cpu() each step")},
            "e2e_u8": None if not r["e2e_u8"] else {
                "value": world * B / (r["e2e_u8"]["ms"] * 1e-3), "unit": "clips/s", "ms_per_step": r["e2e_u8"]["ms"],
                "h2d_bytes_per_step": r["e2e_u8"]["h2d_bytes"], "d2h_bytes_per_step": r["d2h_bytes"],
                "note": "SURVEY 8(f) row f1: the same step fed uint8 (N,T,S,S,21) frames from pinned host "
                        "memory, as the dataset holds them before ToTensor/Normalize (chalearn_dataset.py:41-46); "
                        "model(staging.RawClips) normalises, splits 0:5/5:20, subsamples the Slow pathway and writes the "
                        "stem operands in one kernel (sfb_stage_clips_u8), bit-identical to the fp32 path"},
            "gpu_launches": r["launches"], "clocks": r["clocks"],
            "step_tensor_roofline": {"algorithmic_tflop_per_clip": flops_clip / 1e3, "achieved_tflops_per_gpu": tf_gpu,
                                     "peak_sustained": peaks["sustained"], "frac_sustained": tf_gpu / peaks["sustained"],
                                     "peak_burst": peaks["burst"], "frac_burst": tf_gpu / peaks["burst"],
                                     "peaks": peaks["src"]}}
    if train:
        line["config"]["optimizer"] = "video_classification_b200.optim.Adam(lr=1e-4): sfb_adam_step, one launch per step"
    if r["replica_check"] is not None:
        line["replica_check"] = r["replica_check"]
    if r.get("ddp_timeline"):
        line["ddp_timeline"] = r["ddp_timeline"]
    eng_ = r["model"]._engine
    line["config"]["cuda_graphs"] = {"enabled": bool(eng_.use_graphs), "captured": sum(len(p_.graphs) for p_ in eng_.plans.values()),
                                     "errors": eng_.graph_errors[:2]}
    if kernels:
        # dominant kernel = conv_tma_kernel<BN,BRES>: sfb_conv_gemm forward and input-gradient launches are the same
        # CUDA kernel (csrc/conv_tma.cuh); the weight gradient is a different kernel and listed in the breakdown
        fam = [kernels[k] for k in ("sfb_conv_gemm[fprop]", "sfb_conv_gemm[dgrad]", "sfb_conv_dgrad_bn") if k in kernels]
        if fam:
            ms = sum(v["ms"] for v in fam)
            tfl = sum(v["algorithmic_tflop"] for v in fam)
            calls = sum(v["launches"] for v in fam)
            # kernel launches, not API calls: a strided dgrad is one conv_tma_kernel launch per parity class
            nk = sum(v.get("kernel_launches", v["launches"]) for v in fam)
            ach = tfl * 1e3 / ms
            line["roofline"] = {"kernel": "conv_tma_kernel (sfb_conv_gemm fprop + dgrad, sfb_conv_dgrad_bn)", "bound": "tensor",
                                "achieved": ach,
                                "peak": peaks["sustained"], "unit": "TFLOP/s", "frac": ach / peaks["sustained"],
                                "traffic": None, "launches_per_step": nk, "api_calls_per_step": calls, "ms_per_step": ms,
                                "algorithmic_tflop_per_step": tfl,
                                "peak_source": f"{peaks['src']} bf16_tflops_sustained (kernel timed inside a long step)",
                                "share_of_step": ms / sum(v["ms"] for v in kernels.values())}
            line["roofline"]["algorithmic_bytes_per_launch"] = sum(v.get("algorithmic_bytes", 0.0) for v in fam) / nk
            if train and args.workload == "hyb":
                line["roofline"].update(conv_traffic_from_ncu())
        line["kernel_breakdown_ms_per_step"] = {k: round(v["ms"], 3) for k, v in sorted(kernels.items(), key=lambda t: -t[1]["ms"])}
        line["kernel_tflops"] = {k: round(v["tflops"], 1) for k, v in kernels.items() if "tflops" in v}
    if adam is not None:
        line["optimizer_step"] = adam
    lib = library_baseline_from_profiles()
    if lib is not None and args.workload == "hyb" and train:
        line["library_baseline"] = lib
    if world == 1 and not args.no_cpu_baseline:
        if _ORIG_AFFINITY:
            os.sched_setaffinity(0, _ORIG_AFFINITY)  # the CPU baseline uses every host core, not the GPU's NUMA node only
        n = args.cpu_clips
        sec, threads = cpu_steps(args.mode, args.workload, n, 2, 1)
        what = "train steps" if train else "eval forwards"
        line["cpu_baseline"] = {"value": n / sec, "unit": "clips/s", "cores": threads, "kind": "port",
                                "cpu": cpu_model_string(),
                                "sample": f"2 {what} (after 1 warm-up) of {n} clips of this workload, fp32 eager PyTorch on "
                                          "the host cores; oracle restatement of the reference path"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
