"""Execution engine of the B200 SlowFast: compiles the parameter tree of my_slowfast.SlowFastB200 (or of
res3d.SlowOnlyB200) into a static list of libsfb200 kernel calls (forward, and backward for training) per input shape,
with every activation / gradient buffer planned up front, and replays the lists as CUDA graphs.

Data flow (training), per conv layer L with BatchNorm:
    conv_gemm        : raw y_L (bf16 NDHWC) + per-channel sum / sum-of-squares fused in the GEMM epilogue
    bn_add_relu_raw  : finalises the batch statistics (mean/var -> scale, shift, running-stat update) and materialises
                       z_L = relu(y_L*scale+shift), the next conv's operand -- one HBM-bound pass (training BatchNorm needs
                       the statistics of the WHOLE tensor before anything can be normalised)
    block tail       : out = relu(bn_c(y_c) + shortcut) + packed ReLU mask          (one fused pass per ResBlock)
Inference: BatchNorm folded into per-channel (scale, shift); affine + residual + ReLU run in the GEMM epilogue, no
elementwise pass at all.
The Fast->Slow lateral conv's activated output lands directly behind the Slow stage output in one channel-concatenated
buffer (torch.cat of reference model/my_slowfast.py:343 never happens, in either direction).
Backward: the input-gradient GEMM that produces dz of a conv -> BN -> ReLU layer masks it and reduces that layer's
dgamma / dbeta in its epilogue (sfb_conv_dgrad_bn); bn_bwd_apply turns the masked gradient into dy; conv_wgrad accumulates
fp32 weight gradients into one flat arena (reverse-forward order, ready for bucketed all-reduce); block tails use the
packed mask (bn_bwd_reduce / apply, or sfb_bn_bwd_tail2 for both BatchNorms of a projection-shortcut tail).
"""
import ctypes as C
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed
import torch.nn as nn

from . import _ffi
from ._ffi import ConvGeom, Epilogue, Prologue, lib
from . import ops as _ops
from .ops import Act, conv_out_dims, round_up


def _probe():
    """no-op plan entry: lets a debug capture snapshot a buffer at this point of the op list"""
    return 0


def _f32(n, dev, zero=False):
    return (torch.zeros if zero else torch.empty)(n, dtype=torch.float32, device=dev)


class BufferPool:
    """Plan-time recycling of bf16 scratch buffers (ops run in stream order, so static reuse is safe)."""

    def __init__(self, dev):
        self.dev = dev
        self.free_list: Dict[int, List[torch.Tensor]] = {}
        self.total_bytes = 0

    def alloc(self, shape) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= s
        lst = self.free_list.get(n)
        if lst:
            return lst.pop().view(shape)
        self.total_bytes += n * 2
        return torch.empty(shape, dtype=torch.bfloat16, device=self.dev)

    def release(self, t: torch.Tensor):
        self.free_list.setdefault(t.numel(), []).append(t)


class ConvL:
    """One Conv3d (+ its BatchNorm3d) of the graph: parameters, packed bf16 operands, BN scratch, gradient views."""

    def __init__(self, name, conv: nn.Conv3d, bn: nn.BatchNorm3d, needs_dgrad=True, conv_name=None):
        self.name = name
        self.conv_name = conv_name or name  # path of the nn.Conv3d in the module tree (debug captures)
        self.conv, self.bn = conv, bn
        self.k = tuple(conv.kernel_size)
        self.s = tuple(conv.stride)
        self.p = tuple(conv.padding)
        self.Ci, self.Co = conv.in_channels, conv.out_channels
        self.Cg = round_up(self.Ci, 16)    # gathered-operand width: TMA boxes / MMA K steps are 16 channels
        self.Cgt = round_up(self.Co, 16)   # same for the dgrad operand (dY)
        self.taps = self.k[0] * self.k[1] * self.k[2]
        self.needs_dgrad = needs_dgrad
        self.wp = self.wpt = None
        self.wps = None  # stems: space-to-depth packing for the row-window kernels (csrc/stem_win.cuh)
        # pixel folding (narrow-channel layers, csrc/elementwise.cuh): f W-neighbouring pixels viewed as one
        self.fold_f = 1
        self.ldi, self.ldo = round_up(self.Ci, 16), round_up(self.Co, 16)
        self.kf = (self.k[0], self.k[1], 1 if self.k[2] == 1 else 3)
        self.pf = tuple(v // 2 for v in self.kf)
        self.wf = self.wpf = self.wptf = self.dwf = self.sf = None
        self.is_stem = False
        self.packed_version = -1
        # set by the engine
        self.stats = self.scale = self.shift = self.mean = self.invstd = None
        self.dw = self.dgamma = self.dbeta = None


class FinalAct:
    """A materialised (post-activation) tensor plus, in training plans, its gradient buffer."""

    def __init__(self, act: Act):
        self.act = act
        self.grad: Optional[Act] = None  # allocated at the first gradient contribution (backward plan)
        self.ginit = False  # has any contribution been written to grad yet (plan-time flag)
        self.cat = None     # CatInfo when this tensor lives inside a Slow|lateral concat buffer


class CatInfo:
    """Channel-concatenated Slow-stage input: [slow_out (final) | lateral conv output (raw, BN+ReLU lazily applied)]."""

    def __init__(self, cat: Act, c_slow: int):
        self.cat, self.c_slow = cat, c_slow
        self.dcat: Optional[Act] = None
        self.X: Optional[FinalAct] = None         # the whole buffer as the next stage's input
        self.slow_out: Optional[FinalAct] = None  # channels [0, c_slow)


class Plan:
    def __init__(self):
        self.fwd: List[Tuple] = []
        self.bwd: List[Tuple] = []
        self.keep = []  # ctypes structs / tensors that must outlive the op lists
        self.x_in: List[Act] = []
        self.mask_bits = []  # (block output Act, packed ReLU mask) pairs of a training plan
        self.stats_folded = {}  # conv name -> its fused BatchNorm sums are in the pixel-folded layout
        self.logits = None
        self.dlogits = None
        self.act_bytes = 0
        self.seed_dev = None  # int64[1] on the device: the head dropout's seed of the current step
        self.bn_flat = None   # this plan's BatchNorm scratch (SlowFastEngine._bind_bn_arena)
        self.stats_flat = None
        self.gen = 0  # bumped by every training forward; backward checks it saw the latest one
        # CUDA graphs of the two op lists (SlowFastEngine._graphed): {"fwd"|"bwd": (CUDAGraph, kernels per replay)},
        # eager runs so far, and whether a capture failed (then the list keeps running eagerly)
        # eval plans: the BatchNorm folds (running statistics -> scale / shift, ~110 tiny launches) are kept apart from the
        # forward list and re-run only when a BatchNorm parameter or buffer may have changed (SlowFastEngine._fold_key)
        self.folds: List[Tuple] = []
        self.fold_key = None
        self.graphs = {}
        self.graph_runs = {}
        self.bwd_segments = None  # data parallel: the backward list cut at the bucket launches
        self.graph_failed = {}


class SlowFastEngine:
    def __init__(self, model, dry_run: bool = False):
        """dry_run=True plans on the `meta` device without touching the GPU (host-logic tests only; it cannot
        execute anything)."""
        self.model = model
        cfg = model.graph_cfg
        self.cfg = cfg
        self.dry_run = dry_run
        p0 = next(model.parameters())
        if not p0.is_cuda and not dry_run:
            raise _ffi.SfbError("the B200 SlowFast runs on CUDA only: call model.cuda() first (no CPU fallback)")
        self.dev = torch.device("meta") if dry_run else p0.device
        self.lib = lib()
        self.stream = C.c_void_p(0)
        self.impl = _ffi.IMPL_TC
        self._pack_cache = {}
        self.plans: Dict[Tuple, Plan] = {}
        self.head_mod = model.blocks[6] if int(cfg.get("npath", 2)) == 2 else model.blocks[5]
        self.dropout_p = self.head_mod.dropout.p if self.head_mod.dropout is not None else 0.0
        self.step_seed = 0
        self.ddp = None  # set by parallel.attach_ddp
        # BatchNorm backward reduction fused into the producing dgrad's epilogue (SFB_NO_BN_FUSE=1: separate passes)
        import os as _os
        self.fuse_bn_reduce = not _os.environ.get("SFB_NO_BN_FUSE")
        self.graph_failed_all = False
        self.graph_errors = []
        # static op lists replayed as CUDA graphs from the third run of a plan on (SFB_GRAPHS=0: always eager launches)
        self.use_graphs = _os.environ.get("SFB_GRAPHS", "1") != "0"
        self.graphs_with_ddp = _os.environ.get("SFB_GRAPHS_DDP", "0") == "1"
        self.graph_after = 2
        self.prof = None     # list: bench.py sets it to time every plan entry with CUDA events
        self.capture = None  # dict: tools/layer_diag.py sets it to snapshot tagged intermediates (debug only)
        self._collect_layers()
        self._alloc_arenas()

    # ------------------------------------------------------------------ graph description
    def _collect_layers(self):
        m = self.model
        cfg = self.cfg
        self.fuse = cfg["fuse"]
        L: List[ConvL] = []
        # number of pathways: 2 = the SlowFast tree (blocks.N.multipathway_blocks.P...), 1 = the Slow-only 3D-ResNet tree
        # of pytorchvideo's create_resnet (blocks.0 stem, blocks.1-4 stages, blocks.5 head; res3d.py)
        P = self.P = int(cfg.get("npath", 2))
        stem_path = (lambda p: "blocks.0") if P == 1 else (lambda p: f"blocks.0.multipathway_blocks.{p}")
        stage_path = (lambda si, p: f"blocks.{si}") if P == 1 else (lambda si, p: f"blocks.{si}.multipathway_blocks.{p}")
        self.stems = []
        for p in range(P):
            st = m.get_submodule(stem_path(p))
            self.stems.append(ConvL(stem_path(p), st.conv, st.norm, needs_dgrad=False, conv_name=stem_path(p) + ".conv"))
        for st_ in self.stems:
            st_.is_stem = True
        L += self.stems
        self.laterals: List[Optional[ConvL]] = []
        for si in range(5):
            fu = m.blocks[si].multipathway_fusion if P == 2 else None
            if self.fuse and fu is not None and not isinstance(fu, nn.Identity):
                self.laterals.append(ConvL(f"blocks.{si}.multipathway_fusion", fu.conv_fast_to_slow[0], fu.norm[0],
                                           conv_name=f"blocks.{si}.multipathway_fusion.conv_fast_to_slow.0"))
            else:
                self.laterals.append(None)
        self.stages = []  # [stage][pathway] -> list of block dicts
        for si in range(1, 5):
            per_path = []
            for p in range(P):
                blks = []
                for bi, rb in enumerate(m.get_submodule(stage_path(si, p)).res_blocks):
                    nm = f"{stage_path(si, p)}.res_blocks.{bi}"
                    b2 = rb.branch2
                    d = dict(a=ConvL(nm + ".a", b2.conv_a, b2.norm_a, conv_name=nm + ".branch2.conv_a"),
                             b=ConvL(nm + ".b", b2.conv_b, b2.norm_b, conv_name=nm + ".branch2.conv_b"),
                             c=ConvL(nm + ".c", b2.conv_c, b2.norm_c, conv_name=nm + ".branch2.conv_c"),
                             s=ConvL(nm + ".s", rb.branch1_conv, rb.branch1_norm, conv_name=nm + ".branch1_conv")
                             if rb.branch1_conv is not None else None, name=nm)
                    blks.append(d)
                per_path.append(blks)
            self.stages.append(per_path)
        # forward order of all conv layers (used for arenas and the reverse-order gradient layout)
        order: List[ConvL] = list(self.stems)
        if self.laterals[0] is not None:
            order.append(self.laterals[0])
        for si in range(4):
            for p in reversed(range(P)):  # execution order: Fast first (its output feeds the lateral of the Slow concat buffer)
                for d in self.stages[si][p]:
                    if d["s"] is not None:
                        order.append(d["s"])
                    order += [d["a"], d["b"], d["c"]]
            if self.laterals[si + 1] is not None:
                order.append(self.laterals[si + 1])
        self.layers = order
        import os as _os
        if not _os.environ.get("SFB_NO_FOLD"):
            for l in order:
                if not l.is_stem:
                    l.fold_f = _ops.fold_factor(l.ldi, l.ldo, l.k, l.s, 8)
        self.proj = self.head_mod.proj
        self.bn_modules = [l.bn for l in self.layers]
        self.bn_epoch = 0
        self._bn_tensors = [t for bn in self.bn_modules for t in (bn.running_mean, bn.running_var)]
        self._ptr_sig = None
        self._objs_cache = None

    def _bind_bn_arena(self, plan: "Plan"):
        """BatchNorm scratch of ONE plan: [all raw sums | per layer scale, shift, mean, invstd].  Every plan owns its
        arena (≈ 0.2 MB), so a forward of another shape, or an eval forward (whose sfb_bn_fold rewrites scale / shift
        from the running statistics), between a training forward and its backward cannot touch the coefficients that
        backward reads -- the reference's autograd graph allows any such interleaving.  The layer attributes
        l.stats / l.scale / ... are (re)bound to the plan being built or inspected."""
        if plan.bn_flat is None:
            n_stats = sum(2 * l.Co + (2 * l.fold_f * l.ldo if l.fold_f > 1 else 0) for l in self.layers)
            n_rest = sum(4 * l.Co for l in self.layers)
            plan.bn_flat = _f32(n_stats + n_rest, self.dev, zero=True)
            plan.stats_flat = plan.bn_flat[:n_stats]
        flat = plan.bn_flat
        o_s, o_r = 0, plan.stats_flat.numel()
        for l in self.layers:
            c = l.Co
            l.stats = flat[o_s:o_s + 2 * c]
            o_s += 2 * c
            if l.fold_f > 1:
                l.sf = flat[o_s:o_s + 2 * l.fold_f * l.ldo]
                o_s += 2 * l.fold_f * l.ldo
            l.scale, l.shift = flat[o_r:o_r + c], flat[o_r + c:o_r + 2 * c]
            l.mean, l.invstd = flat[o_r + 2 * c:o_r + 3 * c], flat[o_r + 3 * c:o_r + 4 * c]
            o_r += 4 * c

    def _alloc_arenas(self):
        dev = self.dev
        # ---- flat gradient arena, reverse forward order (head first) ----
        live: List[torch.nn.Parameter] = []
        for l in self.layers:
            live += [l.conv.weight, l.bn.weight, l.bn.bias]
        live += [self.proj.weight, self.proj.bias]
        self.live_params = live
        offs, o = {}, 0
        for prm in reversed(live):
            offs[id(prm)] = o
            o += round_up(prm.numel(), 8)
        # folded weight gradients (scratch behind the parameters: zeroed with the arena, outside every DDP bucket)
        fold_offs = {}
        for l in self.layers:
            if l.fold_f > 1:
                n = l.fold_f * l.ldo * l.fold_f * l.ldi * l.kf[0] * l.kf[1] * l.kf[2]
                fold_offs[id(l)] = (o, n)
                o += round_up(n, 8)
        self.grad_flat = _f32(o, dev, zero=True)
        for l in self.layers:
            if l.fold_f > 1:
                st, n = fold_offs[id(l)]
                l.dwf = self.grad_flat[st:st + n].view(l.fold_f * l.ldo, l.fold_f * l.ldi, *l.kf)
        self.grad_views = {}
        for prm in live:
            st = offs[id(prm)]
            self.grad_views[id(prm)] = self.grad_flat[st:st + prm.numel()].view(prm.shape)
        self.grad_offsets = offs
        self._grad_view_specs = [(tuple(prm.shape), tuple(self.grad_views[id(prm)].stride()), offs[id(prm)]) for prm in live]
        for l in self.layers:
            l.dw = self.grad_views[id(l.conv.weight)]
            l.dgamma = self.grad_views[id(l.bn.weight)]
            l.dbeta = self.grad_views[id(l.bn.bias)]

    def invalidate_weights(self):
        """public hook: call after writing parameters or BatchNorm buffers through anything that does not move their
        version counters (raw pointers, p.data = ...); load_state_dict and the model's own training steps call it / are
        tracked already"""
        for l in self.layers:
            l.packed_version = -1
        self._pack_cache = {}
        self.bn_epoch += 1

    def _current_objects(self):
        """parameters and BatchNorm buffers as the module tree holds them NOW (dict lookups: nn.Module.__getattr__ costs
        a microsecond per access, this runs on every forward)"""
        par, buf = [], []
        for l in self.layers:
            bp = l.bn._parameters
            par += [l.conv._parameters["weight"], bp["weight"], bp["bias"]]
            bb = l.bn._buffers
            buf += [bb["running_mean"], bb["running_var"]]
        pp = self.proj._parameters
        par += [pp["weight"], pp["bias"]]
        return par, buf

    def params_replaced(self) -> bool:
        """True when a Parameter OBJECT of the module tree is no longer the one this engine was built around
        (`load_state_dict(assign=True)`, `module.weight = nn.Parameter(...)`): gradients, the arena views and every plan
        refer to the old objects, so the model builds a new engine."""
        self._objs_cache = self._current_objects()  # reused by the address check of the same forward
        return any(a is not b for a, b in zip(self._objs_cache[0], self.live_params))

    def _check_baked_pointers(self):
        """Plans (and the CUDA graphs captured from them) hold the device addresses of every parameter and BatchNorm
        buffer.  Rebinding storage -- `p.data = ...`, a buffer attribute assigned anew, a swapped-in EMA copy -- would leave
        the kernels reading the old memory without any error, so the addresses are compared on every forward (~550
        `data_ptr()` calls, ~0.1 ms of host time hidden behind the GPU) and everything address-dependent is rebuilt when
        one moved."""
        par, buf = self._objs_cache or self._current_objects()
        self._objs_cache = None
        self._bn_tensors = buf
        sig = tuple([t.data_ptr() for t in par] + [t.data_ptr() for t in buf])
        if sig != self._ptr_sig:
            if self._ptr_sig is not None:
                self.plans.clear()
                self.invalidate_weights()
            self._ptr_sig = sig

    def _fold_key(self):
        """changes whenever a folded BatchNorm coefficient may be stale: any training forward of this engine (running
        statistics), invalidate_weights() / moved storage, or an in-place update of gamma / beta / the buffers that
        autograd can see"""
        v = 0
        for l in self.layers:
            bp = l.bn._parameters
            v += bp["weight"]._version + bp["bias"]._version
        for t in self._bn_tensors:
            v += t._version
        return (self.bn_epoch, v)

    # ------------------------------------------------------------------ weights
    def _pack_tables(self, training: bool):
        """Device-resident job tables that re-derive EVERY operand in two launches (level 0: straight from the
        master weights, incl. the folded fp32 weights; level 1: operands packed from those).  Built once per
        (mode, kernel impl) -- buffers and master-weight storage are stable; load_state_dict / optimizers write in
        place -- and dropped by invalidate_weights()."""
        key = (bool(training), self.impl)
        tabs = self._pack_cache.get(key)

        def ptr_sig():  # master weights AND every operand buffer the table writes (they are reallocated when impl changes)
            out = []
            for l in self.layers:
                out.append(l.conv.weight.data_ptr())
                for b in (l.wp, l.wpt, l.wps, l.wf, l.wpf, l.wptf):
                    out.append(b.data_ptr() if b is not None else 0)
            return out
        if tabs is not None and tabs["ptrs"] == ptr_sig():
            return tabs
        P = _ffi.PackJob
        lv0, lv1 = [], []
        for l in self.layers:
            w = l.conv.weight
            if not w.is_contiguous():
                return None
            src = w.data_ptr()
            kT, kH, kW = l.k
            rp, kp = l.wp.shape
            lv0.append(P(_ffi.PACK_WEIGHT, l.Co, l.Ci, kT, kH, kW, 1, 1, 1, l.Cg, 0, 0, 0, 0, rp, kp, src, l.wp.data_ptr()))
            dg = training and l.needs_dgrad
            if dg:
                rpt, kpt = l.wpt.shape
                if self.impl == _ffi.IMPL_TC:
                    lv0.append(P(_ffi.PACK_WEIGHT_DGRAD, l.Co, l.Ci, kT, kH, kW, l.s[0], l.s[1], l.s[2], l.Cgt, 0, 0, 0, 0,
                                 rpt, kpt, src, l.wpt.data_ptr()))
                else:
                    lv0.append(P(_ffi.PACK_WEIGHT, l.Co, l.Ci, kT, kH, kW, 1, 1, 1, l.Cgt, 1, 0, 0, 0, rpt, kpt, src,
                                 l.wpt.data_ptr()))
            if l.wps is not None:
                lv0.append(P(_ffi.PACK_STEM, l.Co, l.Ci, 1, 7, 7, 1, 2, 2, l.wps.shape[1] // 16, 0, 0, 0, 0, l.wps.shape[0],
                             l.wps.shape[1], src, l.wps.data_ptr()))
            if l.fold_f > 1:
                f = l.fold_f
                lv0.append(P(_ffi.PACK_FOLD, l.Co, l.Ci, kT, kH, kW, 1, 1, 1, 0, 0, f, l.ldo, l.ldi, 0, 0, src, l.wf.data_ptr()))
                fk = l.kf
                lv1.append(P(_ffi.PACK_WEIGHT, f * l.ldo, f * l.ldi, fk[0], fk[1], fk[2], 1, 1, 1, f * l.ldi, 0, 0, 0, 0,
                             l.wpf.shape[0], l.wpf.shape[1], l.wf.data_ptr(), l.wpf.data_ptr()))
                if dg:
                    if self.impl == _ffi.IMPL_TC:
                        lv1.append(P(_ffi.PACK_WEIGHT_DGRAD, f * l.ldo, f * l.ldi, fk[0], fk[1], fk[2], 1, 1, 1, f * l.ldo, 0,
                                     0, 0, 0, l.wptf.shape[0], l.wptf.shape[1], l.wf.data_ptr(), l.wptf.data_ptr()))
                    else:
                        lv1.append(P(_ffi.PACK_WEIGHT, f * l.ldo, f * l.ldi, fk[0], fk[1], fk[2], 1, 1, 1, f * l.ldo, 1,
                                     0, 0, 0, l.wptf.shape[0], l.wptf.shape[1], l.wf.data_ptr(), l.wptf.data_ptr()))
        levels = []
        for jobs in (lv0, lv1):
            if not jobs:
                continue
            arr = (P * len(jobs))(*jobs)
            nbytes = int(self.lib.sfb_pack_table_bytes(len(jobs)))
            tab = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
            nblk = C.c_int32(0)
            _ffi.check(self.lib.sfb_pack_plan(arr, len(jobs), tab.data_ptr(), nbytes, C.byref(nblk)), "sfb_pack_plan")
            levels.append((tab, len(jobs), nblk.value))
        tabs = {"levels": levels, "ptrs": ptr_sig()}
        self._pack_cache[key] = tabs
        return tabs

    def refresh_weights(self, training: bool):
        """(Re)pack bf16 GEMM operands of every conv whose fp32 master weight changed (optimizer step / load).
        Version counters catch ordinary in-place updates; optimizers that write through raw pointers or fused
        multi-tensor kernels do not always move them (torch.optim.Adam(fused=True) left them unchanged in this build:
        the ncu launch lists r1z/r2e show no re-pack), so the first forward after ANY backward re-packs everything --
        one redundant batched launch pair if no optimizer ran in between, never a stale operand."""
        if getattr(self, "_bwd_since_pack", False):
            self._bwd_since_pack = False
            for l in self.layers:
                l.packed_version = -1
        stale = [l for l in self.layers
                 if not (l.packed_version == l.conv.weight._version and l.wp is not None and
                         getattr(l, "packed_impl", None) == self.impl and
                         (l.wpt is not None or not (training and l.needs_dgrad)))]
        if not stale:
            return
        # every operand buffer exists (a first pass went through the per-layer path) and all weights changed --
        # the per-step case after an optimizer update: two batched launches instead of ~330 small ones
        if (len(stale) == len(self.layers) and not self.dry_run and all(l.wp is not None for l in self.layers) and
                all(l.wpt is not None for l in self.layers if training and l.needs_dgrad) and
                all(getattr(l, "packed_impl", None) == self.impl for l in self.layers)):
            tabs = self._pack_tables(training)
            if tabs is not None:
                for tab, n, nblk in tabs["levels"]:
                    _ffi.check(self.lib.sfb_pack_run(tab.data_ptr(), n, nblk, self.stream), "sfb_pack_run")
                for l in self.layers:
                    l.packed_version = l.conv.weight._version
                return
        for l in self.layers:
            w = l.conv.weight
            if l.packed_version == w._version and l.wp is not None and getattr(l, "packed_impl", None) == self.impl and \
                    (l.wpt is not None or not (training and l.needs_dgrad)):
                continue
            l.packed_impl = self.impl
            rp, kp = _ffi.packed_weight_dims(l.Co, l.Cg, l.taps)
            if l.wp is None:
                l.wp = torch.empty((rp, kp), dtype=torch.bfloat16, device=self.dev)
            s2d_ok = l.is_stem and l.k == (1, 7, 7) and l.s == (1, 2, 2) and l.p == (0, 3, 3) and \
                _ops.stem_s2d_eligible(l.Ci, l.Co, l.k, l.s, l.p, 2, 2)
            if s2d_ok and l.wps is None:
                cs = 32 if 4 * l.Ci <= 32 else 64
                l.wps = torch.empty((_ops.stem_rows_pad(l.Co), 16 * cs), dtype=torch.bfloat16, device=self.dev)
            if l.fold_f > 1 and l.wf is None:
                f_, tf_ = l.fold_f, l.kf[0] * l.kf[1] * l.kf[2]
                l.wf = torch.empty((f_ * l.ldo, f_ * l.ldi) + l.kf, dtype=torch.float32, device=self.dev)
                rpf, kpf = _ffi.packed_weight_dims(f_ * l.ldo, f_ * l.ldi, tf_)
                l.wpf = torch.empty((rpf, kpf), dtype=torch.bfloat16, device=self.dev)
                rpt_, kpt_ = _ffi.packed_weight_dims_dgrad(f_ * l.ldi, f_ * l.ldo, l.kf, (1, 1, 1))
                l.wptf = torch.empty((rpt_, kpt_), dtype=torch.bfloat16, device=self.dev)
            if self.dry_run:
                rpt, kpt = _ffi.packed_weight_dims_dgrad(l.Ci, l.Cgt, l.k, l.s)
                l.wpt = torch.empty((rpt, kpt), dtype=torch.bfloat16, device=self.dev)
                l.packed_version = w._version
                continue
            wd = w.detach()
            if not wd.is_contiguous():
                wd = wd.contiguous()
            _ffi.check(self.lib.sfb_pack_weight(wd.data_ptr(), l.Co, l.Ci, l.taps, l.Cg, 0, l.wp.data_ptr(), rp, kp, self.stream),
                       "sfb_pack_weight")
            if l.fold_f > 1:
                f = l.fold_f
                tf = l.kf[0] * l.kf[1] * l.kf[2]
                _ffi.check(self.lib.sfb_fold_weight(wd.data_ptr(), l.Co, l.Ci, l.k[0], l.k[1], l.k[2], f, l.ldo, l.ldi,
                                                    l.wf.data_ptr(), self.stream), "sfb_fold_weight")
                _ffi.check(self.lib.sfb_pack_weight(l.wf.data_ptr(), f * l.ldo, f * l.ldi, tf, f * l.ldi, 0, l.wpf.data_ptr(),
                                                    l.wpf.shape[0], l.wpf.shape[1], self.stream), "sfb_pack_weight(folded)")
                if training and l.needs_dgrad:
                    if self.impl == _ffi.IMPL_TC:
                        rpt, kpt = _ffi.packed_weight_dims_dgrad(f * l.ldi, f * l.ldo, l.kf, (1, 1, 1))
                    else:
                        rpt, kpt = _ffi.packed_weight_dims(f * l.ldi, f * l.ldo, tf)
                    if l.wptf is None or tuple(l.wptf.shape) != (rpt, kpt):
                        l.wptf = torch.empty((rpt, kpt), dtype=torch.bfloat16, device=self.dev)
                    if self.impl == _ffi.IMPL_TC:
                        _ffi.check(self.lib.sfb_pack_weight_dgrad(l.wf.data_ptr(), f * l.ldo, f * l.ldi, l.kf[0], l.kf[1], l.kf[2],
                                                                  1, 1, 1, f * l.ldo, l.wptf.data_ptr(), rpt, kpt, self.stream),
                                   "sfb_pack_weight_dgrad(folded)")
                    else:
                        _ffi.check(self.lib.sfb_pack_weight(l.wf.data_ptr(), f * l.ldo, f * l.ldi, tf, f * l.ldo, 1,
                                                            l.wptf.data_ptr(), rpt, kpt, self.stream), "sfb_pack_weight(T, folded)")
            if l.wps is not None:
                _ffi.check(self.lib.sfb_pack_weight_stem(wd.data_ptr(), l.Co, l.Ci, l.wps.shape[1] // 16, l.wps.data_ptr(),
                                                         l.wps.shape[0], self.stream), "sfb_pack_weight_stem")
            if training and l.needs_dgrad:
                if self.impl == _ffi.IMPL_TC:
                    # class-major operand: strided convs back-propagate as one stride-1 launch per parity class
                    rpt, kpt = _ffi.packed_weight_dims_dgrad(l.Ci, l.Cgt, l.k, l.s)
                    if l.wpt is None or tuple(l.wpt.shape) != (rpt, kpt):
                        l.wpt = torch.empty((rpt, kpt), dtype=torch.bfloat16, device=self.dev)
                    _ffi.check(self.lib.sfb_pack_weight_dgrad(wd.data_ptr(), l.Co, l.Ci, l.k[0], l.k[1], l.k[2], l.s[0], l.s[1],
                                                              l.s[2], l.Cgt, l.wpt.data_ptr(), rpt, kpt, self.stream),
                               "sfb_pack_weight_dgrad")
                else:
                    rpt, kpt = _ffi.packed_weight_dims(l.Cg, l.Cgt, l.taps)
                    if l.wpt is None or tuple(l.wpt.shape) != (rpt, kpt):
                        l.wpt = torch.empty((rpt, kpt), dtype=torch.bfloat16, device=self.dev)
                    _ffi.check(self.lib.sfb_pack_weight(wd.data_ptr(), l.Co, l.Ci, l.taps, l.Cgt, 1, l.wpt.data_ptr(), rpt, kpt,
                                                        self.stream), "sfb_pack_weight(T)")
            l.packed_version = w._version

    # ------------------------------------------------------------------ op emission helpers
    def _emit(self, lst, fn, *args, tag=None):
        """tag = (name, Act | tensor): what this op just finished writing (debug captures only)"""
        lst.append((fn, args, tag))

    def _conv(self, plan, lst, L: ConvL, x: Act, pro, out: Act, *, stats=False, transposed=False, residual: Act = None,
              wp=None, tag=None, epi=None, bnred=None):
        """x is gathered with its channel count rounded up to 16 (the pad lanes of every buffer are zero); a dgrad
        (transposed) writes the full padded width of its output so those lanes stay zero in recycled buffers.
        epi = (scale, shift, relu): folded-BatchNorm epilogue (eval mode).
        bnred = (Lb, y): this dgrad produces the gradient reaching layer Lb's ReLU(BatchNorm(y)); the epilogue masks it
        with the forward ReLU and accumulates Lb's dgamma / dbeta (sfb_conv_dgrad_bn) -- no bn_bwd_reduce pass."""
        k, s, p = L.k, L.s, L.p
        N, Ti, Hi, Wi = x.dims
        _, To, Ho, Wo = out.dims
        f = self._fold_of(L, x, out, transposed) if (wp is None and pro is None) else 1
        if f > 1 and epi is not None and residual is not None and not (
                residual.coff == 0 and residual.ld == out.ld and residual.dims == out.dims):
            f = 1  # the residual must fold the same way as the output
        if f > 1:
            # the same buffers seen as (N,T,H,W/f,f*ld): 128-byte rows for TMA, block-banded weights
            k, p = L.kf, L.pf
            cg, co = f * x.ld, f * out.ld
            g = ConvGeom(N, Ti, Hi, Wi // f, cg, To, Ho, Wo // f, co, k[0], k[1], k[2], 1, 1, 1, p[0], p[1], p[2],
                         cg, 0, co, 0, int(transposed))
            if epi is not None:
                # inference: folded-BatchNorm coefficients per FOLDED column (j*ld + c -> channel c; pad lanes 0, so
                # the pad lanes of the output stay 0 through the ReLU), re-derived right after the layer's sfb_bn_fold
                scf = _f32(co, self.dev, zero=True)
                shf = _f32(co, self.dev, zero=True)
                plan.keep += [scf, shf]

                def _replicate(scf=scf, shf=shf, sc=epi[0], sh=epi[1], n=L.Co, f=f, ld=out.ld):
                    scf.view(f, ld)[:, :n].copy_(sc)
                    shf.view(f, ld)[:, :n].copy_(sh)
                    return 0
                plan.folds.append((_replicate, (), None))
                ep = Epilogue(scf.data_ptr(), shf.data_ptr(), residual.ptr() if residual else 0,
                              f * residual.ld if residual else 0, 0, int(epi[2]), 0)
            else:
                assert residual is None or (residual.buf is out.buf and residual.coff == 0)
                ep = Epilogue(0, 0, out.ptr() if residual else 0, co if residual else 0, 0, 0, L.sf.data_ptr() if stats else 0)
            wp = L.wptf if transposed else L.wpf
        else:
            cg = round_up(x.C, 16)
            co = round_up(out.C, 16) if transposed else out.C
            assert x.coff + cg <= x.ld and out.coff + co <= out.ld
            g = ConvGeom(N, Ti, Hi, Wi, cg, To, Ho, Wo, co, k[0], k[1], k[2], s[0], s[1], s[2], p[0], p[1], p[2],
                         x.ld, x.coff, out.ld, out.coff, int(transposed))
            ep = Epilogue(epi[0].data_ptr() if epi else 0, epi[1].data_ptr() if epi else 0,
                          residual.ptr() if residual else 0, residual.ld if residual else 0,
                          residual.coff if residual else 0, int(epi[2]) if epi else 0, L.stats.data_ptr() if stats else 0)
            wp = wp if wp is not None else (L.wpt if transposed else L.wp)
        pr = Prologue(pro[0].data_ptr(), pro[1].data_ptr(), int(pro[2])) if pro is not None else Prologue(0, 0, 0)
        g.real_k = (L.Co if transposed else L.Ci) * L.taps  # algorithmic reduction length (no channel padding)
        g.real_n = L.Ci if transposed else L.Co
        g.real_m = N * Ti * Hi * Wi if transposed else N * To * Ho * Wo  # rows of the forward conv's output
        plan.keep += [g, pr, ep]
        if bnred is not None:
            Lb, y = bnred
            assert transposed and residual is None and y.dims == out.dims and y.ld == out.ld and y.coff == out.coff
            fz = _ffi.BnBwdFuse(y.ptr(), f * y.ld if f > 1 else y.ld, y.coff, Lb.scale.data_ptr(), Lb.shift.data_ptr(),
                                Lb.mean.data_ptr(), Lb.invstd.data_ptr(), Lb.dgamma.data_ptr(), Lb.dbeta.data_ptr(),
                                Lb.Co, y.ld if f > 1 else 0)
            plan.keep.append(fz)
            self._emit(lst, self.lib.sfb_conv_dgrad_bn, C.byref(g), x.ptr(), wp.data_ptr(), wp.shape[0], wp.shape[1],
                       out.ptr(), C.byref(fz), self.stream, tag=tag)
            return
        self._emit(lst, self.lib.sfb_conv_gemm, C.byref(g), x.ptr(), C.byref(pr), wp.data_ptr(), wp.shape[0], wp.shape[1],
                   out.ptr(), C.byref(ep), self.impl, self.stream,
                   tag=tag if tag is not None else (None if transposed else (L.conv_name + ".y", out)))
        if stats:
            plan.stats_folded[L.conv_name] = f > 1  # the consumer reads the folded sums directly (sfb_bn_add_relu_raw)

    def _fold_of(self, L: ConvL, x: Act, out: Act, transposed: bool) -> int:
        """fold factor to run this conv pass with (1 = unfolded): needs whole dense buffers on both sides"""
        f = L.fold_f
        if f <= 1:
            return 1
        a, b = (out, x) if transposed else (x, out)  # a = conv input side, b = conv output side
        if a.coff or b.coff or a.ld != L.ldi or b.ld != L.ldo or a.dims != b.dims or a.dims[3] % f:
            return 1
        return f

    def _wgrad(self, plan, lst, L: ConvL, x: Act, pro, dy: Act):
        k, s, p = L.k, L.s, L.p
        N, Ti, Hi, Wi = x.dims
        _, To, Ho, Wo = dy.dims
        f = self._fold_of(L, x, dy, False) if pro is None else 1
        if f > 1:
            k, p = L.kf, L.pf
            g = ConvGeom(N, Ti, Hi, Wi // f, f * x.ld, To, Ho, Wo // f, f * dy.ld, k[0], k[1], k[2], 1, 1, 1, p[0], p[1], p[2],
                         f * x.ld, 0, f * dy.ld, 0, 0)
            g.real_k, g.real_n, g.real_m = L.Ci * L.taps, L.Co, N * To * Ho * Wo
            pr = Prologue(0, 0, 0)
            plan.keep += [g, pr]
            if self.impl == _ffi.IMPL_TC:
                # the band map back to the reference layout is applied in the kernel's epilogue
                self._emit(lst, self.lib.sfb_conv_wgrad_folded, C.byref(g), x.ptr(), dy.ptr(), f * dy.ld, 0, L.dw.data_ptr(),
                           L.Co, L.Ci, L.k[2], f, L.ldo, L.ldi, self.stream)
            else:
                self._emit(lst, self.lib.sfb_conv_wgrad, C.byref(g), x.ptr(), C.byref(pr), dy.ptr(), f * dy.ld, 0,
                           L.dwf.data_ptr(), f * x.ld, self.impl, self.stream)
                self._emit(lst, self.lib.sfb_unfold_wgrad, L.dwf.data_ptr(), L.Co, L.Ci, L.k[0], L.k[1], L.k[2], f, L.ldo,
                           L.ldi, L.dw.data_ptr(), self.stream)
            if self.ddp is not None:
                self.ddp.note_grad_ready(plan, lst, L.conv.weight)
            return
        g = ConvGeom(N, Ti, Hi, Wi, round_up(x.C, 16), To, Ho, Wo, dy.C, k[0], k[1], k[2], s[0], s[1], s[2], p[0], p[1], p[2],
                     x.ld, x.coff, dy.ld, dy.coff, 0)
        pr = Prologue(pro[0].data_ptr(), pro[1].data_ptr(), int(pro[2])) if pro is not None else Prologue(0, 0, 0)
        g.real_k, g.real_n = L.Ci * L.taps, L.Co
        plan.keep += [g, pr]
        self._emit(lst, self.lib.sfb_conv_wgrad, C.byref(g), x.ptr(), C.byref(pr), dy.ptr(), dy.ld, dy.coff,
                   L.dw.data_ptr(), L.Ci, self.impl, self.stream)
        if self.ddp is not None:
            self.ddp.note_grad_ready(plan, lst, L.conv.weight)

    def _bn_stats_ready(self, plan, lst, L: ConvL, count: int, training: bool):
        bn = L.bn
        if training:
            self._emit(lst, self.lib.sfb_bn_finalize, L.stats.data_ptr(), L.Co, count, bn.weight.data_ptr(), bn.bias.data_ptr(),
                       float(bn.eps), float(bn.momentum if bn.momentum is not None else 0.1), bn.running_mean.data_ptr(),
                       bn.running_var.data_ptr(), L.scale.data_ptr(), L.shift.data_ptr(), L.mean.data_ptr(),
                       L.invstd.data_ptr(), self.stream)
        else:
            self._emit(plan.folds, self.lib.sfb_bn_fold, bn.weight.data_ptr(), bn.bias.data_ptr(), bn.running_mean.data_ptr(),
                       bn.running_var.data_ptr(), float(bn.eps), L.Co, L.scale.data_ptr(), L.shift.data_ptr(), self.stream)

    def _bn_bwd(self, plan, lst, L: ConvL, dz: Act, y: Act, mask, dy: Act, g_out: Optional[Act] = None,
                g_acc=False, do_reduce=True, premasked=False):
        """mask: None (ReLU mask recomputed from y), an Act (sign of that tensor) or a uint8 tensor (packed bits
        written by the forward residual tail).  premasked: dz already is the masked gradient and dgamma / dbeta are
        complete (the producing dgrad ran with the fused reduction, _conv(bnred=...)): only the apply pass remains."""
        if premasked:
            do_reduce = False
        bn = L.bn
        if isinstance(mask, torch.Tensor):
            m = (0, 0, 0, mask.data_ptr())
        else:
            m = (mask.ptr(), mask.ld, mask.coff, 0) if mask is not None else (0, 0, 0, 0)
        if do_reduce:
            self._emit(lst, self.lib.sfb_bn_bwd_reduce, dz.ptr(), dz.ld, dz.coff, y.ptr(), y.ld, y.coff, *m,
                       L.scale.data_ptr(), L.shift.data_ptr(), L.mean.data_ptr(), L.invstd.data_ptr(), 1, y.M, y.C,
                       L.dgamma.data_ptr(), L.dbeta.data_ptr(), self.stream)
        go = (g_out.ptr(), g_out.ld, g_out.coff) if g_out is not None else (0, 0, 0)
        self._emit(lst, self.lib.sfb_bn_bwd_apply, dz.ptr(), dz.ld, dz.coff, y.ptr(), y.ld, y.coff, *m,
                   L.scale.data_ptr(), L.shift.data_ptr(), L.mean.data_ptr(), L.invstd.data_ptr(), bn.weight.data_ptr(),
                   L.dgamma.data_ptr(), L.dbeta.data_ptr(), 0 if premasked else 1, y.M, y.C, dy.ptr(), dy.ld, dy.coff, *go,
                   int(g_acc), self.stream, tag=(L.conv_name + ".dy", dy))
        if self.ddp is not None:
            self.ddp.note_grad_ready(plan, lst, bn.weight)
            self.ddp.note_grad_ready(plan, lst, bn.bias)

    def _bn_raw(self, plan, L: ConvL, count: int):
        """SfbBnRaw of layer L's training statistics (the conv's fused sums; folded layout when L runs pixel-folded)"""
        bn = L.bn
        folded = plan.stats_folded.get(L.conv_name, False)
        q = _ffi.BnRaw((L.sf if folded else L.stats).data_ptr(), int(count), bn.weight.data_ptr(), bn.bias.data_ptr(),
                       float(bn.eps), float(bn.momentum if bn.momentum is not None else 0.1),
                       bn.running_mean.data_ptr(), bn.running_var.data_ptr(), L.scale.data_ptr(), L.shift.data_ptr(),
                       L.mean.data_ptr(), L.invstd.data_ptr(), L.fold_f if folded else 0, L.ldo if folded else 0)
        plan.keep.append(q)
        return q

    def _apply_raw(self, plan, lst, L: ConvL, y: Act, out: Act, tag=None):
        """training: out = relu(BatchNorm(y)) with the finalize of L's batch statistics fused into the same launch"""
        q = self._bn_raw(plan, L, y.M)
        self._emit(lst, self.lib.sfb_bn_add_relu_raw, y.ptr(), y.ld, y.coff, C.byref(q), 0, 0, 0, None,
                   out.ptr(), out.ld, out.coff, 0, y.M, y.C, 1, self.stream, tag=tag)

    def _apply(self, lst, L: ConvL, y: Act, out: Act, tag=None):
        """out = relu(y * scale_L + shift_L)   (training: BatchNorm needs the batch statistics of the whole tensor
        before it can normalise, so the activated operand of the next conv is materialised by one HBM-bound pass)"""
        self._emit(lst, self.lib.sfb_bn_add_relu, y.ptr(), y.ld, y.coff, L.scale.data_ptr(), L.shift.data_ptr(), 0, 0, 0, 0, 0,
                   out.ptr(), out.ld, out.coff, 0, y.M, y.C, 1, self.stream, tag=tag)

    # ------------------------------------------------------------------ plan construction
    def build_plan(self, N, Ts, Tf, H, W, training) -> Plan:
        cfg = self.cfg
        dev = self.dev
        plan = Plan()
        self._bind_bn_arena(plan)
        pool = BufferPool(dev)
        fwd = plan.fwd

        def new_act(n, t, h, w, c, pad=True):
            """activation buffer; channel counts that are not a multiple of 16 get zeroed pad lanes (TMA operand width)"""
            ld = round_up(c, 16) if pad else c
            a = Act.empty(n, t, h, w, ld, device=dev, C=c, zero=(ld != c))
            plan.act_bytes += a.buf.numel() * 2
            return a

        def new_grad(like: Act):
            n, t, h, w = like.dims
            return Act(pool.alloc((n, t, h, w, like.ld)), like.coff, like.C)

        def fold(L):
            bn = L.bn
            self._emit(plan.folds, self.lib.sfb_bn_fold, bn.weight.data_ptr(), bn.bias.data_ptr(), bn.running_mean.data_ptr(),
                       bn.running_var.data_ptr(), float(bn.eps), L.Co, L.scale.data_ptr(), L.shift.data_ptr(), self.stream)

        def conv_bn_act(L, x: Act, out_z: Optional[Act], *, relu=True, residual=None, keep_raw=True):
            """conv + BatchNorm (+ReLU).  training: raw y (+ fused statistics) -> finalize -> [apply -> z];
            eval: folded BatchNorm (+residual, ReLU) in the GEMM epilogue, nothing else.
            returns (y_raw | None, z | None)"""
            n_, t_, h_, w_ = x.dims
            to, ho, wo = conv_out_dims((t_, h_, w_), L.k, L.s, L.p)
            if training:
                y = new_act(n_, to, ho, wo, L.Co)
                self._conv(plan, fwd, L, x, None, y, stats=True)
                if out_z is not None:
                    self._apply_raw(plan, fwd, L, y, out_z, tag=(L.conv_name + ".z", out_z))
                return y, out_z  # out_z is None: the residual tail finalises this layer's statistics
            fold(L)
            assert out_z is not None
            self._conv(plan, fwd, L, x, None, out_z, epi=(L.scale, L.shift, relu), residual=residual,
                       tag=(L.conv_name + ".z", out_z))
            return None, out_z

        T_in = (Ts, Tf)
        P = self.P
        # ---------------- stems ----------------
        stem_rec = []
        pooled: List[FinalAct] = []
        cat_bufs = {}
        for p in range(P):
            L = self.stems[p]
            To, Ho, Wo = conv_out_dims((T_in[p], H, W), L.k, L.s, L.p)
            y0 = new_act(N, To, Ho, Wo, L.Co, pad=False)  # biggest tensors of the net: consumed by the pool only
            if L.wps is not None and _ops.stem_s2d_eligible(L.Ci, L.Co, L.k, L.s, L.p, H, W):
                # space-to-depth clip + row-window kernel: every input row is staged once for all 49 taps
                cs = L.wps.shape[1] // 16
                x_in = torch.empty((N, T_in[p], H // 2, W // 2, cs), dtype=torch.bfloat16, device=dev)
                plan.act_bytes += x_in.numel() * 2
                plan.x_in.append(x_in)
                self._emit(fwd, self.lib.sfb_stem_conv_fprop, x_in.data_ptr(), N, T_in[p], H // 2, W // 2, L.Ci, cs,
                           L.wps.data_ptr(), L.wps.shape[0], y0.ptr(), y0.ld, y0.coff, L.Co,
                           L.stats.data_ptr() if training else 0, self.stream, tag=(L.conv_name + ".y", y0))
            else:
                x_in = new_act(N, T_in[p], H, W, L.Ci)
                plan.x_in.append(x_in)
                self._conv(plan, fwd, L, x_in, None, y0, stats=training)
            self._bn_stats_ready(plan, fwd, L, y0.M, training)
            Hp, Wp = (Ho - 1) // 2 + 1, (Wo - 1) // 2 + 1
            if p == 0 and self.laterals[0] is not None:
                lat = self.laterals[0]
                ci = CatInfo(new_act(N, To, Hp, Wp, L.Co + lat.Co), L.Co)
                cat_bufs[0] = ci
                out = FinalAct(ci.cat.view(0, L.Co))
                out.cat = ci
                ci.slow_out = out
            else:
                out = FinalAct(new_act(N, To, Hp, Wp, L.Co))
            idx = torch.empty((N, To, Hp, Wp, L.Co), dtype=torch.uint8, device=dev) if training else None
            self._emit(fwd, self.lib.sfb_maxpool_fwd, y0.ptr(), y0.ld, y0.coff, L.scale.data_ptr(), L.shift.data_ptr(),
                       N, To, Ho, Wo, L.Co, out.act.ptr(), out.act.ld, out.act.coff, idx.data_ptr() if training else 0,
                       self.stream, tag=(L.name + ".O", out.act))
            pooled.append(out)
            stem_rec.append(dict(L=L, x_in=x_in, y0=y0, out=out, idx=idx))

        lateral_rec = {}

        def do_lateral(si, slow_out: FinalAct, fast_out: FinalAct):
            """lateral conv after stage index si (0 = stems): relu(bn(conv(fast))) lands behind the Slow tensor in
            the concat buffer (torch.cat of reference model/my_slowfast.py:343 never happens)"""
            lat = self.laterals[si]
            ci = cat_bufs[si]
            assert slow_out is ci.slow_out
            zlat = ci.cat.view(ci.c_slow, lat.Co)
            ylat, _ = conv_bn_act(lat, fast_out.act, zlat)
            X = FinalAct(ci.cat)
            X.cat = ci
            ci.X = X
            lateral_rec[si] = dict(L=lat, x=fast_out, ylat=ylat, zlat=zlat, ci=ci)
            return X

        # ---------------- residual stages ----------------
        block_recs = []
        cur: List[FinalAct] = pooled
        if self.laterals[0] is not None:
            cur = [do_lateral(0, cur[0], cur[1]), cur[1]]
        for si in range(4):
            nxt: List[FinalAct] = [None] * P
            for p in reversed(range(P)):  # Fast first: its output feeds the lateral that completes the Slow concat buffer
                X = cur[p]
                blks = self.stages[si][p]
                for bi, d in enumerate(blks):
                    a, b, c, s_ = d["a"], d["b"], d["c"], d["s"]
                    n_, t_, h_, w_ = X.act.dims
                    ta, ha, wa = conv_out_dims((t_, h_, w_), a.k, a.s, a.p)
                    za = new_act(n_, ta, ha, wa, a.Co)
                    ya, _ = conv_bn_act(a, X.act, za)
                    tb, hb, wb = conv_out_dims((ta, ha, wa), b.k, b.s, b.p)
                    zb = new_act(n_, tb, hb, wb, b.Co)
                    yb, _ = conv_bn_act(b, za, zb)
                    last = bi == len(blks) - 1
                    if last and p == 0 and self.laterals[si + 1] is not None:
                        lat = self.laterals[si + 1]
                        ci = CatInfo(new_act(n_, tb, hb, wb, c.Co + lat.Co), c.Co)
                        cat_bufs[si + 1] = ci
                        O = FinalAct(ci.cat.view(0, c.Co))
                        O.cat = ci
                        ci.slow_out = O
                    else:
                        O = FinalAct(new_act(n_, tb, hb, wb, c.Co))
                    otag = (d["name"] + ".O", O.act)
                    yc = y1 = None
                    if training:
                        # packed ReLU mask of O for the backward tail (1 byte per 8 channels instead of re-reading O)
                        O.bits = torch.empty(O.act.M * (c.Co // 8), dtype=torch.uint8, device=dev)
                        plan.keep.append(O.bits)
                        plan.mask_bits.append((O.act, O.bits))
                        plan.act_bytes += O.bits.numel()
                        yc, _ = conv_bn_act(c, zb, None)
                        if s_ is not None:
                            y1, _ = conv_bn_act(s_, X.act, None)
                            qc, qs = self._bn_raw(plan, c, yc.M), self._bn_raw(plan, s_, y1.M)
                            self._emit(fwd, self.lib.sfb_bn_add_relu_raw, yc.ptr(), yc.ld, yc.coff, C.byref(qc),
                                       y1.ptr(), y1.ld, y1.coff, C.byref(qs),
                                       O.act.ptr(), O.act.ld, O.act.coff, O.bits.data_ptr(), yc.M, c.Co, 1, self.stream, tag=otag)
                        else:
                            qc = self._bn_raw(plan, c, yc.M)
                            self._emit(fwd, self.lib.sfb_bn_add_relu_raw, yc.ptr(), yc.ld, yc.coff, C.byref(qc),
                                       X.act.ptr(), X.act.ld, X.act.coff, None,
                                       O.act.ptr(), O.act.ld, O.act.coff, O.bits.data_ptr(), yc.M, c.Co, 1, self.stream, tag=otag)
                    else:
                        # eval: shortcut = folded-BN conv (no ReLU) or X itself; conv_c adds it and applies the ReLU
                        if s_ is not None:
                            S = new_act(n_, tb, hb, wb, s_.Co)
                            conv_bn_act(s_, X.act, S, relu=False)
                            res = S
                        else:
                            res = X.act
                        fold(c)
                        self._conv(plan, fwd, c, zb, None, O.act, epi=(c.scale, c.shift, True), residual=res, tag=otag)
                    block_recs.append(dict(si=si, p=p, d=d, X=X, ya=ya, za=za, yb=yb, zb=zb, yc=yc, y1=y1, O=O))
                    X = O
                nxt[p] = X
            cur = nxt
            if self.laterals[si + 1] is not None:
                Xs = do_lateral(si + 1, cur[0], cur[1])
                block_recs.append(dict(lateral=si + 1))
                cur = [Xs, cur[1]]

        # ---------------- head ----------------
        feat_dim = sum(fa.act.C for fa in cur)
        ncls = self.proj.out_features
        assert feat_dim == self.proj.in_features
        feat = _f32(N * feat_dim, dev).view(N, feat_dim)
        logits = _f32(N * ncls, dev).view(N, ncls)
        plan.logits = logits
        hk = cfg["head_pool_kernel_sizes"]
        drop = float(self.dropout_p) if training else 0.0
        plan.seed_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        seed_ptr = plan.seed_dev.data_ptr() if drop > 0.0 else None
        f_off = 0
        for p in range(P):
            a = cur[p].act
            n_, t_, h_, w_ = a.dims
            if hk[p][0] > t_ or hk[p][1] > h_ or hk[p][2] > w_:
                raise _ffi.SfbError(f"head pool kernel {hk[p]} larger than the final feature map {(t_, h_, w_)} "
                                    f"(same failure as nn.AvgPool3d in the reference)")
            self._emit(fwd, self.lib.sfb_head_pool_fwd, a.ptr(), a.ld, a.coff, n_, t_, h_, w_, a.C, hk[p][0], hk[p][1], hk[p][2],
                       drop, 0, seed_ptr, feat.data_ptr(), feat_dim, f_off, self.stream)
            f_off += a.C
        self._emit(fwd, self.lib.sfb_fc_fwd, feat.data_ptr(), self.proj.weight.data_ptr(), self.proj.bias.data_ptr(), N,
                   feat_dim, ncls, logits.data_ptr(), self.stream, tag=("logits", logits))
        plan.keep += [feat, logits]
        plan.feat = feat
        if not training:
            return plan

        # =============================== backward ===============================
        bwd = plan.bwd
        if self.ddp is not None:
            self.ddp.begin_plan()
        dlogits = _f32(N * ncls, dev).view(N, ncls)
        dfeat = _f32(N * feat_dim, dev).view(N, feat_dim)
        plan.dlogits = dlogits
        plan.keep += [dfeat]
        self._emit(bwd, self.lib.sfb_fc_bwd, dlogits.data_ptr(), feat.data_ptr(), self.proj.weight.data_ptr(), N, feat_dim, ncls,
                   dfeat.data_ptr(), self.grad_views[id(self.proj.weight)].data_ptr(),
                   self.grad_views[id(self.proj.bias)].data_ptr(), self.stream, tag=("dfeat", dfeat))
        if self.ddp is not None:
            self.ddp.note_grad_ready(plan, bwd, self.proj.weight)
            self.ddp.note_grad_ready(plan, bwd, self.proj.bias)

        def ensure_grad(fa: FinalAct):
            if fa.grad is not None:
                return
            if fa.cat is not None:
                ci = fa.cat
                ci.dcat = new_grad(ci.cat)
                ci.X.grad = ci.dcat
                ci.slow_out.grad = ci.dcat.view(0, ci.c_slow)
            else:
                fa.grad = new_grad(fa.act)

        def dgrad_into(L: ConvL, dy: Act, X: FinalAct):
            """dX (+)= conv_transpose(dy, W_L)"""
            ensure_grad(X)
            self._conv(plan, bwd, L, dy, None, X.grad, transposed=True, residual=X.grad if X.ginit else None)
            X.ginit = True

        def lateral_backward(si):
            lr = lateral_rec[si]
            lat, ci = lr["L"], lr["ci"]
            assert ci.X.ginit, "concat buffer without gradient contribution"
            ci.slow_out.ginit = True  # the Slow channels of dcat are now complete as well
            dz = ci.dcat.view(ci.c_slow, lat.Co)
            self._bn_bwd(plan, bwd, lat, dz, lr["ylat"], None, dz)  # in place: dz -> dy; ReLU mask recomputed from y
            self._wgrad(plan, bwd, lat, lr["x"].act, None, dz)
            dgrad_into(lat, dz, lr["x"])

        f_off = 0
        for p in range(P):
            fa = cur[p]
            ensure_grad(fa)
            a, ga = fa.act, fa.grad
            n_, t_, h_, w_ = a.dims
            self._emit(bwd, self.lib.sfb_head_pool_bwd, dfeat.data_ptr(), feat_dim, f_off, n_, t_, h_, w_, a.C,
                       hk[p][0], hk[p][1], hk[p][2], drop, 0, seed_ptr, ga.ptr(), ga.ld, ga.coff, self.stream)
            fa.ginit = True
            f_off += a.C

        for rec in reversed(block_recs):
            if "lateral" in rec:
                lateral_backward(rec["lateral"])
                continue
            d, X, O = rec["d"], rec["X"], rec["O"]
            a, b, c, s_ = d["a"], d["b"], d["c"], d["s"]
            ya, za, yb, zb, yc, y1 = rec["ya"], rec["za"], rec["yb"], rec["zb"], rec["yc"], rec["y1"]
            dO = O.grad
            assert O.ginit, "block output without gradient contribution"
            self._emit(bwd, _probe, tag=(d["name"] + ".dO", dO))
            dyc = new_grad(yc)
            if s_ is not None:
                # projection shortcut: both BatchNorms of the tail in one reduction + one apply pass (they share dO * mask)
                dy1 = new_grad(y1)
                sides = []
                for Lx, yx, dyx in ((c, yc, dyc), (s_, y1, dy1)):
                    sides.append(_ffi.BnTailSide(yx.ptr(), yx.ld, yx.coff, Lx.mean.data_ptr(), Lx.invstd.data_ptr(),
                                                 Lx.bn.weight.data_ptr(), Lx.dgamma.data_ptr(), Lx.dbeta.data_ptr(),
                                                 dyx.ptr(), dyx.ld, dyx.coff))
                plan.keep += sides
                self._emit(bwd, self.lib.sfb_bn_bwd_tail2, dO.ptr(), dO.ld, dO.coff, O.bits.data_ptr(), C.byref(sides[0]),
                           C.byref(sides[1]), yc.M, yc.C, self.stream, tag=(c.conv_name + ".dy", dyc))
                if self.ddp is not None:
                    for Lx in (c, s_):
                        self.ddp.note_grad_ready(plan, bwd, Lx.bn.weight)
                        self.ddp.note_grad_ready(plan, bwd, Lx.bn.bias)
            else:
                # identity shortcut: the masked gradient g also flows straight to the block input
                ensure_grad(X)
                self._bn_bwd(plan, bwd, c, dO, yc, O.bits, dyc, g_out=X.grad, g_acc=X.ginit)
                X.ginit = True
            self._wgrad(plan, bwd, c, zb, None, dyc)
            dzb = new_grad(yb)
            # the dgrad that produces dz of conv_b / conv_a also masks it with the layer's forward ReLU and reduces its
            # BatchNorm's dgamma / dbeta in the epilogue (the y tile arrives by TMA): bn_bwd_reduce is gone here
            fuse = self.fuse_bn_reduce and self.impl == _ffi.IMPL_TC
            self._conv(plan, bwd, c, dyc, None, dzb, transposed=True, bnred=(b, yb) if fuse else None)
            pool.release(dyc.buf)
            self._bn_bwd(plan, bwd, b, dzb, yb, None, dzb, premasked=fuse)  # in place: masked dz -> dy
            self._wgrad(plan, bwd, b, za, None, dzb)
            dza = new_grad(ya)
            self._conv(plan, bwd, b, dzb, None, dza, transposed=True, bnred=(a, ya) if fuse else None)
            pool.release(dzb.buf)
            self._bn_bwd(plan, bwd, a, dza, ya, None, dza, premasked=fuse)
            self._wgrad(plan, bwd, a, X.act, None, dza)
            # conv_a (stride 1) writes every pixel of dX first; the strided shortcut then only touches its own
            # sub-lattice with a reduce-add (no zero fill of the pixels a stride-2 conv never reads)
            dgrad_into(a, dza, X)
            pool.release(dza.buf)
            if s_ is not None:
                self._wgrad(plan, bwd, s_, X.act, None, dy1)
                dgrad_into(s_, dy1, X)
                pool.release(dy1.buf)
            pool.release(O.cat.dcat.buf if O.cat is not None else dO.buf)

        if self.laterals[0] is not None:
            lateral_backward(0)
        for p in range(P):
            sr = stem_rec[p]
            L, y0, out = sr["L"], sr["y0"], sr["out"]
            assert out.ginit
            n_, t_, h_, w_ = y0.dims
            dz0 = new_grad(y0)
            g = out.grad
            # (a fused max-pool-backward + BatchNorm-backward pair that rebuilds the pool gradient in both passes was
            # measured slower -- 3.1 ms against 0.8 + 1.0 ms: the gather through the arg-max bytes is the expensive part
            # and ran twice; profiles/r3j_bench.json)
            self._emit(bwd, self.lib.sfb_maxpool_bwd, g.ptr(), g.ld, g.coff, sr["idx"].data_ptr(), n_, t_, h_, w_, L.Co,
                       dz0.ptr(), dz0.ld, dz0.coff, self.stream)
            self._bn_bwd(plan, bwd, L, dz0, y0, None, dz0)
            if isinstance(sr["x_in"], Act):
                self._wgrad(plan, bwd, L, sr["x_in"], None, dz0)
            else:
                xs2 = sr["x_in"]
                self._emit(bwd, self.lib.sfb_stem_conv_wgrad, xs2.data_ptr(), xs2.shape[0], xs2.shape[1], xs2.shape[2],
                           xs2.shape[3], L.Ci, xs2.shape[4], dz0.ptr(), dz0.ld, dz0.coff, L.Co, L.dw.data_ptr(), self.stream)
                if self.ddp is not None:
                    self.ddp.note_grad_ready(plan, bwd, L.conv.weight)
            pool.release(dz0.buf)
            pool.release(out.cat.dcat.buf if out.cat is not None else g.buf)
            plan.keep.append(sr["idx"])
        plan.act_bytes += pool.total_bytes
        plan.keep.append(pool)
        return plan

    # ------------------------------------------------------------------ execution
    def _graphed(self, plan: Plan, which: str, body) -> bool:
        """Run one op list (`body()` enqueues it on the current stream: the ~350 C-ABI launches of a forward or
        backward pass plus the arena zero fills and, under data parallelism, the bucket all-reduces on the side stream)
        as a CUDA graph.  Every buffer, tensor map and kernel argument of a plan is static, so the list is captured once
        -- at the third run of the plan, after the kernels' one-time attribute calls -- and replayed afterwards: one
        graph launch instead of several hundred stream launches, no per-kernel launch gap on the device.  What changes
        from step to step stays outside: input layout (caller's pointers), the dropout seed word, dlogits.
        Returns False when the list must run eagerly (profiling / debug capture, gloo, a failed capture)."""
        if not self.use_graphs or self.dry_run or self.prof is not None or self.capture is not None:
            return False
        if self.graph_failed_all or plan.graph_failed.get(which):
            return False
        hit = plan.graphs.get(which)
        if hit is None:
            plan.graph_runs[which] = plan.graph_runs.get(which, 0) + 1
            if plan.graph_runs[which] <= self.graph_after:
                return False
            cur = torch.cuda.current_stream(self.dev)
            side = torch.cuda.Stream(self.dev)
            side.wait_stream(cur)
            g = torch.cuda.CUDAGraph()
            k0 = _ffi.launch_count()
            saved = self.stream.value
            try:
                with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
                    self.stream.value = torch.cuda.current_stream(self.dev).cuda_stream
                    body()
            except Exception as e:  # noqa: BLE001 -- capture is an optimisation: fall back to stream launches
                plan.graph_failed[which] = True
                self.graph_errors.append(f"{which}: {type(e).__name__}: {e}")
                self.stream.value = saved
                torch.cuda.synchronize(self.dev)
                return False
            self.stream.value = saved
            cur.wait_stream(side)
            hit = plan.graphs[which] = (g, _ffi.launch_count() - k0)
        g, nk = hit
        g.replay()
        self.lib.sfb_launch_count_add(nk)
        return True

    def _run(self, ops):
        cap = self.capture
        if self.prof is not None:
            return self._run_profiled(ops)
        for fn, args, tag in ops:
            rc = fn(*args)
            if rc:
                _ffi.check(rc, getattr(fn, "__name__", "libsfb200 call"))
            if cap is not None and tag is not None:
                t = tag[1]
                cap[tag[0]] = (t.torch_view() if isinstance(t, Act) else t).float().clone()

    def _run_profiled(self, ops):
        """bench support: CUDA events around every plan entry (on the launching stream); self.prof collects
        (entry name, start event, end event, op args) -- resolved to milliseconds by profile_summary()"""
        st = torch.cuda.current_stream(self.dev)
        for fn, args, tag in ops:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            k0 = _ffi.launch_count()
            e0.record(st)
            rc = fn(*args)
            e1.record(st)
            if rc:
                _ffi.check(rc, getattr(fn, "__name__", "libsfb200 call"))
            # kernel launches of this entry (a strided dgrad is one launch per parity class)
            self.prof.append((getattr(fn, "__name__", "op"), e0, e1, args, _ffi.launch_count() - k0))

    def profile_ops(self):
        """-> list of (entry name, layer description, ms, algorithmic GFLOP) per recorded plan entry"""
        torch.cuda.synchronize(self.dev)
        rows = []
        for name, e0, e1, args, _nk in self.prof:
            desc, gf = "", 0.0
            name = "sfb_conv_wgrad" if name == "sfb_conv_wgrad_folded" else name
            if name in ("sfb_conv_gemm", "sfb_conv_wgrad", "sfb_conv_dgrad_bn"):
                g = args[0]._obj
                kind = "wgrad" if name == "sfb_conv_wgrad" else ("dgrad+bn" if name == "sfb_conv_dgrad_bn" else
                                                                  ("dgrad" if g.transposed else "fprop"))
                M = getattr(g, "real_m", None) or (g.N * g.Ti * g.Hi * g.Wi if g.transposed else g.N * g.To * g.Ho * g.Wo)
                gf = 2.0 * M * g.real_n * g.real_k / 1e9
                desc = (f"{kind} in({g.Ti},{g.Hi},{g.Wi})x{g.Cg} out({g.To},{g.Ho},{g.Wo})x{g.Co} "
                        f"k({g.kT},{g.kH},{g.kW}) s({g.sT},{g.sH},{g.sW})")
            if name in ("sfb_stem_conv_fprop", "sfb_stem_conv_wgrad"):
                ci, co = args[5], (args[12] if name == "sfb_stem_conv_fprop" else args[10])
                gf = 2.0 * args[1] * args[2] * args[3] * args[4] * co * ci * 49 / 1e9
                desc = f"stem s2d in({args[2]},{2 * args[3]},{2 * args[4]})x{ci} out x{co} k(1,7,7) s(1,2,2)"
            ms = e0.elapsed_time(e1)
            # HBM-bound BatchNorm passes: tensors moved (bf16 [M][C]) and the bandwidth that implies
            nt = 0
            if name == "sfb_bn_add_relu":      # y (+ r) -> out
                M, Cc, nt = args[14], args[15], 2 + (1 if args[5] else 0)
            elif name == "sfb_bn_add_relu_raw":
                M, Cc, nt = args[12], args[13], 2 + (1 if args[4] else 0)
            elif name == "sfb_bn_bwd_reduce":  # dz, y (+ bf16 mask)
                M, Cc, nt = args[15], args[16], 2 + (1 if args[6] else 0)
            elif name == "sfb_bn_bwd_apply":   # dz, y (+ mask) -> dy (+ g_out (+ old g))
                M, Cc = args[18], args[19]
                nt = 3 + (1 if args[6] else 0) + (1 if args[23] else 0) + (1 if args[23] and args[26] else 0)
            if nt:
                desc = f"M={M} C={Cc} tensors={nt} {nt * M * Cc * 2 / ms / 1e6:7.0f} GB/s"
            rows.append((name, desc, ms, gf))
        return rows

    def profile_summary(self):
        """-> {entry name: (launches, total ms)}, plus conv FLOPs per class, from the events in self.prof"""
        torch.cuda.synchronize(self.dev)
        out, flops = {}, {}
        for name, e0, e1, args, nk in self.prof:
            ms = e0.elapsed_time(e1)
            name = "sfb_conv_wgrad" if name == "sfb_conv_wgrad_folded" else name
            key = name
            if name in ("sfb_conv_gemm", "sfb_conv_wgrad", "sfb_conv_dgrad_bn"):
                g = args[0]._obj
                if name == "sfb_conv_dgrad_bn":
                    key = "sfb_conv_dgrad_bn"
                elif name == "sfb_conv_gemm" and g.transposed:
                    key = "sfb_conv_gemm[dgrad]"
                elif name == "sfb_conv_gemm":
                    key = "sfb_conv_gemm[fprop]"
                # algorithmic work: rows of the forward conv's OUTPUT (for dgrad that is the gathered tensor dY)
                M = getattr(g, "real_m", None) or (g.N * g.Ti * g.Hi * g.Wi if g.transposed else g.N * g.To * g.Ho * g.Wo)
                flops[key] = flops.get(key, 0.0) + 2.0 * M * g.real_n * g.real_k
                # algorithmic bytes: gathered tensor + written tensor (bf16), + the old value when accumulating in place
                byt = 2.0 * g.N * (g.Ti * g.Hi * g.Wi * g.Cg + g.To * g.Ho * g.Wo * g.Co)
                if name == "sfb_conv_gemm" and args[7]._obj.residual:
                    byt += 2.0 * g.N * g.To * g.Ho * g.Wo * g.Co
                if name == "sfb_conv_dgrad_bn":  # the y tile the fused BatchNorm-backward reduction reads
                    byt += 2.0 * g.N * g.To * g.Ho * g.Wo * g.Co
                flops[key + "#bytes"] = flops.get(key + "#bytes", 0.0) + byt
                flops[key + "#kernels"] = flops.get(key + "#kernels", 0.0) + nk
            if name in ("sfb_stem_conv_fprop", "sfb_stem_conv_wgrad"):
                ci, co = args[5], (args[12] if name == "sfb_stem_conv_fprop" else args[10])
                flops[name] = flops.get(name, 0.0) + 2.0 * args[1] * args[2] * args[3] * args[4] * co * ci * 49
            n, t = out.get(key, (0, 0.0))
            out[key] = (n + 1, t + ms)
        return out, flops

    # ------------------------------------------------------------------ debug / test support
    def tagged_buffers(self, plan: Plan) -> Dict[str, object]:
        """name -> Act | tensor of every tagged forward intermediate (raw conv outputs `<conv>.y`, block outputs
        `<block>.O`, stem pool outputs, logits)"""
        return {tag[0]: tag[1] for _, _, tag in plan.fwd if tag is not None}

    def force_forward_state(self, plan: Plan, tensors: Dict[str, torch.Tensor], feat: Optional[torch.Tensor] = None):
        """TEST SUPPORT (teacher forcing): overwrite the saved forward state of a training plan with externally
        computed values -- NDHWC tensors keyed like tagged_buffers(); BatchNorm batch statistics are re-derived
        from the forced raw conv outputs -- so that a following backward() can be compared with a reference
        backward evaluated on the very same forward state (no ReLU-mask divergence)."""
        bufs = self.tagged_buffers(plan)
        self._bind_bn_arena(plan)  # l.mean / l.scale / ... below are this plan's
        by_conv = {l.conv_name: l for l in self.layers}
        for name, t in tensors.items():
            dst = bufs[name]
            view = dst.torch_view() if isinstance(dst, Act) else dst
            view.copy_(t.to(view.dtype))
            if name.endswith(".y"):
                L = by_conv[name[:-2]]
                yf = view.float().reshape(-1, L.Co)
                mean = yf.mean(0)
                var = yf.var(0, unbiased=False)
                invstd = torch.rsqrt(var + L.bn.eps)
                L.mean.copy_(mean)
                L.invstd.copy_(invstd)
                L.scale.copy_(L.bn.weight.detach() * invstd)
                L.shift.copy_(L.bn.bias.detach() - mean * L.bn.weight.detach() * invstd)
                # the raw sums the fused-finalize kernels read (re-run below for the activated operands)
                cnt = float(yf.shape[0])
                sums = torch.cat([mean * cnt, (var + mean * mean) * cnt])
                if plan.stats_folded.get(L.conv_name, False):
                    L.sf.zero_()
                    L.sf[:L.Co].copy_(sums[:L.Co])
                    L.sf[L.fold_f * L.ldo:L.fold_f * L.ldo + L.Co].copy_(sums[L.Co:])
                else:
                    L.stats.copy_(sums)
        if feat is not None:
            plan.feat.copy_(feat)
        # activated operands (z = relu(bn(y))) and the stem pools (+ arg-max indices) derive from the forced raw
        # outputs: recompute them with the engine's own kernels
        self._run([op for op in plan.fwd if op[0] is self.lib.sfb_maxpool_fwd or
                   (op[0] in (self.lib.sfb_bn_add_relu, self.lib.sfb_bn_add_relu_raw) and op[2] is not None and
                    op[2][0].endswith(".z"))])
        # packed ReLU masks of the block outputs, from the forced O tensors
        for O_act, bits in plan.mask_bits:
            o = O_act.torch_view().reshape(-1, O_act.C // 8, 8) > 0
            w = (2 ** torch.arange(8, device=o.device, dtype=torch.int32)).view(1, 1, 8)
            bits.copy_((o.to(torch.int32) * w).sum(-1).to(torch.uint8).reshape(-1))

    def get_plan(self, xs, xf, training) -> Plan:
        N, Cs, Ts, H, W = xs.shape
        if self.P == 1:
            if Cs != self.cfg["input_channels"][0]:
                raise ValueError(f"expected {self.cfg['input_channels'][0]} input channels, got {Cs}")
            key = (N, Ts, 0, H, W, bool(training), self.impl)
            plan = self.plans.get(key)
            if plan is None:
                self.refresh_weights(training)
                plan = self.build_plan(N, Ts, 0, H, W, training)
                self.plans[key] = plan
            return plan
        N2, Cf, Tf, H2, W2 = xf.shape
        if (N2, H2, W2) != (N, H, W):
            raise ValueError("slow and fast clips must share batch and spatial size")
        if (Cs, Cf) != tuple(self.cfg["input_channels"]):
            raise ValueError(f"expected input channels {self.cfg['input_channels']}, got {(Cs, Cf)}")
        lat_kT, lat_sT = self.cfg["fusion_conv_kernel"][0], self.cfg["fusion_conv_stride"][0]
        t_lat = (Tf + 2 * (lat_kT // 2) - lat_kT) // lat_sT + 1
        if self.fuse and t_lat != Ts:
            raise RuntimeError(f"Sizes of tensors must match except in dimension 1. Expected size {Ts} but got size "
                               f"{t_lat} (lateral output vs Slow pathway; same failure as torch.cat at reference "
                               f"model/my_slowfast.py:343)")
        key = (N, Ts, Tf, H, W, bool(training), self.impl)
        plan = self.plans.get(key)
        if plan is None:
            self.refresh_weights(training)
            plan = self.build_plan(N, Ts, Tf, H, W, training)
            self.plans[key] = plan
        return plan

    def _stage_raw(self, raw, training: bool) -> Plan:
        """uint8 clips -> the stems' s2d operands (sfb_stage_clips_u8; reference dataset/chalearn_dataset.py:41-46,
        79-88,113-118 + train.py:136-140 in one device pass)"""
        raw.validate()
        if raw.frames.device != self.dev or (raw.crop_yx is not None and raw.crop_yx.device != self.dev):
            raise _ffi.SfbError("inputs must live on the model's CUDA device")
        N, T, H, W, Ct = raw.frames.shape
        pws = raw.pathways()
        Cs_, Cf_ = self.cfg["input_channels"]
        if (pws[0][1], pws[1][1]) != (Cs_, Cf_):
            raise ValueError(f"expected input channels {(Cs_, Cf_)}, got {(pws[0][1], pws[1][1])}")
        if max(pws[0][0] + pws[0][1], pws[1][0] + pws[1][1]) > Ct:
            raise ValueError("channel ranges exceed the channels of the raw frames")
        # shape-only stand-ins: get_plan reads .shape
        xs = torch.empty((N, Cs_, T // raw.slow_t_stride, H, W), device="meta")
        xf = torch.empty((N, Cf_, T, H, W), device="meta")
        plan = self.get_plan(xs, xf, training)
        if any(isinstance(a, Act) for a in plan.x_in):
            raise _ffi.SfbError("uint8 staging needs both stems in space-to-depth form ((1,7,7)/s(1,2,2) stems, even H, W)")
        arr = (_ffi.StagePathway * 2)()
        for i, ((c0, Cc, ts), a) in enumerate(zip(pws, plan.x_in)):
            arr[i].c0, arr[i].C, arr[i].t_stride, arr[i].Cs, arr[i].out = c0, Cc, ts, a.shape[4], a.data_ptr()
        _ffi.check(self.lib.sfb_stage_clips_u8(raw.frames.data_ptr(), N, T, H, W, Ct,
                                               None if raw.crop_yx is None else raw.crop_yx.data_ptr(),
                                               float(raw.mean), float(raw.std), arr, 2, self.stream), "sfb_stage_clips_u8")
        return plan

    def forward(self, xs, xf, training: bool) -> Tuple[torch.Tensor, Plan]:
        """xs, xf: fp32 NCTHW clips (any strides), or xs = staging.RawClips and xf = None"""
        self.stream.value = torch.cuda.current_stream(self.dev).cuda_stream
        self._check_baked_pointers()
        if xf is None and not torch.is_tensor(xs):
            self.refresh_weights(training)
            plan = self._stage_raw(xs, training)
            return self._forward_staged(plan, training)
        ins = (xs,) if self.P == 1 else (xs, xf)
        if any(x.dtype != torch.float32 for x in ins):
            raise TypeError("inputs must be float32 (as produced by the reference loader)")
        if any(x.device != self.dev for x in ins):
            raise _ffi.SfbError("inputs must live on the model's CUDA device")
        plan = self.get_plan(xs, xf, training)
        self.refresh_weights(training)
        for x, a in zip(ins, plan.x_in):
            N, Cc, T, H, W = x.shape
            sn, sc, st, sh, sw = x.stride()
            if isinstance(a, Act):
                _ffi.check(self.lib.sfb_layout_input(x.data_ptr(), sn, sc, st, sh, sw, N, Cc, T, H, W, a.ptr(), a.ld, self.stream),
                           "sfb_layout_input")
            else:
                _ffi.check(self.lib.sfb_layout_input_s2d(x.data_ptr(), sn, sc, st, sh, sw, N, Cc, T, H, W, a.data_ptr(),
                                                         a.shape[4], self.stream), "sfb_layout_input_s2d")
        return self._forward_staged(plan, training)

    def _forward_staged(self, plan: Plan, training: bool) -> Tuple[torch.Tensor, Plan]:
        if training:
            plan.gen += 1
            if self.dropout_p > 0.0:
                plan.seed_dev.fill_(self._draw_dropout_seed())

        else:
            key = self._fold_key()
            if plan.fold_key != key:
                self._run(plan.folds)
                plan.fold_key = key

        def body():
            if training:
                plan.stats_flat.zero_()
            self._run(plan.fwd)

        if not self._graphed(plan, "fwd", body):
            body()
        if training:
            self.bn_epoch += 1  # running statistics moved (written through raw pointers: no version counter sees it)
        if training:
            torch._foreach_add_([bn.num_batches_tracked for bn in self.bn_modules], 1)
        return plan.logits, plan

    def _draw_dropout_seed(self) -> int:
        """64-bit seed of this step's head dropout mask, drawn from torch's default generator -- so torch.manual_seed
        reproduces a run, successive steps / runs / resumed runs differ, as with the reference's nn.Dropout
        (ResNetBasicHead; it draws from the process's generator at every call) -- and mixed with the data-parallel
        rank so that replicas seeded alike do not drop the same units of their local sample n."""
        v = int(torch.randint(-(1 << 62), 1 << 62, (1,), dtype=torch.int64))
        self.step_seed += 1
        rank = 0
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            rank = torch.distributed.get_rank()
        z = (v ^ ((rank + 1) * 0x9E3779B97F4A7C15)) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        z ^= z >> 31
        return z - (1 << 64) if z >= (1 << 63) else z  # as a signed int64 for Tensor.fill_

    def backward(self, plan: Plan, dlogits: torch.Tensor):
        self.stream.value = torch.cuda.current_stream(self.dev).cuda_stream
        plan.dlogits.copy_(dlogits)
        self._bwd_since_pack = True  # an optimizer step may follow: see refresh_weights
        # A parameter whose .grad still aliases the arena (autograd adopted last step's view and the caller did not
        # zero_grad) would be wiped by the arena zero fill below and then added to itself: give it storage of its own,
        # so that gradient accumulation over several backward passes behaves as with the reference's modules.
        for prm in self.live_params:
            g = prm.grad
            if g is not None and g.data_ptr() == self.grad_views[id(prm)].data_ptr():
                prm.grad = g.clone()

        if self.ddp is None or (self.graphs_with_ddp and self.ddp.bucketer.use_avg):
            def body():
                self.grad_flat.zero_()
                if self.ddp is not None:
                    self.ddp.begin_backward(self)
                self._run(plan.bwd)
                if self.ddp is not None:
                    self.ddp.finish_backward(self)

            if not self._graphed(plan, "bwd", body):
                body()
        else:
            # data parallel: the bucket all-reduces are launched by the host between graph SEGMENTS (capturing NCCL's
            # side-stream collectives inside one graph hung on this pool's driver / NCCL 2.28 -- profiles/ r3h; gloo
            # collectives are host-driven anyway): the op list is cut at every bucket launch, each piece is its own graph
            if plan.bwd_segments is None:
                segs, cur_ops = [], []
                for op in plan.bwd:
                    if getattr(op[0], "is_bucket_launch", False):
                        segs.append((cur_ops, op[0]))
                        cur_ops = []
                    else:
                        cur_ops.append(op)
                segs.append((cur_ops, None))
                plan.bwd_segments = segs
            self.ddp.begin_backward(self)
            for i, (ops, launch) in enumerate(plan.bwd_segments):
                def body(ops=ops, first=(i == 0)):
                    if first:
                        self.grad_flat.zero_()
                    self._run(ops)

                if not ops and i > 0:
                    pass
                elif not self._graphed(plan, ("bwd", i), body):
                    body()
                if launch is not None:
                    launch()
            self.ddp.finish_backward(self)
        # FRESH view objects, referenced by nobody else once returned: autograd then adopts them as .grad (no copy) instead
        # of cloning every gradient -- 332 device-to-device copies per step with views that the engine also holds
        flat = self.grad_flat
        return [torch.as_strided(flat, shp, std, off) for shp, std, off in self._grad_view_specs]


class _SlowFastFn(torch.autograd.Function):
    """forward + backward of the whole network as ONE autograd node (no per-layer dispatcher overhead)."""

    @staticmethod
    def forward(ctx, engine: SlowFastEngine, xs, xf, *params):
        training = engine.model.training
        logits, plan = engine.forward(xs, xf, training)
        ctx.engine, ctx.plan, ctx.gen = engine, plan, plan.gen
        ctx.training = training
        return logits.clone()

    @staticmethod
    def backward(ctx, dlogits):
        if not ctx.training:
            raise RuntimeError("backward through an eval-mode forward of the B200 SlowFast is not supported: "
                               "call model.train() before the forward pass")
        if ctx.plan.gen != ctx.gen:
            raise RuntimeError("a newer training forward of the same shape overwrote the saved activations; "
                               "run backward before the next forward (one forward in flight per input shape)")
        grads = ctx.engine.backward(ctx.plan, dlogits.contiguous().float())
        return (None, None, None) + tuple(grads)


def run_model(engine: SlowFastEngine, xs, xf):
    if torch.is_grad_enabled() and any(p.requires_grad for p in engine.live_params):
        return _SlowFastFn.apply(engine, xs, xf, *engine.live_params)
    logits, _ = engine.forward(xs, xf, engine.model.training)
    return logits.clone()
