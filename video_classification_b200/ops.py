"""Tensor-level wrappers over the C ABI (include/sfb200.h).

`Act` is a channel view (buffer, channel offset, channels) of a bf16 NDHWC buffer -- the unit every
kernel takes as (ptr, ld, coff).  torch owns memory and streams; all arithmetic happens in libsfb200.so.
"""
import ctypes as C
from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from . import _ffi
from ._ffi import ConvGeom, Epilogue, Prologue, check, lib, ptr, stream_ptr


@dataclass
class Act:
    buf: torch.Tensor  # (N, T, H, W, ld) bf16 contiguous
    coff: int
    C: int

    @property
    def ld(self):
        return self.buf.shape[-1]

    @property
    def dims(self):
        return tuple(self.buf.shape[:4])

    @property
    def M(self):
        n, t, h, w = self.dims
        return n * t * h * w

    def ptr(self):
        return self.buf.data_ptr()

    def torch_view(self):
        return self.buf[..., self.coff:self.coff + self.C]

    @staticmethod
    def empty(N, T, H, W, ld, device="cuda", coff=0, C=None, zero=False):
        f = torch.zeros if zero else torch.empty
        return Act(f((N, T, H, W, ld), dtype=torch.bfloat16, device=device), coff, ld if C is None else C)

    def view(self, coff, C):
        return Act(self.buf, coff, C)


def round_up(x, m):
    return (x + m - 1) // m * m


def conv_out_dims(dims, k, s, p):
    return tuple((d + 2 * pp - kk) // ss + 1 for d, kk, ss, pp in zip(dims, k, s, p))


def layout_input(x: torch.Tensor, ld: int) -> Act:
    """fp32 (N,C,T,H,W) with arbitrary strides -> bf16 NDHWC buffer with `ld` channels (zero padded)."""
    assert x.dtype == torch.float32 and x.dim() == 5 and x.is_cuda
    N, Cc, T, H, W = x.shape
    out = Act.empty(N, T, H, W, ld, device=x.device, C=ld)
    sn, sc, st, sh, sw = x.stride()
    check(lib().sfb_layout_input(x.data_ptr(), sn, sc, st, sh, sw, N, Cc, T, H, W, out.ptr(), ld, stream_ptr()),
          "sfb_layout_input")
    return out


def pack_weight(w: torch.Tensor, Cg: int, transposed: bool = False, out: Optional[torch.Tensor] = None):
    """fp32 OIDHW conv weight -> packed bf16 [rows_pad][k_pad] (k = tap*Cg + c)."""
    assert w.dtype == torch.float32 and w.is_cuda and w.is_contiguous()
    Co, Ci = w.shape[0], w.shape[1]
    taps = w[0, 0].numel()
    rows = Ci if transposed else Co
    rows_pad, k_pad = _ffi.packed_weight_dims(rows, Cg, taps)
    if out is None:
        out = torch.empty((rows_pad, k_pad), dtype=torch.bfloat16, device=w.device)
    assert tuple(out.shape) == (rows_pad, k_pad)
    check(lib().sfb_pack_weight(w.data_ptr(), Co, Ci, taps, Cg, int(transposed), out.data_ptr(), rows_pad, k_pad,
                                stream_ptr()), "sfb_pack_weight")
    return out


def pack_weight_dgrad(w: torch.Tensor, Cg: int, stride, out: Optional[torch.Tensor] = None):
    """fp32 OIDHW conv weight -> class-major packed bf16 operand of the tensor-core input-gradient GEMM."""
    assert w.dtype == torch.float32 and w.is_cuda and w.is_contiguous()
    Co, Ci = w.shape[0], w.shape[1]
    k = tuple(w.shape[2:])
    rows_pad, k_pad = _ffi.packed_weight_dims_dgrad(Ci, Cg, k, stride)
    if out is None:
        out = torch.empty((rows_pad, k_pad), dtype=torch.bfloat16, device=w.device)
    assert tuple(out.shape) == (rows_pad, k_pad)
    check(lib().sfb_pack_weight_dgrad(w.data_ptr(), Co, Ci, k[0], k[1], k[2], stride[0], stride[1], stride[2], Cg,
                                      out.data_ptr(), rows_pad, k_pad, stream_ptr()), "sfb_pack_weight_dgrad")
    return out


# ---------------------------------------------------------------- pixel folding (narrow-channel layers)
def fold_factor(ld_in: int, ld_out: int, k, s, W: int) -> int:
    """how many W-neighbouring pixels to view as one (1 = do not fold): rows of >= 128 bytes for TMA on both sides"""
    if tuple(s) != (1, 1, 1) or k[2] not in (1, 3):
        return 1
    m = min(ld_in, ld_out)
    if m >= 64:
        return 1
    f = 64 // m
    while f > 1 and (W % f or f * max(ld_in, ld_out) > 512):
        f //= 2
    return f


def folded(a: Act, f: int) -> Act:
    """the same memory with f neighbouring W pixels as one pixel of f*ld channels (a must be a whole dense buffer)"""
    N, T, H, W = a.dims
    assert a.coff == 0 and W % f == 0
    return Act(a.buf.view(N, T, H, W // f, f * a.ld), 0, f * a.ld)


def fold_weight(w: torch.Tensor, f: int, ldo: int, ldi: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    Co, Ci, kT, kH, kW = w.shape
    kWf = 1 if kW == 1 else 3
    if out is None:
        out = torch.empty((f * ldo, f * ldi, kT, kH, kWf), dtype=torch.float32, device=w.device)
    check(lib().sfb_fold_weight(w.data_ptr(), Co, Ci, kT, kH, kW, f, ldo, ldi, out.data_ptr(), stream_ptr()), "sfb_fold_weight")
    return out


def unfold_wgrad(dwf: torch.Tensor, f: int, ldo: int, ldi: int, dw: torch.Tensor):
    Co, Ci, kT, kH, kW = dw.shape
    check(lib().sfb_unfold_wgrad(dwf.data_ptr(), Co, Ci, kT, kH, kW, f, ldo, ldi, dw.data_ptr(), stream_ptr()), "sfb_unfold_wgrad")
    return dw


def unfold_stats(sf: torch.Tensor, f: int, ld: int, C_: int, stats: torch.Tensor):
    check(lib().sfb_unfold_stats(sf.data_ptr(), f, ld, C_, stats.data_ptr(), stream_ptr()), "sfb_unfold_stats")
    return stats


def stem_s2d_eligible(C_in, C_out, k, s, p, H, W) -> bool:
    """can the (1,7,7)/s(1,2,2)/p(0,3,3) stem take the space-to-depth row-window kernels (csrc/stem_win.cuh)?"""
    return (tuple(k) == (1, 7, 7) and tuple(s) == (1, 2, 2) and tuple(p) == (0, 3, 3) and H % 2 == 0 and W % 2 == 0
            and W // 2 <= 125 and 4 * C_in <= 64 and C_out <= (64 if 4 * C_in <= 32 else 32) and C_out % 8 == 0)


def stem_s2d_channels(C_in: int) -> int:
    cs = lib().sfb_stem_s2d_channels(C_in)
    if cs < 0:
        check(cs, "sfb_stem_s2d_channels")
    return cs


def stem_rows_pad(C_out: int) -> int:
    return 16 if C_out <= 16 else (32 if C_out <= 32 else 64)


def layout_input_s2d(x: torch.Tensor) -> torch.Tensor:
    """fp32 (N,C,T,H,W), arbitrary strides -> s2d bf16 (N,T,H/2,W/2,Cs)"""
    assert x.dtype == torch.float32 and x.dim() == 5 and x.is_cuda
    N, Cc, T, H, W = x.shape
    cs = stem_s2d_channels(Cc)
    out = torch.empty((N, T, H // 2, W // 2, cs), dtype=torch.bfloat16, device=x.device)
    sn, sc, st, sh, sw = x.stride()
    check(lib().sfb_layout_input_s2d(x.data_ptr(), sn, sc, st, sh, sw, N, Cc, T, H, W, out.data_ptr(), cs, stream_ptr()),
          "sfb_layout_input_s2d")
    return out


def stage_clips_u8(frames: torch.Tensor, pathways, crop_yx: Optional[torch.Tensor] = None, mean: float = 0.45,
                   std: float = 0.225, outs=None):
    """uint8 (N,T,H,W,Ct) raw frames -> one s2d bf16 (N,T/t_stride,H/2,W/2,Cs) operand per pathway.
    pathways: sequence of (c0, C, t_stride).  Replaces the host-side ToTensor/Normalize/RandomCrop/permute/split of
    reference dataset/chalearn_dataset.py:41-46,79-88,113-118 and train.py:136-140."""
    assert frames.dtype == torch.uint8 and frames.dim() == 5 and frames.is_cuda and frames.is_contiguous()
    N, T, H, W, Ct = frames.shape
    pws = (_ffi.StagePathway * len(pathways))()
    res = []
    for i, (c0, Cc, ts) in enumerate(pathways):
        cs = stem_s2d_channels(Cc)
        if outs is not None:
            o = outs[i]
            assert o.dtype == torch.bfloat16 and o.is_contiguous() and tuple(o.shape) == (N, T // ts, H // 2, W // 2, cs)
        else:
            o = torch.empty((N, T // max(ts, 1), H // 2, W // 2, cs), dtype=torch.bfloat16, device=frames.device)
        res.append(o)
        pws[i].c0, pws[i].C, pws[i].t_stride, pws[i].Cs, pws[i].out = c0, Cc, ts, cs, o.data_ptr()
    cptr = None
    if crop_yx is not None:
        assert crop_yx.dtype == torch.int32 and crop_yx.is_cuda and crop_yx.is_contiguous() and tuple(crop_yx.shape) == (N, 2)
        cptr = crop_yx.data_ptr()
    check(lib().sfb_stage_clips_u8(frames.data_ptr(), N, T, H, W, Ct, cptr, mean, std, pws, len(pathways), stream_ptr()),
          "sfb_stage_clips_u8")
    return res


def pack_weight_stem(w: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    assert w.dtype == torch.float32 and w.is_cuda and w.is_contiguous() and tuple(w.shape[2:]) == (1, 7, 7)
    Co, Ci = w.shape[0], w.shape[1]
    cs, rp = stem_s2d_channels(Ci), stem_rows_pad(Co)
    if out is None:
        out = torch.empty((rp, 16 * cs), dtype=torch.bfloat16, device=w.device)
    check(lib().sfb_pack_weight_stem(w.data_ptr(), Co, Ci, cs, out.data_ptr(), rp, stream_ptr()), "sfb_pack_weight_stem")
    return out


def stem_conv_fprop(xs2d: torch.Tensor, Ci: int, wp: torch.Tensor, out: Act, stats: Optional[torch.Tensor] = None):
    N, T, H2, W2, cs = xs2d.shape
    check(lib().sfb_stem_conv_fprop(xs2d.data_ptr(), N, T, H2, W2, Ci, cs, wp.data_ptr(), wp.shape[0], out.ptr(), out.ld,
                                    out.coff, out.C, ptr(stats), stream_ptr()), "sfb_stem_conv_fprop")
    return out


def stem_conv_wgrad(xs2d: torch.Tensor, Ci: int, dy: Act, dw: torch.Tensor):
    assert dw.dtype == torch.float32 and dw.is_contiguous()
    N, T, H2, W2, cs = xs2d.shape
    check(lib().sfb_stem_conv_wgrad(xs2d.data_ptr(), N, T, H2, W2, Ci, cs, dy.ptr(), dy.ld, dy.coff, dy.C, dw.data_ptr(),
                                    stream_ptr()), "sfb_stem_conv_wgrad")
    return dw


def make_geom(x: Act, out: Act, k, s, p, transposed=False) -> ConvGeom:
    N, Ti, Hi, Wi = x.dims
    N2, To, Ho, Wo = out.dims
    assert N == N2
    return ConvGeom(N, Ti, Hi, Wi, x.C, To, Ho, Wo, out.C, k[0], k[1], k[2], s[0], s[1], s[2], p[0], p[1], p[2],
                    x.ld, x.coff, out.ld, out.coff, int(transposed))


def conv_gemm(x: Act, wp: torch.Tensor, out: Act, k, s, p, *, transposed=False,
              pro: Optional[Tuple[torch.Tensor, torch.Tensor, bool]] = None,
              epi_affine: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, residual: Optional[Act] = None,
              relu=False, stats: Optional[torch.Tensor] = None, impl=_ffi.IMPL_TC):
    g = make_geom(x, out, k, s, p, transposed)
    pr = Prologue(ptr(pro[0]), ptr(pro[1]), int(pro[2])) if pro is not None else Prologue(0, 0, 0)
    ep = Epilogue(ptr(epi_affine[0]) if epi_affine else 0, ptr(epi_affine[1]) if epi_affine else 0,
                  residual.ptr() if residual else 0, residual.ld if residual else 0,
                  residual.coff if residual else 0, int(relu), ptr(stats))
    check(lib().sfb_conv_gemm(C.byref(g), x.ptr(), C.byref(pr), wp.data_ptr(), wp.shape[0], wp.shape[1], out.ptr(),
                              C.byref(ep), impl, stream_ptr()), "sfb_conv_gemm")
    return out


def conv_dgrad_bn(dy: Act, wpt: torch.Tensor, out: Act, k, s, p, y: Act, scale, shift, mean, invstd, dgamma, dbeta,
                  C_real: Optional[int] = None, fold_ld: int = 0):
    """out = masked input gradient of a conv whose INPUT was relu(BatchNorm(y)); dgamma / dbeta (fp32 [C]) accumulate the
    BatchNorm backward sums (sfb_conv_dgrad_bn).  y has the dims and view layout of `out`."""
    g = make_geom(dy, out, k, s, p, True)
    assert y.dims == out.dims
    fz = _ffi.BnBwdFuse(y.ptr(), y.ld, y.coff, ptr(scale), ptr(shift), ptr(mean), ptr(invstd), ptr(dgamma), ptr(dbeta),
                        int(C_real if C_real is not None else out.C), int(fold_ld))
    check(lib().sfb_conv_dgrad_bn(C.byref(g), dy.ptr(), wpt.data_ptr(), wpt.shape[0], wpt.shape[1], out.ptr(), C.byref(fz),
                                  stream_ptr()), "sfb_conv_dgrad_bn")
    return out


def conv_wgrad(x: Act, dy: Act, dw: torch.Tensor, k, s, p, *, pro=None, impl=_ffi.IMPL_TC):
    """dw (fp32, (Co, Ci, kT, kH, kW), contiguous) += wgrad."""
    assert dw.dtype == torch.float32 and dw.is_contiguous()
    g = make_geom(x, dy, k, s, p, False)
    pr = Prologue(ptr(pro[0]), ptr(pro[1]), int(pro[2])) if pro is not None else Prologue(0, 0, 0)
    check(lib().sfb_conv_wgrad(C.byref(g), x.ptr(), C.byref(pr), dy.ptr(), dy.ld, dy.coff, dw.data_ptr(),
                               dw.shape[1], impl, stream_ptr()), "sfb_conv_wgrad")
    return dw


def conv_wgrad_folded(xf: Act, dyf: Act, dw: torch.Tensor, kf, pf, f, ldo, ldi):
    """dw (fp32 reference layout (Co, Ci, kT, kH, kW)) += wgrad of the pixel-folded conv (xf / dyf = folded views)"""
    assert dw.dtype == torch.float32 and dw.is_contiguous()
    g = make_geom(xf, dyf, kf, (1, 1, 1), pf, False)
    check(lib().sfb_conv_wgrad_folded(C.byref(g), xf.ptr(), dyf.ptr(), dyf.ld, dyf.coff, dw.data_ptr(), dw.shape[0],
                                      dw.shape[1], dw.shape[4], f, ldo, ldi, stream_ptr()), "sfb_conv_wgrad_folded")
    return dw


def bn_finalize(stats, count, gamma, beta, eps, momentum, running_mean, running_var, scale, shift, mean, invstd):
    check(lib().sfb_bn_finalize(stats.data_ptr(), gamma.numel(), count, gamma.data_ptr(), beta.data_ptr(), eps,
                                momentum, ptr(running_mean), ptr(running_var), scale.data_ptr(), shift.data_ptr(),
                                mean.data_ptr(), invstd.data_ptr(), stream_ptr()), "sfb_bn_finalize")


def bn_fold(gamma, beta, rm, rv, eps, scale, shift):
    check(lib().sfb_bn_fold(gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(), eps, gamma.numel(),
                            scale.data_ptr(), shift.data_ptr(), stream_ptr()), "sfb_bn_fold")


def bn_add_relu(y: Act, scale, shift, r: Optional[Act], r_scale, r_shift, out: Act, relu=True, mask_bits=None):
    """mask_bits: optional uint8 tensor [M, C/8] receiving the packed ReLU mask of `out` (bit i = channel 8*cv+i)"""
    check(lib().sfb_bn_add_relu(y.ptr(), y.ld, y.coff, scale.data_ptr(), shift.data_ptr(),
                                r.ptr() if r else 0, r.ld if r else 0, r.coff if r else 0, ptr(r_scale), ptr(r_shift),
                                out.ptr(), out.ld, out.coff, ptr(mask_bits), y.M, y.C, int(relu), stream_ptr()),
          "sfb_bn_add_relu")
    return out


def bn_raw(stats, count, gamma, beta, eps, momentum, running_mean, running_var, scale, shift, save_mean, save_invstd,
           fold_f=0, fold_ld=0):
    """SfbBnRaw descriptor (keep the returned object alive until the launch has been enqueued)"""
    return _ffi.BnRaw(stats.data_ptr(), int(count), gamma.data_ptr(), beta.data_ptr(), float(eps), float(momentum),
                      ptr(running_mean), ptr(running_var), scale.data_ptr(), shift.data_ptr(), save_mean.data_ptr(),
                      save_invstd.data_ptr(), int(fold_f), int(fold_ld))


def bn_add_relu_raw(y: Act, y_bn, r: Optional[Act], r_bn, out: Act, relu=True, mask_bits=None):
    """sfb_bn_add_relu with the BatchNorm finalize fused in (y_bn / r_bn: bn_raw(...) descriptors, r_bn may be None)"""
    check(lib().sfb_bn_add_relu_raw(y.ptr(), y.ld, y.coff, C.byref(y_bn), r.ptr() if r else 0, r.ld if r else 0,
                                    r.coff if r else 0, C.byref(r_bn) if r_bn is not None else None, out.ptr(), out.ld,
                                    out.coff, ptr(mask_bits), y.M, y.C, int(relu), stream_ptr()), "sfb_bn_add_relu_raw")
    return out


def bn_bwd_reduce(dz: Act, y: Act, mask: Optional[Act], scale, shift, mean, invstd, relu, dgamma, dbeta, mask_bits=None):
    check(lib().sfb_bn_bwd_reduce(dz.ptr(), dz.ld, dz.coff, y.ptr(), y.ld, y.coff,
                                  mask.ptr() if mask else 0, mask.ld if mask else 0, mask.coff if mask else 0,
                                  ptr(mask_bits), ptr(scale), ptr(shift), mean.data_ptr(), invstd.data_ptr(), int(relu), y.M, y.C,
                                  dgamma.data_ptr(), dbeta.data_ptr(), stream_ptr()), "sfb_bn_bwd_reduce")


def bn_bwd_apply(dz: Act, y: Act, mask: Optional[Act], scale, shift, mean, invstd, gamma, dgamma, dbeta, relu,
                 dy: Act, g_out: Optional[Act] = None, g_accumulate=False, mask_bits=None):
    check(lib().sfb_bn_bwd_apply(dz.ptr(), dz.ld, dz.coff, y.ptr(), y.ld, y.coff,
                                 mask.ptr() if mask else 0, mask.ld if mask else 0, mask.coff if mask else 0,
                                 ptr(mask_bits), ptr(scale), ptr(shift), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(),
                                 dgamma.data_ptr(), dbeta.data_ptr(), int(relu), y.M, y.C, dy.ptr(), dy.ld, dy.coff,
                                 g_out.ptr() if g_out else 0, g_out.ld if g_out else 0, g_out.coff if g_out else 0,
                                 int(g_accumulate), stream_ptr()), "sfb_bn_bwd_apply")


NOALLOC_MAX_C = 2048  # SFB_NOALLOC_MAX_C in include/sfb200.h


def bn_bwd_noalloc(dz: Act, y: Act, scale, shift, mean, invstd, gamma, relu, dy: Act, dgamma_out=None, dbeta_out=None):
    """sfb_bn_bwd_noalloc: BatchNorm(+ReLU) backward whose per-channel sums live in the library's static device array
    (no allocation anywhere in the call; C <= NOALLOC_MAX_C; calls that overlap on one GPU race -- include/sfb200.h)."""
    check(lib().sfb_bn_bwd_noalloc(dz.ptr(), dz.ld, dz.coff, y.ptr(), y.ld, y.coff, ptr(scale), ptr(shift),
                                   mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(), int(relu), y.M, y.C,
                                   dy.ptr(), dy.ld, dy.coff, ptr(dgamma_out), ptr(dbeta_out), stream_ptr()),
          "sfb_bn_bwd_noalloc")
    return dy


def maxpool_fwd(y: Act, scale, shift, out: Act, idx: Optional[torch.Tensor]):
    N, T, H, W = y.dims
    check(lib().sfb_maxpool_fwd(y.ptr(), y.ld, y.coff, ptr(scale), ptr(shift), N, T, H, W, y.C, out.ptr(), out.ld,
                                out.coff, ptr(idx), stream_ptr()), "sfb_maxpool_fwd")
    return out


def maxpool_bwd(dout: Act, idx: torch.Tensor, dz: Act):
    N, T, H, W = dz.dims
    check(lib().sfb_maxpool_bwd(dout.ptr(), dout.ld, dout.coff, idx.data_ptr(), N, T, H, W, dz.C, dz.ptr(), dz.ld,
                                dz.coff, stream_ptr()), "sfb_maxpool_bwd")
    return dz


def bn_bwd_tail2(dz: Act, mask_bits: torch.Tensor, sides):
    """sides = two tuples (y, mean, invstd, gamma, dgamma, dbeta, dy): both BatchNorms of a projection-shortcut tail in one
    reduction + one apply pass (sfb_bn_bwd_tail2)"""
    st = [_ffi.BnTailSide(y.ptr(), y.ld, y.coff, ptr(mean), ptr(invstd), ptr(gamma), ptr(dgamma), ptr(dbeta), dy.ptr(), dy.ld,
                          dy.coff) for (y, mean, invstd, gamma, dgamma, dbeta, dy) in sides]
    check(lib().sfb_bn_bwd_tail2(dz.ptr(), dz.ld, dz.coff, mask_bits.data_ptr(), C.byref(st[0]), C.byref(st[1]), dz.M, dz.C,
                                 stream_ptr()), "sfb_bn_bwd_tail2")


def head_pool_fwd(x: Act, k, dropout_p, seed, feat: torch.Tensor, f_off: int, seed_dev: torch.Tensor = None):
    N, T, H, W = x.dims
    check(lib().sfb_head_pool_fwd(x.ptr(), x.ld, x.coff, N, T, H, W, x.C, k[0], k[1], k[2], float(dropout_p), seed,
                                  seed_dev.data_ptr() if seed_dev is not None else None,
                                  feat.data_ptr(), feat.shape[1], f_off, stream_ptr()), "sfb_head_pool_fwd")


def head_pool_bwd(dfeat: torch.Tensor, f_off: int, k, dropout_p, seed, dx: Act, seed_dev: torch.Tensor = None):
    N, T, H, W = dx.dims
    check(lib().sfb_head_pool_bwd(dfeat.data_ptr(), dfeat.shape[1], f_off, N, T, H, W, dx.C, k[0], k[1], k[2],
                                  float(dropout_p), seed, seed_dev.data_ptr() if seed_dev is not None else None,
                                  dx.ptr(), dx.ld, dx.coff, stream_ptr()), "sfb_head_pool_bwd")


def fc_fwd(feat, w, b, logits):
    N, K = feat.shape
    check(lib().sfb_fc_fwd(feat.data_ptr(), w.data_ptr(), ptr(b), N, K, w.shape[0], logits.data_ptr(), stream_ptr()),
          "sfb_fc_fwd")
    return logits


def fc_bwd(dlogits, feat, w, dfeat, dw, db):
    N, K = feat.shape
    check(lib().sfb_fc_bwd(dlogits.data_ptr(), feat.data_ptr(), w.data_ptr(), N, K, w.shape[0], dfeat.data_ptr(),
                           dw.data_ptr(), ptr(db), stream_ptr()), "sfb_fc_bwd")
