"""ctypes binding of libsfb200.so (C ABI declared in include/sfb200.h).

The product path has no fallback: if the shared library is missing or a call fails, an
exception is raised (SfbError).  torch is used only to own device memory and streams.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libsfb200.so")

IMPL_TC = 0
IMPL_SIMT = 1
IMPL_TC_GATHER = 2


class SfbError(RuntimeError):
    pass


class ConvGeom(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "N", "Ti", "Hi", "Wi", "Cg", "To", "Ho", "Wo", "Co", "kT", "kH", "kW", "sT", "sH", "sW",
        "pT", "pH", "pW", "g_ld", "g_coff", "o_ld", "o_coff", "transposed")]


class Prologue(C.Structure):
    _fields_ = [("scale", C.c_void_p), ("shift", C.c_void_p), ("relu", C.c_int32)]


class Epilogue(C.Structure):
    _fields_ = [("scale", C.c_void_p), ("shift", C.c_void_p), ("residual", C.c_void_p),
                ("res_ld", C.c_int32), ("res_coff", C.c_int32), ("relu", C.c_int32), ("stats", C.c_void_p)]


class BnBwdFuse(C.Structure):
    """SfbBnBwdFuse (include/sfb200.h): BatchNorm(+ReLU) backward reduction fused into the dgrad epilogue"""
    _fields_ = [("y", C.c_void_p), ("y_ld", C.c_int32), ("y_coff", C.c_int32), ("scale", C.c_void_p),
                ("shift", C.c_void_p), ("mean", C.c_void_p), ("invstd", C.c_void_p), ("dgamma", C.c_void_p),
                ("dbeta", C.c_void_p), ("C", C.c_int32), ("fold_ld", C.c_int32)]


class BnTailSide(C.Structure):
    """SfbBnTailSide (include/sfb200.h): one of the two BatchNorms of a projection-shortcut block tail"""
    _fields_ = [("y", C.c_void_p), ("y_ld", C.c_int32), ("y_coff", C.c_int32), ("mean", C.c_void_p), ("invstd", C.c_void_p),
                ("gamma", C.c_void_p), ("dgamma", C.c_void_p), ("dbeta", C.c_void_p), ("dy", C.c_void_p),
                ("dy_ld", C.c_int32), ("dy_coff", C.c_int32)]


_i32, _i64, _u64, _f32, _vp = C.c_int32, C.c_int64, C.c_uint64, C.c_float, C.c_void_p


class PackJob(C.Structure):
    """SfbPackJob (include/sfb200.h): one weight-operand packing job of a batched table"""
    _fields_ = [(n, C.c_int32) for n in (
        "kind", "Co", "Ci", "kT", "kH", "kW", "sT", "sH", "sW", "Cg", "transposed", "f", "ldo", "ldi", "rows_pad",
        "k_pad")] + [("src", C.c_void_p), ("dst", C.c_void_p)]


class AdamJob(C.Structure):
    """SfbAdamJob (include/sfb200.h): one parameter tensor of the batched optimizer step"""
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("n", C.c_int64)]


PACK_WEIGHT, PACK_WEIGHT_DGRAD, PACK_FOLD, PACK_STEM = 0, 1, 2, 3


class BnRaw(C.Structure):
    """SfbBnRaw (include/sfb200.h): raw BatchNorm sums + where the finalised coefficients go"""
    _fields_ = [("stats", C.c_void_p), ("count", C.c_int64), ("gamma", C.c_void_p), ("beta", C.c_void_p),
                ("eps", C.c_float), ("momentum", C.c_float), ("running_mean", C.c_void_p), ("running_var", C.c_void_p),
                ("scale", C.c_void_p), ("shift", C.c_void_p), ("save_mean", C.c_void_p), ("save_invstd", C.c_void_p),
                ("fold_f", C.c_int32), ("fold_ld", C.c_int32)]

class StagePathway(C.Structure):
    """SfbStagePathway (include/sfb200.h): one destination of sfb_stage_clips_u8"""
    _fields_ = [("c0", C.c_int32), ("C", C.c_int32), ("t_stride", C.c_int32), ("Cs", C.c_int32), ("out", C.c_void_p)]


# name -> argtypes; every function returns int (status) unless listed in _SPECIAL
SIGNATURES = {
    "sfb_packed_weight_dims": [_i32, _i32, _i32, C.POINTER(_i32), C.POINTER(_i32)],
    "sfb_pack_weight": [_vp, _i32, _i32, _i32, _i32, _i32, _vp, _i32, _i32, _vp],
    "sfb_layout_input": [_vp, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _i32, _vp, _i32, _vp],
    "sfb_conv_gemm": [C.POINTER(ConvGeom), _vp, C.POINTER(Prologue), _vp, _i32, _i32, _vp, C.POINTER(Epilogue), _i32, _vp],
    "sfb_conv_dgrad_bn": [C.POINTER(ConvGeom), _vp, _vp, _i32, _i32, _vp, C.POINTER(BnBwdFuse), _vp],
    "sfb_conv_wgrad": [C.POINTER(ConvGeom), _vp, C.POINTER(Prologue), _vp, _i32, _i32, _vp, _i32, _i32, _vp],
    "sfb_bn_finalize": [_vp, _i32, _i64, _vp, _vp, _f32, _f32, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "sfb_bn_fold": [_vp, _vp, _vp, _vp, _f32, _i32, _vp, _vp, _vp],
    "sfb_bn_add_relu": [_vp, _i32, _i32, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _i32, _i32, _vp, _i64, _i32, _i32, _vp],
    "sfb_bn_bwd_reduce": [_vp, _i32, _i32, _vp, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _i32, _vp, _vp, _vp],
    "sfb_bn_bwd_apply": [_vp, _i32, _i32, _vp, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                         _i32, _i64, _i32, _vp, _i32, _i32, _vp, _i32, _i32, _i32, _vp],
    "sfb_bn_bwd_noalloc": [_vp, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _i32, _vp, _i32, _i32,
                           _vp, _vp, _vp],
    "sfb_maxpool_fwd": [_vp, _i32, _i32, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _i32, _i32, _vp, _vp],
    "sfb_maxpool_bwd": [_vp, _i32, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _i32, _i32, _vp],
    "sfb_bn_bwd_tail2": [_vp, _i32, _i32, _vp, C.POINTER(BnTailSide), C.POINTER(BnTailSide), _i64, _i32, _vp],
    "sfb_head_pool_fwd": [_vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _f32, _u64, _vp, _vp, _i32, _i32, _vp],
    "sfb_head_pool_bwd": [_vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _f32, _u64, _vp, _vp, _i32, _i32, _vp],
    "sfb_fc_fwd": [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp],
    "sfb_fc_bwd": [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp],
    "sfb_packed_weight_dims_dgrad": [_i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, C.POINTER(_i32), C.POINTER(_i32)],
    "sfb_pack_weight_dgrad": [_vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _i32, _i32, _vp],
    "sfb_fold_weight": [_vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp],
    "sfb_unfold_wgrad": [_vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp],
    "sfb_unfold_stats": [_vp, _i32, _i32, _i32, _vp, _vp],
    "sfb_layout_input_s2d": [_vp, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _i32, _vp, _i32, _vp],
    "sfb_stage_clips_u8": [_vp, _i32, _i32, _i32, _i32, _i32, _vp, C.c_float, C.c_float, C.POINTER(StagePathway), _i32, _vp],
    "sfb_video_scores": [_vp, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp],
    "sfb_sparse_fuse": [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp],
    "sfb_sparse_fuse_bwd": [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp],
    "sfb_pack_weight_stem": [_vp, _i32, _i32, _i32, _vp, _i32, _vp],
    "sfb_stem_conv_fprop": [_vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp],
    "sfb_stem_conv_wgrad": [_vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _vp],
    "sfb_bn_add_relu_raw": [_vp, _i32, _i32, C.POINTER(BnRaw), _vp, _i32, _i32, C.POINTER(BnRaw), _vp, _i32, _i32, _vp, _i64,
                            _i32, _i32, _vp],
    "sfb_conv_wgrad_folded": [C.POINTER(ConvGeom), _vp, _vp, _i32, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp],
    "sfb_pack_plan": [C.POINTER(PackJob), _i32, _vp, _i64, C.POINTER(_i32)],
    "sfb_pack_run": [_vp, _i32, _i32, _vp],
    "sfb_adam_plan": [C.POINTER(AdamJob), _i32, _vp, _i64, C.POINTER(_i32)],
    "sfb_adam_step": [_vp, _i32, _i32, _i64, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, _vp],
}
_SPECIAL = {
    "sfb_strerror": ([_i32], C.c_char_p),
    "sfb_last_cuda_error": ([], C.c_char_p),
    "sfb_version": ([], _i32),
    "sfb_launch_count": ([], _i64),
    "sfb_launch_count_add": ([_i64], _i64),
    "sfb_stem_s2d_channels": ([_i32], _i32),
    "sfb_pack_table_bytes": ([_i32], _i64),
    "sfb_adam_table_bytes": ([_i32], _i64),
}
ALL_SYMBOLS = sorted(list(SIGNATURES) + list(_SPECIAL))

_lib = None


def lib():
    """Load libsfb200.so (once).  Raises SfbError if it has not been built -- there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SfbError(f"{LIB_PATH} not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(the sm_100a CUDA extension is mandatory; there is no CPU/PyTorch fallback)")
        l = C.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(l, name)
            fn.argtypes = args
            fn.restype = C.c_int
        for name, (args, res) in _SPECIAL.items():
            fn = getattr(l, name)
            fn.argtypes = args
            fn.restype = res
        _lib = l
    return _lib


def check(status, what=""):
    if status != 0:
        l = lib()
        msg = l.sfb_strerror(status).decode()
        extra = l.sfb_last_cuda_error().decode()
        raise SfbError(f"{what}: {msg} [{status}] {extra}")


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream


def packed_weight_dims(Co, Cg, taps):
    r, k = _i32(), _i32()
    check(lib().sfb_packed_weight_dims(Co, Cg, taps, C.byref(r), C.byref(k)), "sfb_packed_weight_dims")
    return r.value, k.value


def packed_weight_dims_dgrad(Ci, Cg, k, s):
    r, kp = _i32(), _i32()
    check(lib().sfb_packed_weight_dims_dgrad(Ci, Cg, k[0], k[1], k[2], s[0], s[1], s[2], C.byref(r), C.byref(kp)),
          "sfb_packed_weight_dims_dgrad")
    return r.value, kp.value


def launch_count():
    return int(lib().sfb_launch_count())


def ptr(t):
    return 0 if t is None else t.data_ptr()
