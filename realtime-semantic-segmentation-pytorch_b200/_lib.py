"""ctypes binding of librtseg.so — the only door between Python and the kernels.

Every prototype below mirrors include/rtseg.h one to one.  The loader fails
loudly: there is no Python/PyTorch re-implementation behind these calls, so a
missing or stale library is an error, never a silent fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

c_i32, c_i64, c_f32, c_vp, c_sz, c_u64 = C.c_int32, C.c_int64, C.c_float, C.c_void_p, C.c_size_t, C.c_uint64

F32, BF16 = 0, 1

ACT_CODES = {
    "none": 0, "relu": 1, "relu6": 2, "leakyrelu": 3, "prelu": 4, "elu": 5, "celu": 6, "selu": 7,
    "hardswish": 8, "hardtanh": 9, "gelu": 10, "silu": 11, "sigmoid": 12, "tanh": 13,
}
RES_NONE, RES_PRE, RES_POST = 0, 1, 2


class ConvDesc(C.Structure):
    """struct rtseg_conv_desc (include/rtseg.h)."""

    _fields_ = [
        ("n", c_i32), ("h", c_i32), ("w", c_i32),
        ("cin", c_i32), ("cout", c_i32),
        ("kh", c_i32), ("kw", c_i32),
        ("stride", c_i32), ("dil", c_i32),
        ("pad_h", c_i32), ("pad_w", c_i32),
        ("groups", c_i32), ("dtype", c_i32),
        ("act", c_i32), ("act_p0", c_f32), ("act_p1", c_f32),
        ("res_mode", c_i32), ("act2", c_i32),
        ("x_ctotal", c_i32), ("x_coff", c_i32),
        ("y_ctotal", c_i32), ("y_coff", c_i32),
        ("r_ctotal", c_i32), ("r_coff", c_i32),
    ]


_DP = C.POINTER(ConvDesc)

# name -> (restype, argtypes); keep in the order of include/rtseg.h
PROTOTYPES = {
    "rtseg_version": (c_i32, []),
    "rtseg_last_error": (C.c_char_p, []),
    "rtseg_device_check": (c_i32, []),
    "rtseg_set_device": (c_i32, [c_i32]),
    "rtseg_launch_count": (c_u64, []),
    "rtseg_out_hw": (c_i32, [_DP, C.POINTER(c_i32), C.POINTER(c_i32)]),
    "rtseg_conv_dw": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_conv_dense": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_conv_tc_supported": (c_i32, [_DP]),
    "rtseg_conv_tc": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_conv_pair_tc_supported": (c_i32, [_DP, c_i32]),
    "rtseg_conv_pair_tc": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "rtseg_tc_timeout_flag": (c_i32, [c_vp]),
    "rtseg_conv_tc_plan_info": (c_i32, [_DP, C.POINTER(c_i32)]),
    "rtseg_dw_tma_launch_count": (c_u64, []),
    "rtseg_stem": (c_i32, [_DP, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_stem_u8": (c_i32, [_DP, c_i32, c_vp, C.POINTER(c_f32), C.POINTER(c_f32), c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_maxpool3x3s2": (c_i32, [_DP, c_vp, c_vp, c_vp]),
    "rtseg_maxpool2x2_idx": (c_i32, [c_i32] * 5 + [c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_vp]),
    "rtseg_maxunpool2x2": (c_i32, [c_i32] * 7 + [c_vp, c_i32, c_i32, c_vp, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_deconv3x3s2": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_deconv3x3s2_tc_supported": (c_i32, [_DP]),
    "rtseg_deconv3x3s2_tc": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_upsample_bilinear": (c_i32, [c_i32] * 8 + [c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_upsample_argmax": (c_i32, [c_i32] * 9 + [c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "rtseg_argmax_nhwc": (c_i32, [c_i64, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_argmax_nchw": (c_i32, [c_i32, c_i32, c_i64, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_resize_nearest_i64": (c_i32, [c_i32] * 5 + [c_vp, c_vp, c_vp]),
    "rtseg_confmat_update": (c_i32, [c_i64, c_i32, c_i32, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_ohem_workspace_bytes": (c_sz, []),
    "rtseg_ohem_fwd": (c_i32, [c_i32] * 6 + [c_vp, c_vp, c_i32, c_f32, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_ohem_bwd": (c_i32, [c_i32] * 6 + [c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_ohem_fwd_static_max_pixels": (c_i64, []),
    "rtseg_ohem_fwd_static": (c_i32, [c_i32] * 6 + [c_vp, c_vp, c_i32, c_f32, c_vp, c_vp]),
    "rtseg_nchw_to_nhwc": (c_i32, [c_i32] * 5 + [c_vp, c_i32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_nhwc_to_nchw": (c_i32, [c_i32] * 5 + [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "rtseg_adaptive_avgpool": (c_i32, [c_i32] * 6 + [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "rtseg_avgpool3x3s2": (c_i32, [c_i32] * 6 + [c_vp, c_i32, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_avgpool3x3s2_u8": (c_i32, [c_i32] * 4 + [c_vp, C.POINTER(c_f32), C.POINTER(c_f32), c_i32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_channel_shuffle": (c_i32, [c_i64, c_i32, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_bn_stats": (c_i32, [c_i64, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp]),
    "rtseg_affine_act": (c_i32, [c_i64, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_vp, c_i32, c_f32, c_f32,
                                 c_vp, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_bn_act_bwd_reduce": (c_i32, [c_i64, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32,
                                        c_vp, c_vp, c_vp, c_vp, c_i32, c_f32, c_f32, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_bn_act_bwd_apply": (c_i32, [c_i64, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32,
                                       c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_f32, c_f32, c_vp, c_vp, c_i32, c_i32,
                                       c_vp]),
    "rtseg_bn_finalize": (c_i32, [c_i32, c_vp, c_vp, C.c_double, c_vp, c_vp, C.c_double, C.c_double, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_bn_bwd_finalize": (c_i32, [c_i32, c_vp, c_vp, c_vp, C.c_double, c_vp, c_vp]),
    "rtseg_conv_wgrad": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_conv_wgrad_tc_supported": (c_i32, [_DP]),
    "rtseg_conv_wgrad_tc": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_conv_dw_bwd_data": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_conv_dw_wgrad": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_prelu_slope_grad": (c_i32, [c_i64, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_vp,
                                       c_vp, c_vp]),
    "rtseg_upsample_bilinear_bwd": (c_i32, [c_i32] * 8 + [c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_adaptive_avgpool_bwd": (c_i32, [c_i32] * 6 + [c_vp, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_combine_up": (c_i32, [c_i32] * 5 + [c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_f32,
                                 c_f32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_avgpool2d": (c_i32, [c_i32] * 8 + [c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_dwpw_tc_supported": (c_i32, [_DP, _DP]),
    "rtseg_dwpw_tc": (c_i32, [_DP, c_vp, c_vp, c_vp, c_vp, _DP, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_maxunpool2x2_bwd": (c_i32, [c_i32] * 7 + [c_vp, c_vp, c_vp, c_vp]),
    "rtseg_pyramid_avgpool": (c_i32, [c_i32] * 5 + [c_vp, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rtseg_ppm_branches": (c_i32, [c_i32] * 6 + [c_vp] * 7 + [c_i32, c_i32, c_vp]),
    "rtseg_opt_chunk_elems": (c_i32, []),
    "rtseg_sgd_ema_step": (c_i32, [c_vp, c_i32, c_vp, c_i32, c_vp]),
    "rtseg_maxpool3x3s2_bwd": (c_i32, [c_i32] * 5 + [c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp]),
    "rtseg_colorize_blend": (c_i32, [c_i64, c_vp, c_vp, c_vp, c_f32, c_vp, c_vp]),
    "rtseg_upsample_argmax_host_scratch_bytes": (c_sz, [c_i32] * 7),
    "rtseg_upsample_argmax_host": (c_i32, [c_i32] * 8 + [c_vp, c_vp, c_vp, c_vp]),
}

_lock = threading.Lock()
_lib = None


class RtsegError(RuntimeError):
    pass


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load librtseg.so (building it with nvcc first if it is absent or stale)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = os.environ.get("RTSEG_LIB_PATH") or _build.LIB_PATH   # RTSEG_LIB_PATH: A/B runs of two builds
        if path == _build.LIB_PATH and build_if_missing and _build.sources_newer_than_lib():
            try:
                _build.build_library()
            except Exception as e:  # stale-but-present library is still usable
                if not os.path.exists(path):
                    raise RtsegError(f"librtseg.so is missing and could not be built: {e}") from e
        if not os.path.exists(path):
            raise RtsegError(f"{path} not found: run `python -c 'import __graft_entry__ as g; g.build()'`")
        lib = C.CDLL(path)
        for name, (res, args) in PROTOTYPES.items():
            try:
                fn = getattr(lib, name)
            except AttributeError as e:
                raise RtsegError(f"librtseg.so does not export {name}; rebuild it") from e
            fn.restype = res
            fn.argtypes = args
        if lib.rtseg_version() != 100:
            raise RtsegError(f"librtseg.so version {lib.rtseg_version()} != binding version 100")
        _lib = lib
    return _lib


def last_error() -> str:
    return load().rtseg_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = "") -> None:
    """Raise on a non-zero rtseg_status (mirrors the reference raising from torch ops)."""
    if rc != 0:
        msg = last_error()
        if rc == -2:
            raise NotImplementedError(f"{what}: {msg}")
        raise RtsegError(f"{what} failed (status {rc}): {msg}")


_tls = threading.local()


def sync_device(index: int) -> None:
    """Tell the library's CUDA runtime which device torch made current (cheap when unchanged)."""
    if getattr(_tls, "dev", None) != index:
        check(load().rtseg_set_device(index), "rtseg_set_device")
        _tls.dev = index


def launch_count() -> int:
    return int(load().rtseg_launch_count())
