"""ctypes binding of libraftcorr.so (include/raft_corr.h).  No torch types cross
this boundary: only raw device pointers, sizes and the stream handle."""
from __future__ import annotations

import ctypes as C
import os

RC_MAX_LEVELS = 6
RC_VOL_F32, RC_VOL_BF16, RC_VOL_F16 = 0, 1, 2
RC_MATH_F32, RC_MATH_BF16, RC_MATH_TF32, RC_MATH_F16 = 0, 1, 2, 3
RC_ABI_VERSION = 6
RC_STATIC_SCRATCH_BYTES = 16 << 20

_HERE = os.path.dirname(os.path.abspath(__file__))
# RAFTCORR_LIB: another build of the same library (tools/power_probe.py loads the attribution build this way)
LIB_PATH = os.environ.get("RAFTCORR_LIB") or os.path.join(_HERE, "libraftcorr.so")


class Layout(C.Structure):
    """mirror of struct rc_layout"""
    _fields_ = [
        ("num_levels", C.c_int32), ("B", C.c_int32), ("N", C.c_int32),
        ("tile_h", C.c_int32), ("tile_w", C.c_int32),
        ("H", C.c_int32 * RC_MAX_LEVELS), ("W", C.c_int32 * RC_MAX_LEVELS),
        ("tiles_y", C.c_int32 * RC_MAX_LEVELS), ("tiles_x", C.c_int32 * RC_MAX_LEVELS),
        ("slab", C.c_int64 * RC_MAX_LEVELS), ("offset", C.c_int64 * RC_MAX_LEVELS),
        ("total_elems", C.c_int64), ("elem_bytes", C.c_int32),
    ]


# name -> (restype, argtypes); every symbol include/raft_corr.h declares
_p, _i, _sz, _i64 = C.c_void_p, C.c_int, C.c_size_t, C.c_int64
SIGNATURES = {
    "rc_abi_version": (_i, []),
    "rc_last_error": (C.c_char_p, []),
    "rc_pyramid_layout": (_i, [_i, _i, _i, _i, _i, C.POINTER(Layout)]),
    "rc_build_workspace_bytes": (_i64, [_i, _i, _i, _i, _i]),
    "rc_build_pyramid": (_i, [_p, _p, _i, _i, _i, _i, _i, _i, _i, _p, _p, _sz, _p]),
    "rc_build_pyramid_static": (_i, [_p, _p, _i, _i, _i, _i, _i, _i, _p, _p]),
    "rc_pack_features_nhwc16": (_i, [_p, _p, _i, _i, _i, _i, _i, _p, _p, _p]),
    "rc_build_pyramid_nhwc16": (_i, [_p, _p, _i, _i, _i, _i, _i, _i, _i, _p, _p]),
    "rc_lookup_fwd": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _p, _p]),
    "rc_convc1_weight_bytes": (_i64, [_i, _i]),
    "rc_convc1_pack_weights": (_i, [_p, _i, _i, _p, _p]),
    "rc_lookup_convc1_workspace_bytes": (_i64, [_i, _i, _i, _i, _i]),
    "rc_lookup_convc1_fwd": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _p, _p, _i, _p, _i, _p, _sz, _p]),
    "rc_lookup_bwd": (_i, [_p, _i, _p, _p, _i, _i, _i, _i, _i, _p, _p, _p]),
    "rc_build_bwd_workspace_bytes": (_i64, [_i, _i, _i, _i, _i]),
    "rc_build_bwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "rc_alt_fwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _p, _p]),
    "rc_alt_bwd": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p]),
    "rc_alt_feature_pyramid_elems": (_i64, [_i, _i, _i, _i, _i, C.POINTER(_i64)]),
    "rc_alt_pool_features": (_i, [_p, _p, _i, _i, _i, _i, _i, _p, _p, _p]),
    "rc_alt_lookup_fwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p]),
    "rc_alt_features_16": (_i, [_p, _p, _i, _i, _i, _i, _i, _i, _p, _p, _p]),
    "rc_alt_lookup_fwd_tc": (_i, [_p, _p, _i, _p, _i, _i, _i, _i, _i, _i, _p, _p]),
    "rc_lookup_fwd_f16": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _p, _p]),
    "rc_upsample_flow": (_i, [_p, _p, _i, _i, _i, _p, _p]),
    "rc_upsample_flow_f16mask": (_i, [_p, _p, _i, _i, _i, _p, _p]),
    "rc_upsample_flow_bwd": (_i, [_p, _p, _p, _i, _i, _i, _p, _p, _p]),
    "rc_upflow8": (_i, [_p, _i, _i, _i, _i, _p, _p]),
    "rc_upflow8_bwd": (_i, [_p, _i, _i, _i, _i, _p, _p]),
}

_lib = None


class RaftCorrError(RuntimeError):
    """Raised for every non-zero return of the C ABI (the reference raises
    RuntimeError from TORCH_CHECK, alt_cuda_corr/correlation.cpp:19-21)."""


def lib() -> C.CDLL:
    """Load libraftcorr.so once.  Missing library is a hard error, never a fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RaftCorrError(
                f"{LIB_PATH} is missing: build it with `python -m raft_stir_b200.build` "
                "(there is no CPU or PyTorch fallback for the correlation path)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)   # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        got = handle.rc_abi_version()
        if got != RC_ABI_VERSION:
            raise RaftCorrError(f"libraftcorr ABI {got}, binding expects {RC_ABI_VERSION}")
        _lib = handle
    return _lib


def check(code: int, what: str) -> None:
    if code != 0:
        msg = lib().rc_last_error().decode("utf-8", "replace")
        raise RaftCorrError(f"{what} failed ({code}): {msg}")


def pyramid_layout(B: int, H: int, W: int, num_levels: int, vol_dtype: int) -> Layout:
    lay = Layout()
    check(lib().rc_pyramid_layout(B, H, W, num_levels, vol_dtype, C.byref(lay)), "rc_pyramid_layout")
    return lay


def _tiled(buf, lay: Layout, level: int):
    """5-D strided VIEW [B*N, tiles_y, tile_h, tiles_x, tile_w] of one level."""
    import torch
    th, tw = int(lay.tile_h), int(lay.tile_w)
    ty, tx = int(lay.tiles_y[level]), int(lay.tiles_x[level])
    q = int(lay.B) * int(lay.N)
    return torch.as_strided(buf, (q, ty, th, tx, tw), (int(lay.slab[level]), tx * th * tw, tw, th * tw, 1),
                            int(lay.offset[level]))


def level_store(buf, lay: Layout, level: int, values) -> None:
    """Inverse of level_view: write logical [B*N, H_l, W_l] values into the tiled
    buffer, zeros into the padding (the invariant the lookup relies on)."""
    import torch
    v = _tiled(buf, lay, level)
    q, ty, th, tx, tw = v.shape
    full = torch.zeros(q, ty * th, tx * tw, dtype=buf.dtype, device=buf.device)
    full[:, : int(lay.H[level]), : int(lay.W[level])] = values.to(device=buf.device, dtype=buf.dtype)
    v.copy_(full.view(q, ty, th, tx, tw))


def level_view(buf, lay: Layout, level: int):
    """Logical [B*N, H_l, W_l] copy of one pyramid level out of the tiled buffer
    (`buf`: flat torch tensor holding the whole pyramid).  The buffer keeps every
    query's map in 64-byte tiles (see rc_layout in include/raft_corr.h), which no
    strided view can express, so this gathers; it is meant for tests, debugging
    and the `corr_pyramid` attribute, not for the hot path."""
    v = _tiled(buf, lay, level)
    q, ty, th, tx, tw = v.shape
    return v.reshape(q, ty * th, tx * tw)[:, : int(lay.H[level]), : int(lay.W[level])]
