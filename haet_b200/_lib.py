"""ctypes binding of libhaet_b200.so (the C-ABI declared in include/haet_b200.h).

The library is the product: there is no Python/CPU fallback.  If the shared object is missing
or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HAET_B200_LIB") or os.path.join(HERE, "libhaet_b200.so")   # override: development builds (scripts/trace_pass2.py)
MAX_LEVELS = 8

HAET_OK, HAET_EINVAL, HAET_ECUDA, HAET_ENODEV, HAET_ENOSUP = 0, -1, -2, -3, -4


class LayerShape(C.Structure):
    _fields_ = [("B", C.c_int64), ("N", C.c_int64), ("dim", C.c_int32), ("heads", C.c_int32),
                ("dim_head", C.c_int32), ("slice_num", C.c_int32), ("conv_nd", C.c_int32),
                ("grid", C.c_int32 * 3), ("kernel", C.c_int32), ("dimensionality", C.c_int32),
                ("base_temp", C.c_float), ("epsilon", C.c_float)]


class ErwinCfg(C.Structure):
    _fields_ = [("n_levels", C.c_int32),
                ("c_hidden", C.c_int32 * MAX_LEVELS), ("ball_sizes", C.c_int32 * MAX_LEVELS),
                ("enc_num_heads", C.c_int32 * MAX_LEVELS), ("enc_depths", C.c_int32 * MAX_LEVELS),
                ("dec_num_heads", C.c_int32 * MAX_LEVELS), ("dec_depths", C.c_int32 * MAX_LEVELS),
                ("strides", C.c_int32 * MAX_LEVELS),
                ("rotate", C.c_float), ("decode", C.c_int32), ("mlp_ratio", C.c_int32),
                ("dimensionality", C.c_int32), ("attn_bias", C.c_int32),
                ("bn_train", C.c_int32), ("bn_eps", C.c_float), ("bn_stats_d", C.c_void_p)]


class LayerWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("wx_d", "bx_d", "ws_d", "bs_d", "wt_d", "wo_d", "bo_d", "erwin_d")] + \
               [("bt", C.c_float)]


# name -> (restype, argtypes); every symbol of include/haet_b200.h
_P = C.c_void_p
_SH = C.POINTER(LayerShape)
_EC = C.POINTER(ErwinCfg)
SYMBOLS = {
    "haet_version": (C.c_int, []),
    "haet_last_error": (C.c_char_p, []),
    "haet_device_check": (C.c_int, []),
    "haet_spill_bytes": (C.c_int64, [_SH]),
    "haet_slice_workspace_bytes": (C.c_int64, [_SH]),
    "haet_deslice_workspace_bytes": (C.c_int64, [_SH]),
    "haet_partial_floats": (C.c_int64, [_SH]),
    "haet_erwin_weight_floats": (C.c_int64, [_EC]),
    "haet_erwin_workspace_bytes": (C.c_int64, [_EC, C.c_int64]),
    "haet_balltree_slots": (C.c_int64, [C.c_int64, C.c_int64]),
    "haet_slice_reduce": (C.c_int, [_SH, _P, _P, _P, _P, _P, _P, C.c_float, _P, _P, _P, _P]),
    "haet_slice_ln_workspace_bytes": (C.c_int64, [_SH]),
    "haet_slice_reduce_ln": (C.c_int, [_SH, _P, _P, _P, C.c_float, _P, _P, _P, _P, _P, C.c_float, _P, _P, _P, _P]),
    "haet_slice_halo_workspace_bytes": (C.c_int64, [_SH, C.c_int32, C.c_int32, C.c_int32]),
    "haet_slice_reduce_halo": (C.c_int, [_SH, C.c_int32, C.c_int32, _P, _P, _P, C.c_float, _P, _P, _P, _P, _P, C.c_float, _P, _P,
                                         _P, _P]),
    "haet_slice_merge": (C.c_int, [_SH, _P, C.c_int32, _P, _P, _P, _P]),
    "haet_peer_buffer_bytes": (C.c_int64, [_SH]),
    "haet_peer_exchange": (C.c_int, [_SH, C.POINTER(C.c_void_p), C.c_int32, C.c_int32, _P, _P, _P, _P]),
    "haet_balltree_build": (C.c_int, [_P, C.c_int64, C.c_int64, C.c_int32, C.POINTER(C.c_int32),
                                      C.POINTER(C.c_int32), C.c_int32, C.c_double, _P, _P,
                                      C.POINTER(C.c_void_p), _P]),
    "haet_erwin_forward": (C.c_int, [_EC, _P, _P, C.c_int64, _P, _P, C.c_int64, C.POINTER(C.c_void_p), _P, _P,
                                     _P, _P]),
    "haet_erwin_forward_trees": (C.c_int, [_EC, _P, _P, C.c_int64, C.c_int64, _P, _P, C.c_int64, C.POINTER(C.c_void_p), _P,
                                           _P, _P, _P]),
    "haet_erwin_forward_static": (C.c_int, [_EC, _P, _P, C.c_int64, C.c_int64, _P, _P, C.c_int64, C.POINTER(C.c_void_p), _P,
                                            _P, _P]),
    "haet_deslice_out": (C.c_int, [_SH, _P, _P, _P, _P, _P, _P, C.c_float, _P, _P, _P, _P, _P]),
    "haet_deslice_out_res": (C.c_int, [_SH, _P, _P, _P, _P, _P, _P, C.c_float, _P, _P, _P, _P, _P, _P]),
    "haet_block_mlp_workspace_bytes": (C.c_int64, [C.c_int64, C.c_int32, C.c_int32]),
    "haet_block_mlp": (C.c_int, [_P, C.c_int64, C.c_int32, C.c_int32, _P, _P, C.c_float, _P, _P, _P, _P, _P, _P, _P]),
    "haet_model_prologue_workspace_bytes": (C.c_int64, [C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "haet_model_prologue": (C.c_int, [_P, C.c_int32, _P, C.c_int32, C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                      _P, _P, _P, _P, _P, C.c_int32, _P, _P, _P]),
    "haet_time_embed": (C.c_int, [_P, C.c_int64, C.c_int32, _P, _P, _P, _P, _P, _P, _P]),
    "haet_head_ln_linear": (C.c_int, [_P, C.c_int64, C.c_int32, _P, _P, C.c_float, _P, _P, C.c_int32, _P, _P]),
    "haet_layer_scratch_bytes": (C.c_int64, [_SH, _EC]),
    "haet_layer_forward_host": (C.c_int, [_SH, _EC, C.POINTER(LayerWeights), _P, _P, _P, _P]),
    "haet_layer_forward": (C.c_int, [_SH, _EC, C.POINTER(LayerWeights), _P, _P, _P, _P]),
    "haet_launch_count": (C.c_int64, []),
    "haet_set_path": (C.c_int, [C.c_int]),
    "haet_set_stage_mask": (C.c_int, [C.c_int]),
    "haet_set_raise_threshold": (C.c_int, [C.c_float]),
    "haet_debug_gemm": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_int32, C.c_int32, _P, _P]),
}

_LIB = None


class HaetError(RuntimeError):
    pass


def lib():
    """Load libhaet_b200.so (once).  Raises if it has not been built -- no fallback."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise HaetError(f"{LIB_PATH} is missing: build it with `python -m haet_b200.build` "
                            "(nvcc, sm_100a). haet_b200 has no CPU or PyTorch fallback.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        _LIB = l
    return _LIB


def check(rc: int):
    if rc == HAET_OK:
        return
    msg = lib().haet_last_error().decode(errors="replace")
    if rc == HAET_EINVAL:
        raise ValueError(f"haet_b200: {msg}")
    raise HaetError(f"haet_b200 (code {rc}): {msg}")


def make_shape(B, N, dim, heads, dim_head, slice_num, conv_nd=0, grid=(1, 1, 1), kernel=3, dimensionality=3,
               base_temp=0.5, epsilon=1e-6) -> LayerShape:
    g = list(grid) + [1] * (3 - len(grid))
    return LayerShape(B, N, dim, heads, dim_head, slice_num, conv_nd, (C.c_int32 * 3)(*g), kernel, dimensionality,
                      base_temp, epsilon)


def _arr(vals: Sequence[int]):
    vals = list(vals)
    if len(vals) > MAX_LEVELS:
        raise ValueError(f"at most {MAX_LEVELS} Erwin levels are supported")
    return (C.c_int32 * MAX_LEVELS)(*(vals + [0] * (MAX_LEVELS - len(vals))))


def make_erwin_cfg(c_hidden, ball_sizes, enc_num_heads, enc_depths, dec_num_heads, dec_depths, strides, rotate,
                   decode=True, mlp_ratio=4, dimensionality=3, attn_bias=False) -> ErwinCfg:
    # erwin_flash.py:58-60
    assert len(enc_num_heads) == len(enc_depths) == len(ball_sizes)
    assert len(dec_num_heads) == len(dec_depths) == len(strides)
    assert len(strides) == len(ball_sizes) - 1
    return ErwinCfg(len(ball_sizes), _arr(c_hidden), _arr(ball_sizes), _arr(enc_num_heads), _arr(enc_depths),
                    _arr(dec_num_heads), _arr(dec_depths), _arr(strides), float(rotate), int(bool(decode)),
                    int(mlp_ratio), int(dimensionality), int(bool(attn_bias)), 0, 1e-5, None)
