"""ctypes binding of libconvnets_b200.so (the C ABI declared in include/convnets_b200.h).

This is the only place the shared library is loaded.  There is no fallback: if the library is
missing or fails to load, importing a layer that needs it raises immediately.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libconvnets_b200.so")

CNB_F32, CNB_BF16 = 0, 1
CNB_CONV_BIAS, CNB_CONV_RELU = 1, 2
CNB_ALGO_AUTO, CNB_ALGO_SIMT, CNB_ALGO_TC = 0, 1, 2
CNB_FWD, CNB_DGRAD, CNB_WGRAD = 0, 1, 2
CNB_BN_RELU, CNB_BN_ADD = 1, 2


class ConvDesc(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "N", "H", "W", "C", "K", "R", "S", "stride_h", "stride_w", "pad_h", "pad_w",
        "dil_h", "dil_w", "groups", "dtype", "flags", "algo")]


class PoolDesc(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "N", "H", "W", "C", "kh", "kw", "stride_h", "stride_w", "pad_h", "pad_w", "dtype")]


class BnDesc(C.Structure):
    _fields_ = [("M", C.c_int64), ("C", C.c_int32), ("dtype", C.c_int32), ("flags", C.c_int32),
                ("eps", C.c_float), ("momentum", C.c_float), ("num_batches_tracked", C.c_void_p)]


class PackItem(C.Structure):
    """cnb_pack_item (include/convnets_b200.h): one filter of a batched cnb_pack_filters call."""
    _fields_ = [("w", C.c_void_p), ("w_bf16", C.c_void_p), ("wt_bf16", C.c_void_p),
                ("K", C.c_int32), ("RS", C.c_int32), ("C", C.c_int32), ("first_block", C.c_int32)]


class AdamItem(C.Structure):
    """cnb_adam_item (include/convnets_b200.h): one parameter of a cnb_adam_step call."""
    _fields_ = [("p", C.c_void_p), ("g", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p),
                ("w_bf16", C.c_void_p), ("wt_bf16", C.c_void_p),
                ("K", C.c_int32), ("RS", C.c_int32), ("C", C.c_int32), ("first_block", C.c_int32)]


_P = C.c_void_p
_I = C.c_int
_I32 = C.c_int32
_I64 = C.c_int64
_SZ = C.c_size_t

# name -> (restype, argtypes).  Must list every symbol of include/convnets_b200.h
# (tests/test_abi.py parses the header and checks this table and the .so against it).
SIGNATURES = {
    "cnb_version": (C.c_char_p, []),
    "cnb_last_error": (C.c_char_p, []),
    "cnb_launch_count": (C.c_uint64, []),
    "cnb_conv2d_out_dims": (_I, [C.POINTER(ConvDesc), C.POINTER(_I32), C.POINTER(_I32)]),
    "cnb_conv2d_uses_tc": (_I, [C.POINTER(ConvDesc), _I]),
    "cnb_conv2d_workspace_bytes": (_SZ, [C.POINTER(ConvDesc), _I]),
    "cnb_conv2d_bn_partial_rows": (_I, [C.POINTER(ConvDesc)]),
    "cnb_conv2d_fwd": (_I, [C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_conv2d_dgrad": (_I, [C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_conv2d_dgrad_add": (_I, [C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_conv2d_dgrad_partial_rows": (_I, [C.POINTER(ConvDesc)]),
    "cnb_conv2d_dgrad_bn": (_I, [C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_bn_bwd_from_partials": (_I, [C.POINTER(BnDesc), _P, _P, _P, _P, _P, _P, _P, _I32, _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_rowim2col_nchw": (_I, [C.POINTER(ConvDesc), _P, _P, _I32, _P]),
    "cnb_pack_filter": (_I, [_P, _P, _P, _I32, _I32, _I32, _P]),
    "cnb_pack_filter_blocks": (_I64, [_I32, _I32, _I32]),
    "cnb_pack_filters": (_I, [_P, _I32, _I32, _P]),
    "cnb_conv2d_wgrad": (_I, [C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_bn_workspace_bytes": (_SZ, [C.POINTER(BnDesc)]),
    "cnb_bn_fwd_train": (_I, [C.POINTER(BnDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _P,
                              _SZ, _P]),
    "cnb_bn_static_scratch_bytes": (_SZ, []),
    "cnb_bn_fwd_train_static": (_I, [C.POINTER(BnDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cnb_bn_stats_from_partials": (_I, [C.POINTER(BnDesc), _P, _I32, _P, _P, _P, _SZ, _P]),
    "cnb_bn_fwd_train_prefix": (_I, [C.POINTER(BnDesc), _P, _I64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_bn_bwd_prefix": (_I, [C.POINTER(BnDesc), _P, _I64, _P, _P, _P, _P, _P, _P, _P, _I64, _P, _P, _P, _SZ, _P]),
    "cnb_bn_fwd_eval": (_I, [C.POINTER(BnDesc), _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_bn_bwd": (_I, [C.POINTER(BnDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    "cnb_relu_fwd": (_I, [_P, _P, _I64, _I, _P]),
    "cnb_relu_bwd": (_I, [_P, _P, _P, _I64, _I, _P]),
    "cnb_add_relu_fwd": (_I, [_P, _P, _P, _I64, _I, _P]),
    "cnb_add_relu_bwd": (_I, [_P, _P, _P, _I64, _I, _P]),
    "cnb_add": (_I, [_P, _P, _P, _I64, _I, _P]),
    "cnb_maxpool_fwd": (_I, [C.POINTER(PoolDesc), _P, _P, _P, _P, _P]),
    "cnb_maxpool_bwd": (_I, [C.POINTER(PoolDesc), _P, _P, _P, _P, _P]),
    "cnb_avgpool_fwd": (_I, [C.POINTER(PoolDesc), _P, _P, _P]),
    "cnb_avgpool_bwd": (_I, [C.POINTER(PoolDesc), _P, _P, _P]),
    "cnb_gap_fwd": (_I, [_P, _P, _I32, _I32, _I32, _I, _P]),
    "cnb_gap_bwd": (_I, [_P, _P, _I32, _I32, _I32, _I, _P]),
    "cnb_channel_shuffle": (_I, [_P, _P, _I64, _I32, _I32, _I, _P]),
    "cnb_copy_channels": (_I, [_P, _I32, _I32, _P, _I32, _I32, _I64, _I32, _I, _P]),
    "cnb_nchw_to_nhwc": (_I, [_P, _P, _I32, _I32, _I32, _I32, _I32, _I, _P]),
    "cnb_normalize_u8": (_I, [_P, _P, _I32, _I32, _I32, _I32, _I32, _I32, _I32, _I32, _P, _P, _P, _P]),
    "cnb_im2col_nchw": (_I, [C.POINTER(ConvDesc), _P, _P, _I32, _P]),
    "cnb_adam_item_blocks": (_I64, [_I32, _I32, _I32]),
    "cnb_adam_step": (_I, [_P, _I32, _P, _P]),
    "cnb_cast": (_I, [_P, _I, _P, _I, _I64, _P]),
    "cnb_se_scale_fwd": (_I, [_P, _P, _P, _I32, _I32, _I32, _I, _P]),
    "cnb_se_scale_bwd": (_I, [_P, _P, _P, _P, _P, _I32, _I32, _I32, _I, _P]),
    "cnb_sk_fuse_gap_fwd": (_I, [_P, _P, _P, _I32, _I32, _I32, _I, _P]),
    "cnb_sk_select_fwd": (_I, [_P, _P, _P, _P, _P, _I32, _I32, _I32, _I, _P]),
    "cnb_sk_select_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _I, _P]),
}

_lib = None


class CnbError(RuntimeError):
    """A libconvnets_b200 call returned a non-zero status."""


def load() -> C.CDLL:
    """Load the shared library (once) and attach the prototypes.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CnbError(
            f"{LIB_PATH} not found: build it with `python -m convnets_b200.build` "
            "(or __graft_entry__.build()).  There is no CPU / eager fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = _Calls(lib)
    return _lib


class _Calls:
    """Attribute proxy over the CDLL.  Normally a pass-through; between `trace_begin()` and
    `trace_end()` every entry point that enqueues work is bracketed by CUDA events on the
    launching stream so bench.py can attribute device time to kernels (measurement only)."""

    def __init__(self, cdll):
        self._cdll = cdll
        self._trace = None

    def __getattr__(self, name):
        fn = getattr(self._cdll, name)
        if name in _NO_STREAM:
            setattr(self, name, fn)
            return fn

        def call(*args, _fn=fn, _name=name):
            tr = self._trace
            if tr is None:
                return _fn(*args)
            import torch
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            r = _fn(*args)
            e1.record()
            tr.append((_name, _describe(args), e0, e1))
            return r

        setattr(self, name, call)
        return call


_NO_STREAM = {"cnb_version", "cnb_last_error", "cnb_launch_count", "cnb_conv2d_out_dims",
              "cnb_conv2d_uses_tc", "cnb_conv2d_workspace_bytes", "cnb_bn_workspace_bytes",
              "cnb_bn_static_scratch_bytes",
              "cnb_conv2d_bn_partial_rows", "cnb_conv2d_dgrad_partial_rows", "cnb_adam_item_blocks",
              "cnb_pack_filter_blocks"}


def _describe(args):
    """Plain-Python copy of the scalar arguments / descriptor fields of one call."""
    out = []
    nulls = []
    for i, a in enumerate(args):
        obj = getattr(a, "_obj", None)  # byref(struct)
        if isinstance(obj, C.Structure):
            out.append({f: getattr(obj, f) for f, _ in obj._fields_})
        elif isinstance(a, (int, float)):
            out.append(a)
        elif a is None or (isinstance(a, C.c_void_p) and not a.value):
            nulls.append(i)
    out.append({"_nulls": nulls})   # positions of the NULL pointer arguments (optional tensors)
    return out


def trace_begin() -> None:
    load()._trace = []


def trace_end():
    """-> list of (entry point, args, milliseconds); synchronises the device."""
    import torch
    lib = load()
    tr, lib._trace = lib._trace, None
    torch.cuda.synchronize()
    return [(n, d, e0.elapsed_time(e1)) for n, d, e0, e1 in (tr or [])]


def last_error() -> str:
    return load().cnb_last_error().decode("utf-8", "replace")


def check(status: int, what: str) -> None:
    if status != 0:
        raise CnbError(f"{what} failed (status {status}): {last_error()}")


def version() -> str:
    return load().cnb_version().decode()


def launch_count() -> int:
    return int(load().cnb_launch_count())
