"""Loader for ``libdmlb.so`` — the in-tree sm_100a kernel library (C ABI in ``include/dmlb.h``).

The library is built by ``build_native.py`` (``__graft_entry__.build()``) with
``nvcc -gencode arch=compute_100a,code=sm_100a`` and loaded through ``ctypes``; kernels receive raw device
pointers (``tensor.data_ptr()``) and the current CUDA stream, so they are capturable in CUDA graphs.

On a machine with a GPU a missing library is a hard error (no silent PyTorch fallback); on a CPU-only
machine :func:`available` is False and the CPU/gloo paths of the framework are used.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path
from typing import Optional

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libdmlb.so"
_lib: Optional[ctypes.CDLL] = None

c_int, c_ll, c_ull, c_float, c_vp = ctypes.c_int, ctypes.c_longlong, ctypes.c_ulonglong, ctypes.c_float, ctypes.c_void_p

# name -> argument ctypes (all return int unless listed in _RESTYPES)
_SIGNATURES = {
    "dmlb_abi_version": [],
    "dmlb_sm_count": [],
    "dmlb_reset_launch_count": [],
    "dmlb_gemm_bf16": [c_vp, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_vp, c_int, c_int, c_int,
                       c_int, c_int, c_int, c_vp],
    "dmlb_conv_implicit_bf16": [c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp,
                                c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_vp],
    # (A, lda, a_mn, B, ldb, b_mn, M, N, K, ws, ldw, counters, n_counters, bias, relu, y, ldy, pool_h, pool_w, pooled, ldp, split_k, tile_n, stream)
    "dmlb_gemm_bf16_fin": [c_vp, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_int, c_int,
                           c_int, c_vp, c_int, c_int, c_int, c_vp],
    # (kind, act, ld_act, B, Ha, Wa, Ca, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, ws, ldw, counters, n_counters, bias, relu, y, ldy,
    #  pool_h, pool_w, pooled, ldp, split_k, tile_n, stream)
    "dmlb_conv_implicit_fin": [c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp,
                               c_int, c_vp, c_int, c_vp, c_int, c_vp, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_vp],
    "dmlb_fin_pool_supported": [c_int, c_int],
    # (images, labels, index, out_x, out_y, B, H, W, C, mean3, std3, stream)
    "dmlb_gather_normalize_u8": [c_vp] * 5 + [c_int] * 4 + [c_vp] * 3,
    "dmlb_splitk_finish": [c_vp, c_int, c_vp, c_int, c_vp, c_int, c_ll, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_vp],
    "dmlb_gemm_bias_act_splitk": [c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_vp, c_vp, c_int, c_vp, c_int, c_int, c_vp],
    "dmlb_im2col_nhwc_bf16": [c_vp, c_vp] + [c_int] * 12 + [c_vp],
    "dmlb_im2col_nchw_f32": [c_vp, c_vp] + [c_int] * 11 + [c_vp],
    "dmlb_col2im_nhwc_bf16": [c_vp, c_vp] + [c_int] * 12 + [c_vp],
    "dmlb_maxpool_fwd_bf16": [c_vp, c_vp] + [c_int] * 9 + [c_vp],
    "dmlb_maxpool_fwd_idx_bf16": [c_vp, c_vp, c_vp] + [c_int] * 9 + [c_vp],
    "dmlb_maxpool_bwd_idx_bf16": [c_vp, c_vp, c_vp] + [c_int] * 9 + [c_vp],
    "dmlb_act_bwd_bf16": [c_vp, c_vp, c_vp, c_vp] + [c_int] * 11 + [c_vp],
    "dmlb_relu_fwd_bf16": [c_vp, c_vp, c_ll, c_int, c_int, c_int, c_vp],
    "dmlb_relu_bwd_bf16": [c_vp, c_vp, c_vp, c_ll, c_int, c_int, c_int, c_int, c_vp],
    "dmlb_softmax_xent": [c_vp, c_int, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_vp],
    "dmlb_softmax_xent_ring": [c_vp, c_int, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_vp, c_vp, c_int, c_vp],
    "dmlb_scale_rows_bf16": [c_vp, c_vp, c_ll, c_int, c_int, c_vp],
    "dmlb_sgd_flat": [c_vp, c_vp, c_vp, c_vp, c_vp, c_ll, c_vp, c_float, c_float, c_int, c_vp],
    "dmlb_refresh_shadows": [c_vp, c_vp, c_vp, c_int, c_vp],
    "dmlb_cast_f32_bf16": [c_vp, c_vp, c_ll, c_int, c_ll, c_int, c_vp],
    "dmlb_cast_bf16_f32": [c_vp, c_vp, c_ll, c_int, c_int, c_ll, c_vp],
    "dmlb_dropout_fwd_bf16": [c_vp, c_vp, c_vp, c_ll, c_int, c_int, c_int, c_int, c_int, c_float, c_ull, c_vp, c_vp],
    "dmlb_dropout_bwd_bf16": [c_vp, c_vp, c_vp, c_ll, c_int, c_int, c_int, c_float, c_vp],
    "dmlb_counter_bump": [c_vp, c_vp],
    "dmlb_bn_stats": [c_vp, c_int, c_vp, c_ll, c_int, c_vp],
    "dmlb_bn_apply": [c_vp, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_ll, c_int, c_float, c_float, c_int, c_vp, c_vp],
    "dmlb_bn_bwd_reduce": [c_vp, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_ll, c_int, c_int, c_vp],
    "dmlb_bn_bwd_apply": [c_vp, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_ll, c_int, c_vp, c_vp],
    "dmlb_avgpool_fwd_bf16": [c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_vp],
    "dmlb_avgpool_bwd_bf16": [c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_vp],
    # (kind, act, B, Ha, Wa, cin, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, out_is_f32, ldo, bias, relu, split_k, pt_layout, pooled, ldp,
    #  pool_w, tile_n, stream)
    #  pool_w, tile_n, hp, wp, stream)
    "dmlb_conv_smallc_bf16": [c_int, c_vp, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_int,
                              c_int, c_vp, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp],
    "dmlb_conv_smallc_tf32": [c_int, c_vp, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_int,
                              c_int, c_vp, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp],
    # (x, y, B, C, H, W, hp, wp, ph, pw, stream)
    "dmlb_nchw_to_act_bf16": [c_vp, c_vp] + [c_int] * 8 + [c_vp],
    "dmlb_nchw_to_act_f32": [c_vp, c_vp] + [c_int] * 8 + [c_vp],
    # (kind, act, ld_act, B, Ha, Wa, Ca, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, out_is_f32, ldo, bias, relu, cluster_s, tile_n,
    #  pooled, ldp, pool_w, act_y, act_dy, act_ldy, act_dbias, act_pool2, act_relu, act_hx, act_wx, stream)
    "dmlb_conv_implicit_cluster_bf16": [c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                        c_vp, c_int, c_int, c_vp, c_int, c_int, c_int, c_vp, c_int, c_int, c_vp, c_vp, c_int, c_vp, c_int, c_int,
                                        c_int, c_int, c_vp],
    "dmlb_conv_implicit_cluster_tf32": [c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                        c_vp, c_int, c_int, c_vp, c_int, c_int, c_int, c_vp, c_int, c_int, c_vp, c_vp, c_int, c_vp, c_int, c_int,
                                        c_int, c_int, c_vp],
    # (ws, ldw, y, dy, ldy, dbias, B, Hx, Wx, C, pool2, relu, is_f32, stream)
    "dmlb_splitk_finish_actbwd": [c_vp, c_int, c_vp, c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp],
    "dmlb_sgd_flat_ex": [c_vp, c_vp, c_vp, c_vp, c_vp, c_ll, c_vp, c_float, c_float, c_int, c_vp, c_int, c_int, c_ll, c_ll, c_vp],
    "dmlb_land_pull": [c_vp, c_vp, c_vp, c_vp, c_ll, c_vp],
    # ---- tf32 path: fp32 tensors, tcgen05 kind::tf32 ----
    # (A, lda, a_mn, B, ldb, b_mn, M, N, K, out, ldo, bias, relu, split_k, atomic, perm_cin, perm_khw, tile_n, stream)
    "dmlb_gemm_tf32": [c_vp, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int,
                       c_vp],
    # (kind, act, ld_act, B, Ha, Wa, Ca, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, ldo, bias, relu, split_k, atomic, tile_n, stream)
    "dmlb_conv_implicit_tf32": [c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp,
                                c_int, c_vp, c_int, c_int, c_int, c_int, c_vp],
    "dmlb_splitk_finish_f32": [c_vp, c_int, c_vp, c_int, c_vp, c_int, c_ll, c_int, c_int, c_int, c_int, c_int, c_vp, c_int, c_vp],
    "dmlb_im2col_nhwc_f32": [c_vp, c_vp] + [c_int] * 12 + [c_vp],
    "dmlb_im2col_nchw_f32_f32": [c_vp, c_vp] + [c_int] * 11 + [c_vp],
    "dmlb_col2im_nhwc_f32": [c_vp, c_vp] + [c_int] * 12 + [c_vp],
    "dmlb_maxpool_fwd_f32": [c_vp, c_vp] + [c_int] * 9 + [c_vp],
    "dmlb_act_bwd_f32": [c_vp, c_vp, c_vp, c_vp] + [c_int] * 11 + [c_vp],
    "dmlb_relu_fwd_f32": [c_vp, c_vp, c_ll, c_int, c_int, c_int, c_vp],
    "dmlb_relu_bwd_f32": [c_vp, c_vp, c_vp, c_ll, c_int, c_int, c_int, c_int, c_vp],
    "dmlb_scale_rows_f32": [c_vp, c_vp, c_ll, c_int, c_int, c_vp],
    "dmlb_refresh_shadows_f32": [c_vp, c_vp, c_vp, c_int, c_vp],
    "dmlb_dropout_fwd_f32": [c_vp, c_vp, c_vp, c_ll, c_int, c_int, c_int, c_int, c_int, c_float, c_ull, c_vp, c_vp],
    "dmlb_dropout_bwd_f32": [c_vp, c_vp, c_vp, c_ll, c_int, c_int, c_int, c_float, c_vp],
}
_RESTYPES = {"dmlb_last_error": ctypes.c_char_p, "dmlb_launch_count": c_ll, "dmlb_reset_launch_count": None}
_OPTIONAL = set()  # entry points that newer libraries add; absent ones are simply not bound


class NativeError(RuntimeError):
    """A libdmlb entry point reported failure; the message is ``dmlb_last_error()``."""


def available() -> bool:
    """True when the kernel library exists on disk (it may still need a GPU to run)."""
    return LIB_PATH.exists()


def lib() -> ctypes.CDLL:
    """Return the loaded library, loading it on first use. Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise NativeError(
                f"{LIB_PATH} is missing: run `python build_native.py` (or __graft_entry__.build()). "
                "The CUDA path of distributed_ml_pytorch_b200 has no PyTorch fallback on GPU machines.")
        handle = ctypes.CDLL(os.fspath(LIB_PATH), mode=ctypes.RTLD_GLOBAL)
        handle.dmlb_last_error.restype = ctypes.c_char_p
        handle.dmlb_last_error.argtypes = []
        handle.dmlb_launch_count.restype = c_ll
        handle.dmlb_launch_count.argtypes = []
        for name, argtypes in _SIGNATURES.items():
            try:
                fn = getattr(handle, name)
            except AttributeError:
                if name in _OPTIONAL:
                    continue
                raise
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, c_int)
        _lib = handle
    return _lib


def register(name: str, argtypes, restype=c_int, optional: bool = False) -> None:
    """Declare one more entry point (used by sub-modules that own their part of the ABI)."""
    _SIGNATURES[name] = list(argtypes)
    if restype is not c_int:
        _RESTYPES[name] = restype
    if optional:
        _OPTIONAL.add(name)
    if _lib is not None:
        try:
            fn = getattr(_lib, name)
        except AttributeError:
            if optional:
                return
            raise
        fn.argtypes = list(argtypes)
        fn.restype = restype


def call(name: str, *args) -> None:
    """Invoke ``name(*args)`` and raise :class:`NativeError` on a non-zero return code."""
    rc = getattr(lib(), name)(*args)
    if rc != 0:
        msg = lib().dmlb_last_error()
        raise NativeError(f"{name} failed (rc={rc}): {msg.decode(errors='replace') if msg else ''}")


def launch_count() -> int:
    return int(lib().dmlb_launch_count())


def reset_launch_count() -> None:
    lib().dmlb_reset_launch_count()
