"""Thin, allocation-free Python wrappers over the libdmlb C ABI (``include/dmlb.h``).

Each wrapper takes torch tensors that already have the right dtype/layout, extracts raw pointers and
pitches and launches on the current CUDA stream.  Layout conventions (see ``csrc/conv_ops.cu``):

* activations: channels-last, logical ``[B, C, H, W]`` stored ``[B, H, W, ld]`` (``ld >= C``, rows padded to 16 bytes);
* 2-D activations: ``[rows, C]`` with row pitch ``ld``;
* GEMM weights ("shadows"): ``[Cout, Kpad]`` with K ordered ``(kh, kw, cin)``.

Two compute dtypes share every wrapper through the ``dt`` argument: ``torch.bfloat16`` (bf16 operands and activations,
``ld % 8 == 0``) and ``torch.float32`` (the tf32 path: fp32 activations and weights read by the tensor cores as tf32,
``ld % 4 == 0`` — the precision of the reference, example/main.py:41,53-54,70-73).
"""
from __future__ import annotations

import os

import torch

from .. import _native

_call = _native.call


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def pad8(n: int) -> int:
    return (n + 7) // 8 * 8


BF16, F32 = torch.bfloat16, torch.float32


def pad_for(dt, n: int) -> int:
    """Row length padded to a whole number of 16-byte granules of ``dt`` (8 bf16 / 4 fp32)."""
    return (n + 3) // 4 * 4 if dt == torch.float32 else (n + 7) // 8 * 8


def kblock(dt) -> int:
    """Elements of one 128-byte GEMM k-block: 64 bf16 or 32 fp32 (tf32)."""
    return 32 if dt == torch.float32 else 64


def _sfx(dt) -> str:
    return "f32" if dt == torch.float32 else "bf16"


def gemm(a: torch.Tensor, lda: int, a_mn: bool, b: torch.Tensor, ldb: int, b_mn: bool, M: int, N: int, K: int, out: torch.Tensor,
         ldo: int, bias=None, relu: bool = False, split_k: int = 1, atomic: bool = False, perm_cin: int = 0, perm_khw: int = 0,
         tile_n: int = 0) -> None:
    """``out[M,N] (+)= A[M,K] @ B[N,K]^T`` on the tcgen05 path (``dmlb_gemm_bf16``; fp32 operands: ``dmlb_gemm_tf32``)."""
    if a.dtype == torch.float32:
        if b.dtype != torch.float32 or out.dtype != torch.float32:
            raise TypeError("tf32 GEMM: A, B and out must all be fp32")
        _call("dmlb_gemm_tf32", a.data_ptr(), lda, int(a_mn), b.data_ptr(), ldb, int(b_mn), M, N, K, out.data_ptr(), ldo,
              bias.data_ptr() if bias is not None else None, int(relu), split_k, int(atomic), perm_cin, perm_khw, tile_n, stream_ptr())
        return
    _call("dmlb_gemm_bf16", a.data_ptr(), lda, int(a_mn), b.data_ptr(), ldb, int(b_mn), M, N, K, out.data_ptr(),
          int(out.dtype == torch.float32), ldo, bias.data_ptr() if bias is not None else None, int(relu), split_k, int(atomic),
          perm_cin, perm_khw, tile_n, stream_ptr())


def conv_implicit(kind: int, act_ptr: int, ld_act: int, B, Ha, Wa, Ca, other_ptr: int, ld_other: int, KH, KW, SH, SW, PH, PW, n_out: int,
                  out: torch.Tensor, ldo: int, bias=None, relu: bool = False, split_k: int = 1, atomic: bool = False, tile_n: int = 0,
                  dt=torch.bfloat16) -> None:
    """Implicit-GEMM convolution through TMA im2col (``dmlb_conv_implicit_bf16`` / ``_tf32``): kind 0 fprop, 1 wgrad, 2 dgrad."""
    if dt == torch.float32:
        _call("dmlb_conv_implicit_tf32", kind, act_ptr, ld_act, B, Ha, Wa, Ca, other_ptr, ld_other, KH, KW, SH, SW, PH, PW, n_out, out.data_ptr(),
              ldo, bias.data_ptr() if bias is not None else None, int(relu), split_k, int(atomic), tile_n, stream_ptr())
        return
    _call("dmlb_conv_implicit_bf16", kind, act_ptr, ld_act, B, Ha, Wa, Ca, other_ptr, ld_other, KH, KW, SH, SW, PH, PW, n_out, out.data_ptr(),
          int(out.dtype == torch.float32), ldo, bias.data_ptr() if bias is not None else None, int(relu), split_k, int(atomic), tile_n,
          stream_ptr())


def granule(dt) -> int:
    """Channels in one 16-byte granule: the channel pitch of a small-channel (first-layer) activation."""
    return 4 if dt == torch.float32 else 8


def conv_smallc(kind: int, act_ptr: int, B, Ha, Wa, cin, other_ptr: int, ld_other: int, KH, KW, SH, SW, PH, PW, n_out: int, out: torch.Tensor,
                ldo: int, bias=None, relu: bool = False, split_k: int = 1, pt_layout: bool = False, pooled_ptr: int = 0, ldp: int = 0,
                pool_w: int = 0, tile_n: int = 0, dt=torch.bfloat16, hp: int = 0, wp: int = 0) -> None:
    """Small-channel first-layer convolution (``dmlb_conv_smallc_*``): kind 0 fprop (optional fused 2x2 pool), 1 wgrad.
    ``hp, wp`` > 0 selects the row-window form: ``act_ptr`` is the zero-padded buffer ``[B, hp, wp, granule]``."""
    _call("dmlb_conv_smallc_tf32" if dt == torch.float32 else "dmlb_conv_smallc_bf16", kind, act_ptr, B, Ha, Wa, cin, other_ptr, ld_other, KH, KW,
          SH, SW, PH, PW, n_out, out.data_ptr(), int(out.dtype == torch.float32), ldo, bias.data_ptr() if bias is not None else None, int(relu),
          split_k, int(pt_layout), pooled_ptr or None, ldp, pool_w, tile_n, hp, wp, stream_ptr())


def nchw_to_act(x: torch.Tensor, buf: torch.Tensor, hp: int = 0, wp: int = 0, ph: int = 0, pw: int = 0) -> None:
    """fp32 NCHW batch -> channels-last granule buffer (``dmlb_nchw_to_act_*``): ``buf`` is ``[B, H, W, g]``, or the zero-bordered
    ``[B, hp, wp, g]`` of the row-window convolution with the image at rows ``ph..``, columns ``pw..``."""
    B, C, H, W = x.shape
    _call("dmlb_nchw_to_act_" + _sfx(buf.dtype), x.data_ptr(), buf.data_ptr(), B, C, H, W, hp, wp, ph, pw, stream_ptr())


def _pixel_box(R: int, Q: int, P: int) -> bool:
    """Do R consecutive output pixels (aligned to R) form a box of the (wo, ho, n) index space?  (csrc/gemm_tcgen05.cuh pixel_box)"""
    if Q >= R:
        return Q % R == 0
    if R % Q:
        return False
    a = R // Q
    if a <= P:
        return P % a == 0
    return a % P == 0 and a // P <= 256


def rowwin_geometry(H, W, KH, KW, SH, SW, PH, PW, dt):
    """``(Hp, Wp)`` of the zero-padded input buffer when the row-window convolution applies to this geometry, else ``None``."""
    Ho, Wo = (H + 2 * PH - KH) // SH + 1, (W + 2 * PW - KW) // SW + 1
    if Ho <= 0 or Wo <= 0 or not (_pixel_box(128, Wo, Ho) and _pixel_box(kblock(dt), Wo, Ho)):
        return None
    kwb = (KW + 7) // 8
    return max(H + 2 * PH, (Ho - 1) * SH + KH), max(W + 2 * PW, (Wo - 1) * SW + 8 * kwb)


# Measured on B200 (AlexNet, batch 64, 1 GPU): correct, 8 launches fewer per step, and SLOWER — 0.263 ms/step (tf32, every split layer) vs
# 0.198 with the workspace form; restricted to slices of <= 4 k-blocks (bf16 conv2 / conv3) still 0.205 vs 0.188.  A CTA of these kernels
# streams ~0.4-0.5 us per k-block whatever its size, so the serial loop length decides; a cluster caps the split at 8 where the workspace
# form cuts K 13-32 ways, and gang-scheduling 8 CTAs of 165 KB shared memory each delays the start.  Hence opt-in (DMLB_CLUSTER_SPLITK=1).
_cluster = [os.environ.get("DMLB_CLUSTER_SPLITK", "0") == "1"]


def set_cluster_splitk(on: bool) -> bool:
    """Cluster split-K (K-slices of a tile reduce through distributed shared memory, fused epilogue on the sum) instead of an fp32
    workspace + finishing kernel; returns the old setting."""
    old, _cluster[0] = _cluster[0], bool(on)
    return old


def cluster_plan(M: int, N: int, Kd: int, bk: int, bns=(128, 64, 32), sms: int = 148):
    """``(tile_n, slices)`` of the cluster split-K form for a tiny-M / long-K GEMM, or ``None``: every tile's K is cut into 2, 4 or 8
    slices (one cluster), each CTA finalises ``tile_n / slices`` (4..32) columns; pick what fits the SMs with the fewest bytes per CTA."""
    if not _cluster[0]:
        return None
    nkb = (Kd + bk - 1) // bk
    # Measured on B200 (AlexNet, batch 64): a CTA of these kernels streams ~0.4-0.5 us per 128-byte k-block however few bytes it is, so
    # what counts is the serial loop length.  A cluster caps the K-split at 8, the workspace form does not: the cluster form only wins
    # while each slice keeps a short loop (DMLB_CLUSTER_MAXKB k-blocks, default 4); longer reductions stay on split-K + finish.
    max_kb = int(os.environ.get("DMLB_CLUSTER_MAXKB", "4"))
    tiles_m = (M + 127) // 128
    best = None
    for bn in bns:
        if bn > 32 and bn // 2 >= N:          # tile much wider than the matrix
            continue
        tiles = tiles_m * ((N + bn - 1) // bn)
        for s in (8, 4, 2):
            cs = bn // s
            if s > nkb or cs > 32 or cs % 4:
                continue
            if tiles * s > sms or -(-nkb // s) > max_kb:
                continue
            per = -(-nkb // s) * (16384 + bn * 128)
            if best is None or per < best[0]:
                best = (per, bn, s)
    return (best[1], best[2]) if best else None


def conv_implicit_cluster(kind: int, act_ptr: int, ld_act: int, B, Ha, Wa, Ca, other_ptr: int, ld_other: int, KH, KW, SH, SW, PH, PW, n_out: int,
                          out_ptr: int, ldo: int, cluster_s: int, tile_n: int, bias=None, relu: bool = False, pooled_ptr: int = 0, ldp: int = 0,
                          pool_w: int = 0, act=None, dt=torch.bfloat16) -> None:
    """Implicit convolution, cluster split-K form (``dmlb_conv_implicit_cluster_*``): kind 0 fprop (bias / ReLU / optional 2x2 pool),
    kind 2 dgrad (``act = (y_ptr, dy_ptr, ldy, dbias tensor or None, pool2, relu, Hx, Wx)``: producer block's activation backward fused)."""
    ay, ady, aldy, adb, ap2, arelu, ahx, awx = act if act is not None else (0, 0, 0, None, False, False, 0, 0)
    _call("dmlb_conv_implicit_cluster_tf32" if dt == torch.float32 else "dmlb_conv_implicit_cluster_bf16", kind, act_ptr, ld_act, B, Ha, Wa, Ca,
          other_ptr, ld_other, KH, KW, SH, SW, PH, PW, n_out, out_ptr or None, int(dt == torch.float32), ldo,
          bias.data_ptr() if bias is not None else None, int(relu), cluster_s, tile_n, pooled_ptr or None, ldp, pool_w, ay or None, ady or None, aldy,
          adb.data_ptr() if adb is not None else None, int(ap2), int(arelu), ahx, awx, stream_ptr())


def pick_split_k(tiles: int, k_total: int, sms: int = 148, min_kblocks: int = 2, bk: int = 64) -> int:
    """Split-K factor that fills the SMs when the output has few tiles (weight gradients, tiny-M fprop)."""
    nkb = (k_total + bk - 1) // bk
    min_kblocks = min_kblocks * (64 // bk)
    if tiles >= sms or nkb <= min_kblocks:
        return 1
    want = max(1, (sms + tiles - 1) // tiles)
    return max(1, min(want, nkb // min_kblocks))


_workspaces = {}


def splitk_workspace(device, numel: int) -> torch.Tensor:
    """Persistent all-zero fp32 scratch for split-K forward GEMMs (one per device; the finishing kernel re-zeroes
    what the GEMM touched, so every layer can share it).  It only ever grows, and must first be sized outside CUDA
    graph capture (TrainStep's warm-up iterations do that)."""
    key = (device.type, device.index)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < numel:
        ws = torch.zeros(max(numel, 1 << 20), dtype=torch.float32, device=device)
        _workspaces[key] = ws
    return ws


_fin_counters = {}
_FIN_COUNTERS = 4096
# Measured on B200 (AlexNet step, 8 split-K layers): folded finish 0.2228 ms/step, separate finishing kernel 0.2004 ms/step.  The
# folded form serialises fence -> counter atomic -> fence -> 64 KB L2 read-back on ONE SM per tile, which costs more than a
# PDL-overlapped launch that spreads the same read-back over ~100 CTAs.  So it is opt-in (DMLB_FUSED_FINISH=1).
_fused_finish = [os.environ.get("DMLB_FUSED_FINISH", "0") == "1"]


def set_fused_finish(on: bool) -> bool:
    """Fold the split-K finishing pass into the GEMM kernel, or (default) run it as its own launch; returns the old setting."""
    old, _fused_finish[0] = _fused_finish[0], bool(on)
    return old


def fused_finish() -> bool:
    return _fused_finish[0]


def fin_counters(device) -> torch.Tensor:
    """Per-device tile arrival counters of the fused split-K finish (all zero between kernels; shared by every layer because
    the kernels that use them are ordered on one stream).  First use must happen outside CUDA-graph capture."""
    key = (device.type, device.index)
    c = _fin_counters.get(key)
    if c is None:
        c = torch.zeros(_FIN_COUNTERS, dtype=torch.int32, device=device)
        _fin_counters[key] = c
    return c


def fin_pool_supported(H: int, W: int) -> bool:
    return H > 0 and W > 0 and H % 2 == 0 and W % 2 == 0 and 128 % (2 * W) == 0


def gemm_fin(a, lda, a_mn, b, ldb, b_mn, M, N, K, ws: torch.Tensor, ldw: int, bias, relu: bool, y_ptr: int, ldy: int, split_k: int,
             pool_hw=None, pooled_ptr: int = 0, ldp: int = 0, tile_n: int = 0) -> None:
    """Split-K GEMM whose last-arriving CTA per tile finishes it (bias, ReLU, bf16, optional 2x2 pool): ``dmlb_gemm_bf16_fin``."""
    c = fin_counters(ws.device)
    ph, pw = pool_hw if pool_hw is not None else (0, 0)
    _call("dmlb_gemm_bf16_fin", a.data_ptr(), lda, int(a_mn), b.data_ptr(), ldb, int(b_mn), M, N, K, ws.data_ptr(), ldw, c.data_ptr(),
          c.numel(), bias.data_ptr() if bias is not None else None, int(relu), y_ptr, ldy, ph, pw, pooled_ptr or None, ldp, split_k, tile_n,
          stream_ptr())


def conv_implicit_fin(kind: int, act_ptr: int, ld_act: int, B, Ha, Wa, Ca, other_ptr: int, ld_other: int, KH, KW, SH, SW, PH, PW, n_out: int,
                      ws: torch.Tensor, ldw: int, bias, relu: bool, y_ptr: int, ldy: int, split_k: int, pool_hw=None, pooled_ptr: int = 0,
                      ldp: int = 0, tile_n: int = 0) -> None:
    """Implicit-GEMM convolution (fprop / dgrad), split-K, finishing pass folded in: ``dmlb_conv_implicit_fin``."""
    c = fin_counters(ws.device)
    ph, pw = pool_hw if pool_hw is not None else (0, 0)
    _call("dmlb_conv_implicit_fin", kind, act_ptr, ld_act, B, Ha, Wa, Ca, other_ptr, ld_other, KH, KW, SH, SW, PH, PW, n_out, ws.data_ptr(), ldw,
          c.data_ptr(), c.numel(), bias.data_ptr() if bias is not None else None, int(relu), y_ptr, ldy, ph, pw, pooled_ptr or None, ldp,
          split_k, tile_n, stream_ptr())


def splitk_finish(ws: torch.Tensor, ldw: int, bias, relu: bool, y_ptr: int, ldy: int, M: int, N: int, pool2=None, pooled_ptr: int = 0,
                  ldp: int = 0, dt=torch.bfloat16) -> None:
    B, H, W = pool2 if pool2 is not None else (0, 0, 0)
    _call("dmlb_splitk_finish_f32" if dt == torch.float32 else "dmlb_splitk_finish", ws.data_ptr(), ldw, bias.data_ptr() if bias is not None else None, int(relu), y_ptr, ldy, M, N,
          int(pool2 is not None), B, H, W, pooled_ptr or None, ldp, stream_ptr())


def splitk_finish_actbwd(ws: torch.Tensor, ldw: int, y_ptr: int, dy_ptr: int, ldy: int, dbias, B, Hx, Wx, C, pool2: bool, relu: bool,
                         dt=torch.bfloat16) -> None:
    """Split-K dgrad finish fused with the producer block's pool / ReLU backward and bias gradient (``dmlb_splitk_finish_actbwd``)."""
    _call("dmlb_splitk_finish_actbwd", ws.data_ptr(), ldw, y_ptr, dy_ptr, ldy, dbias.data_ptr() if dbias is not None else None, B, Hx, Wx, C,
          int(pool2), int(relu), int(dt == torch.float32), stream_ptr())


def fprop_split(M: int, N: int, K: int, sms: int = 148, bk: int = 64) -> int:
    """Split-K factor for a forward GEMM, 1 when the plain fused-epilogue kernel is the better choice."""
    bn = 128 if N > 64 else (64 if N > 32 else 32)
    tiles = ((M + 127) // 128) * ((N + bn - 1) // bn)
    nkb = (K + bk - 1) // bk
    r = 64 // bk                      # same byte thresholds for both element sizes
    if tiles * 3 > sms or nkb < 8 * r:
        return 1
    return max(1, min(sms // tiles, nkb // (2 * r)))


def im2col_nhwc(x_ptr: int, col: torch.Tensor, B, H, W, C, ldx, KH, KW, SH, SW, PH, PW, Kpad) -> None:
    _call("dmlb_im2col_nhwc_" + _sfx(col.dtype), x_ptr, col.data_ptr(), B, H, W, C, ldx, KH, KW, SH, SW, PH, PW, Kpad, stream_ptr())


def im2col_nchw_f32(x: torch.Tensor, col: torch.Tensor, B, H, W, C, KH, KW, SH, SW, PH, PW, Kpad) -> None:
    _call("dmlb_im2col_nchw_f32_f32" if col.dtype == torch.float32 else "dmlb_im2col_nchw_f32", x.data_ptr(), col.data_ptr(), B, H, W, C, KH, KW, SH, SW, PH, PW, Kpad, stream_ptr())


def col2im_nhwc(dcol: torch.Tensor, dx_ptr: int, B, H, W, C, ldx, KH, KW, SH, SW, PH, PW, Kpad) -> None:
    _call("dmlb_col2im_nhwc_" + _sfx(dcol.dtype), dcol.data_ptr(), dx_ptr, B, H, W, C, ldx, KH, KW, SH, SW, PH, PW, Kpad, stream_ptr())


def maxpool_fwd(x_ptr: int, y_ptr: int, B, H, W, C, ldx, ldo, k, s, p, dt=torch.bfloat16) -> None:
    _call("dmlb_maxpool_fwd_" + _sfx(dt), x_ptr, y_ptr, B, H, W, C, ldx, ldo, k, s, p, stream_ptr())


def maxpool_fwd_idx(x_ptr: int, y_ptr: int, idx: torch.Tensor, B, H, W, C, ldx, ldo, k, s, p) -> None:
    _call("dmlb_maxpool_fwd_idx_bf16", x_ptr, y_ptr, idx.data_ptr(), B, H, W, C, ldx, ldo, k, s, p, stream_ptr())


def maxpool_bwd_idx(dout_ptr: int, idx: torch.Tensor, dx_ptr: int, B, H, W, C, lddx, ldd, k, s, p) -> None:
    _call("dmlb_maxpool_bwd_idx_bf16", dout_ptr, idx.data_ptr(), dx_ptr, B, H, W, C, lddx, ldd, k, s, p, stream_ptr())


def act_bwd(y_ptr: int, dout_ptr: int, dy_ptr: int, dbias, B, H, W, C, ldy, ldd, pool, relu: bool, dt=torch.bfloat16) -> None:
    k, s, p = pool if pool is not None else (1, 1, 0)
    _call("dmlb_act_bwd_" + _sfx(dt), y_ptr, dout_ptr, dy_ptr, dbias.data_ptr() if dbias is not None else None, B, H, W, C, ldy, ldd,
          int(pool is not None), k, s, p, int(relu), stream_ptr())


def relu_fwd(x_ptr, y_ptr, rows, C, ldx, ldy, dt=torch.bfloat16) -> None:
    _call("dmlb_relu_fwd_" + _sfx(dt), x_ptr, y_ptr, rows, C, ldx, ldy, stream_ptr())


def relu_bwd(y_ptr, dy_ptr, dx_ptr, rows, C, ldy, lddy, lddx, dt=torch.bfloat16) -> None:
    _call("dmlb_relu_bwd_" + _sfx(dt), y_ptr, dy_ptr, dx_ptr, rows, C, ldy, lddy, lddx, stream_ptr())


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, metrics: torch.Tensor, dlogits, B, C, ldl, ldd, ring=None) -> None:
    """``ring = (pinned host float32 tensor [slots, 4], device int64 launch counter)``: also post the result to host memory."""
    if ring is not None:
        _call("dmlb_softmax_xent_ring", logits.data_ptr(), int(logits.dtype == torch.bfloat16), labels.data_ptr(), metrics.data_ptr(),
              dlogits.data_ptr() if dlogits is not None else None, B, C, ldl, ldd, ring[0].data_ptr(), ring[1].data_ptr(), ring[0].shape[0],
              stream_ptr())
        return
    _call("dmlb_softmax_xent", logits.data_ptr(), int(logits.dtype == torch.bfloat16), labels.data_ptr(), metrics.data_ptr(),
          dlogits.data_ptr() if dlogits is not None else None, B, C, ldl, ldd, stream_ptr())


def scale_rows(x: torch.Tensor, scale: torch.Tensor, rows, C, ld) -> None:
    _call("dmlb_scale_rows_" + _sfx(x.dtype), x.data_ptr(), scale.data_ptr(), rows, C, ld, stream_ptr())


def sgd_flat(params: torch.Tensor, grads: torch.Tensor, acc, mom, n: int, lr_dev: torch.Tensor, mu: float = 0.0, wd: float = 0.0,
             shadow=None, zero_grads: bool = False) -> None:
    _call("dmlb_sgd_flat", params.data_ptr(), grads.data_ptr(), acc.data_ptr() if acc is not None else None,
          mom.data_ptr() if mom is not None else None, shadow.data_ptr() if shadow is not None else None, n, lr_dev.data_ptr(), float(mu),
          float(wd), int(zero_grads), stream_ptr())


def sgd_flat_ex(params: torch.Tensor, grads: torch.Tensor, acc, mom, n: int, lr_dev: torch.Tensor, mu: float = 0.0, wd: float = 0.0,
                shadow=None, zero_grads: bool = False, table=None) -> None:
    """:func:`sgd_flat`, optionally refreshing the table-backed weight shadows in the same launch:
    ``table = (device uint8 table, n_entries, dtype of the shadows, lo, hi)`` from ``ShadowSet.fused_table()``."""
    tptr, tn, tf32, lo, hi = (table[0].data_ptr(), table[1], int(table[2] == torch.float32), table[3], table[4]) if table else (None, 0, 0, 0, 0)
    _call("dmlb_sgd_flat_ex", params.data_ptr(), grads.data_ptr(), acc.data_ptr() if acc is not None else None,
          mom.data_ptr() if mom is not None else None, shadow.data_ptr() if shadow is not None else None, n, lr_dev.data_ptr(), float(mu),
          float(wd), int(zero_grads), tptr, tn, tf32, lo, hi, stream_ptr())


def land_pull(params: torch.Tensor, pulled: torch.Tensor, snapshot: torch.Tensor, shadow, n: int) -> None:
    """``params = pulled + (params - snapshot); snapshot = pulled``, bf16 mirror refreshed in the same pass (``dmlb_land_pull``)."""
    _call("dmlb_land_pull", params.data_ptr(), pulled.data_ptr(), snapshot.data_ptr(), shadow.data_ptr() if shadow is not None else None, n,
          stream_ptr())


def refresh_shadows(table: torch.Tensor, n_entries: int, dt=torch.bfloat16) -> None:
    """Absolute-address table: base pointers are NULL, offsets are addresses / element size."""
    _call("dmlb_refresh_shadows_f32" if dt == torch.float32 else "dmlb_refresh_shadows", None, None, table.data_ptr(), n_entries, stream_ptr())


def cast_f32_bf16(x: torch.Tensor, y_ptr: int, rows, C, ldx, ldy) -> None:
    _call("dmlb_cast_f32_bf16", x.data_ptr(), y_ptr, rows, C, ldx, ldy, stream_ptr())


def cast_bf16_f32(x_ptr: int, y: torch.Tensor, rows, C, ldx, ldy) -> None:
    _call("dmlb_cast_bf16_f32", x_ptr, y.data_ptr(), rows, C, ldx, ldy, stream_ptr())


def dropout_fwd(x_ptr, y_ptr, mask: torch.Tensor, rows, C, ldx, ldy, rows_per_sample, channel_mode, p, seed, step: torch.Tensor,
                dt=torch.bfloat16) -> None:
    _call("dmlb_dropout_fwd_" + _sfx(dt), x_ptr, y_ptr, mask.data_ptr(), rows, C, ldx, ldy, rows_per_sample, int(channel_mode), float(p),
          int(seed) & 0xFFFFFFFFFFFFFFFF, step.data_ptr(), stream_ptr())


def dropout_bwd(dy_ptr, mask: torch.Tensor, dx_ptr, rows, C, lddy, lddx, p, dt=torch.bfloat16) -> None:
    _call("dmlb_dropout_bwd_" + _sfx(dt), dy_ptr, mask.data_ptr(), dx_ptr, rows, C, lddy, lddx, float(p), stream_ptr())


def counter_bump(counter: torch.Tensor) -> None:
    _call("dmlb_counter_bump", counter.data_ptr(), stream_ptr())


def _p(t):
    return t.data_ptr() if t is not None else None


def bn_stats(x_ptr, ldx, sums: torch.Tensor, M, C) -> None:
    _call("dmlb_bn_stats", x_ptr, ldx, sums.data_ptr(), M, C, stream_ptr())


def bn_apply(x_ptr, ldx, y_ptr, ldy, res_ptr, ldr, sums, mean_rstd_in, gamma, beta, save, running_mean, running_var, M, C, eps, momentum,
             relu, zero=None) -> None:
    _call("dmlb_bn_apply", x_ptr, ldx, y_ptr, ldy, res_ptr or None, ldr, _p(sums), _p(mean_rstd_in), _p(gamma), _p(beta), _p(save),
          _p(running_mean), _p(running_var), M, C, float(eps), float(momentum), int(relu), _p(zero), stream_ptr())


def bn_bwd_reduce(x_ptr, ldx, dy_ptr, lddy, yact_ptr, ldya, g_ptr, ldg, mean_rstd, dsum, M, C, relu) -> None:
    _call("dmlb_bn_bwd_reduce", x_ptr, ldx, dy_ptr, lddy, yact_ptr or None, ldya, g_ptr or None, ldg, mean_rstd.data_ptr(), dsum.data_ptr(), M, C,
          int(relu), stream_ptr())


def bn_bwd_apply(x_ptr, ldx, g_ptr, ldg, dx_ptr, lddx, mean_rstd, gamma, dsum, dgamma, dbeta, M, C, zero=None) -> None:
    _call("dmlb_bn_bwd_apply", x_ptr, ldx, g_ptr, ldg, dx_ptr, lddx, mean_rstd.data_ptr(), _p(gamma), dsum.data_ptr(), _p(dgamma), _p(dbeta), M, C,
          _p(zero), stream_ptr())


def avgpool_fwd(x_ptr, ldx, y_ptr, ldy, B, P, C) -> None:
    _call("dmlb_avgpool_fwd_bf16", x_ptr, ldx, y_ptr, ldy, B, P, C, stream_ptr())


def avgpool_bwd(dy_ptr, lddy, dx_ptr, lddx, B, P, C) -> None:
    _call("dmlb_avgpool_bwd_bf16", dy_ptr, lddy, dx_ptr, lddx, B, P, C, stream_ptr())
