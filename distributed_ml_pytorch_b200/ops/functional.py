"""Autograd-aware functional ops backed by the sm_100a kernels (CUDA) with plain-PyTorch CPU paths.

On CUDA tensors every op here launches kernels from ``libdmlb.so`` only — tcgen05 GEMMs for
convolutions and linear layers, hand-written kernels for pooling / activation / loss / dropout —
so a training step can be captured as one CUDA graph (:mod:`distributed_ml_pytorch_b200.engine`).
On CPU tensors the same functions fall back to ``torch.nn.functional`` in the input dtype, which is what
the CPU/gloo configuration (BASELINE.json config 1) and the non-GPU tests use.

Reference call sites covered: ``nn.Conv2d`` / ``nn.Linear`` / ``F.max_pool2d`` / ``F.relu`` / ``Dropout2d`` /
``F.dropout`` in example/models.py:8-22,28-48 and ``F.cross_entropy`` in example/main.py:71,125.

Layout contract on CUDA (see ``ops/kernels.py``): activations are channels-last with a channel pitch padded
to 16 bytes; they are ordinary ``torch.Tensor`` views of logical shape ``[B, C, H, W]`` (or ``[B, F]``), so user
code such as ``outputs.max(1)`` keeps working.  Their dtype is the layer's *compute dtype*: ``torch.bfloat16``
(bf16 operands, the default) or ``torch.float32`` (the tf32 path: fp32 activations and weights that the tensor
cores read as tf32 — the reference's own precision, example/main.py:41,53-54,70-73).
"""
from __future__ import annotations

import os
import warnings

from types import SimpleNamespace
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import kernels as K

pad8 = K.pad8
pad_for = K.pad_for
BF16, F32 = torch.bfloat16, torch.float32
_state = SimpleNamespace(unit_loss_grad=False)
_warned = set()


def _warn_once(key: str, msg: str) -> None:
    """Say (once per process and call site) that an op left the hand-written sm_100a path for an ATen kernel."""
    if key not in _warned:
        _warned.add(key)
        warnings.warn(f"distributed_ml_pytorch_b200: {msg}", RuntimeWarning, stacklevel=3)


def _vec(dt) -> int:
    return 4 if dt == F32 else 8


def set_unit_loss_grad(flag: bool) -> bool:
    """Tell :func:`cross_entropy` that ``loss.backward()`` is always seeded with 1 (skips one scale kernel)."""
    old = _state.unit_loss_grad
    _state.unit_loss_grad = bool(flag)
    return old


# ------------------------------------------------------------------------------------------------
# layout helpers
# ------------------------------------------------------------------------------------------------
def new_act4(B: int, C: int, H: int, W: int, device, dt=BF16) -> torch.Tensor:
    """Fresh channels-last activation of dtype ``dt``, logical ``[B, C, H, W]``, channel pitch ``pad_for(dt, C)``."""
    buf = torch.empty((B, H, W, pad_for(dt, C)), dtype=dt, device=device)
    return buf[..., :C].permute(0, 3, 1, 2)


def new_act2(B: int, Fdim: int, device, dtype=BF16) -> torch.Tensor:
    buf = torch.empty((B, pad_for(dtype, Fdim)), dtype=dtype, device=device)
    return buf[:, :Fdim]


def act4_pitch(t: torch.Tensor, dt=BF16) -> Optional[int]:
    """Channel pitch of ``t`` if it is a channels-last activation of dtype ``dt`` in our layout, else ``None``."""
    if t.dtype != dt or t.dim() != 4:
        return None
    v = _vec(dt)
    B, C, H, W = t.shape
    s = t.stride()
    if C > 1 and s[1] != 1:
        return None
    if W > 1:
        ld = s[3]
    elif H > 1:
        ld = s[2]
    elif B > 1:
        ld = s[0]
    else:
        ld = pad_for(dt, C)
    if ld < C or ld % v != 0:
        return None
    if (W > 1 and s[3] != ld) or (H > 1 and s[2] != W * ld) or (B > 1 and s[0] != H * W * ld):
        return None
    if (t.data_ptr() & 15) != 0:
        return None
    return ld


def as_act4(t: torch.Tensor, dt=None) -> Tuple[torch.Tensor, int]:
    """Return ``(tensor, pitch)`` in the kernel layout of dtype ``dt`` (default: ``t``'s own dtype when it is a compute
    dtype, else bf16), converting with one ATen copy only when needed."""
    if dt is None:
        dt = t.dtype if t.dtype in (BF16, F32) else BF16
    ld = act4_pitch(t, dt)
    if ld is not None:
        return t, ld
    B, C, H, W = t.shape
    out = new_act4(B, C, H, W, t.device, dt)
    _warn_once("as_act4", f"a {tuple(t.shape)} {t.dtype} tensor with strides {t.stride()} is not in the channels-last kernel layout; "
                          "converting it with an ATen copy")
    out.copy_(t)
    return out, pad_for(dt, C)


def act2_pitch(t: torch.Tensor, dt=BF16) -> Optional[int]:
    if t.dtype != dt or t.dim() != 2:
        return None
    B, Fd = t.shape
    s = t.stride()
    if Fd > 1 and s[1] != 1:
        return None
    ld = s[0] if B > 1 else pad_for(dt, Fd)
    if ld < Fd or ld % _vec(dt) != 0 or (t.data_ptr() & 15) != 0:
        return None
    return ld


def as_act2(t: torch.Tensor, dt=None) -> Tuple[torch.Tensor, int]:
    if dt is None:
        dt = t.dtype if t.dtype in (BF16, F32) else BF16
    ld = act2_pitch(t, dt)
    if ld is not None:
        return t, ld
    B, Fd = t.shape
    out = new_act2(B, Fd, t.device, dt)
    if dt == BF16 and t.dtype == F32 and t.stride(1) == 1:
        K.cast_f32_bf16(t, out.data_ptr(), B, Fd, t.stride(0), pad8(Fd))
    else:
        _warn_once("as_act2", f"a {tuple(t.shape)} {t.dtype} tensor with strides {t.stride()} is converted to the kernel layout with an ATen copy")
        out.copy_(t)
    return out, pad_for(dt, Fd)


# ------------------------------------------------------------------------------------------------
# side stream for weight-gradient GEMMs
# ------------------------------------------------------------------------------------------------
# In the backward pass of a layer the weight-gradient GEMM and the data-gradient chain (dgrad GEMM, col2im, the next
# layer's activation backward ...) are independent.  At the reference's problem sizes each of them occupies a fraction
# of the 148 SMs, so the wgrad GEMMs are issued on a second stream and joined once, at the end of the backward pass
# (autograd end-of-pass callback); inside a CUDA-graph capture this becomes a parallel branch of the graph.
_side = SimpleNamespace(streams={}, pending={}, enabled=True)


def set_wgrad_side_stream(flag: bool) -> bool:
    old = _side.enabled
    _side.enabled = bool(flag)
    return old


def _side_stream(dev) -> "torch.cuda.Stream":
    key = dev.index if dev.index is not None else torch.cuda.current_device()
    st = _side.streams.get(key)
    if st is None:
        st = torch.cuda.Stream(device=dev)
        _side.streams[key] = st
    return st


def join_side_streams() -> None:
    """Make the current stream wait for every outstanding side-stream wgrad (idempotent)."""
    for key, (main, side) in list(_side.pending.items()):
        main.wait_stream(side)
        del _side.pending[key]


class _WgradFork:
    """``with _WgradFork(dev, tensors...):`` runs the body on the side stream after the current stream's work."""

    def __init__(self, dev, *tensors):
        self.dev = dev
        self.tensors = [t for t in tensors if t is not None]
        self.active = _side.enabled

    def __enter__(self):
        if not self.active:
            return self
        self.main = torch.cuda.current_stream(self.dev)
        self.side = _side_stream(self.dev)
        if self.side.cuda_stream == self.main.cuda_stream:
            self.active = False
            return self
        self.side.wait_stream(self.main)
        self.ctx = torch.cuda.stream(self.side)
        self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if not self.active:
            return False
        self.ctx.__exit__(*exc)
        for t in self.tensors:
            t.record_stream(self.side)
        key = (self.main.cuda_stream, self.side.cuda_stream)
        if key not in _side.pending:
            _side.pending[key] = (self.main, self.side)
            try:
                torch.autograd.Variable._execution_engine.queue_callback(join_side_streams)
            except RuntimeError:      # not inside a backward pass: join right away
                join_side_streams()
        return False


# ------------------------------------------------------------------------------------------------
# gradient-ready notifications (lets the synchronous optimizer start reducing a bucket while backward is still running)
# ------------------------------------------------------------------------------------------------
_grad_hooks = []


def add_grad_ready_hook(fn) -> None:
    """``fn(param, stream)`` is called right after the kernel that finalises ``param``'s arena gradient was launched on ``stream``."""
    if fn not in _grad_hooks:
        _grad_hooks.append(fn)


def remove_grad_ready_hook(fn) -> None:
    if fn in _grad_hooks:
        _grad_hooks.remove(fn)


def _notify_grad_ready(param, dev, stream=None) -> None:
    if not _grad_hooks or param is None:
        return
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    for fn in list(_grad_hooks):
        fn(param, st)


def _sink_ok(param: torch.Tensor, sink: Optional[torch.Tensor]) -> bool:
    return sink is not None and param.grad is not None and param.grad.data_ptr() == sink.data_ptr()


# ------------------------------------------------------------------------------------------------
# conv (+bias +ReLU +max-pool) block
# ------------------------------------------------------------------------------------------------
_conv_opts = SimpleNamespace(implicit=True, smallc=os.environ.get("DMLB_SMALLC", "1") == "1")


def smallc_enabled() -> bool:
    """Small-channel first layers (Cin <= 4) run as implicit GEMMs gathering eight filter taps per k-block (no column matrix)."""
    return _conv_opts.smallc


def smallc_input(B: int, C: int, H: int, W: int, dt, device, rowwin=None, padding=(0, 0)) -> torch.Tensor:
    """Empty first-layer input in the layout the small-channel convolution reads: logical ``[B, C, H, W]`` view of a channels-last
    buffer with one zero-padded 16-byte granule per pixel — for the row-window form (``rowwin=(Hp, Wp)``) the buffer additionally has
    a zero border of the convolution's padding.  Fill it with :func:`fill_smallc_input`.
    :class:`~distributed_ml_pytorch_b200.engine.TrainStep` keeps its static input this way."""
    g = K.granule(dt)
    if rowwin is not None:
        Hp, Wp = rowwin
        ph, pw = padding
        buf = torch.zeros((B, Hp, Wp, g), dtype=dt, device=device)
        out = buf[:, ph:ph + H, pw:pw + W, :C].permute(0, 3, 1, 2)
        out._dmlb_rowwin = SimpleNamespace(buf=buf, Hp=Hp, Wp=Wp, ph=ph, pw=pw)
    else:
        buf = torch.zeros((B, H, W, g), dtype=dt, device=device)
        out = buf[..., :C].permute(0, 3, 1, 2)
        out._dmlb_padzero = True
        out._dmlb_buf = buf
    return out


def fill_smallc_input(dst: torch.Tensor, x: torch.Tensor) -> None:
    """``dst`` (from :func:`smallc_input`) <- the fp32 NCHW batch ``x``: one kernel, layout change (and bf16 cast) included."""
    if x.dtype != torch.float32 or not x.is_contiguous():
        x = x.float().contiguous()
    rw = getattr(dst, "_dmlb_rowwin", None)
    if rw is not None:
        K.nchw_to_act(x, rw.buf, rw.Hp, rw.Wp, rw.ph, rw.pw)
    else:
        K.nchw_to_act(x, dst._dmlb_buf)


def set_implicit_conv(flag: bool) -> bool:
    """Toggle the implicit-GEMM (TMA im2col) convolution path; off = explicit im2col / col2im kernels everywhere."""
    old = _conv_opts.implicit
    _conv_opts.implicit = bool(flag)
    return old


_PAIR_CONV = os.environ.get("DMLB_PAIR_CONV", "0") == "1"
_SMALLC_WGRAD_SPLIT = int(os.environ.get("DMLB_SMALLC_WGRAD_SPLIT", "0"))     # tuning knob: K-split of the first layer's weight gradient

# Backward fusion across two conv blocks (K8).  When block P declares that its output feeds exactly one consumer and that consumer is a
# conv block C (``conv_block(..., sole_consumer=True)``; the model author's promise), C's backward finishes its split-K data-gradient
# GEMM with ONE kernel that also runs P's pool / ReLU backward and bias gradient (`dmlb_splitk_finish_actbwd`).  C then hands autograd a
# zero-stride token of the right shape instead of dx, and P's backward, recognising the token, picks up its ready-made dy.
_fuse = SimpleNamespace(enabled=os.environ.get("DMLB_FUSE_ACT_BWD", "1") == "1", pending={})


def set_fused_act_bwd(flag: bool) -> bool:
    old = _fuse.enabled
    _fuse.enabled = bool(flag)
    return old


class _ConvBlockFn(torch.autograd.Function):
    """Convolution (+bias +ReLU +max-pool) on the tcgen05 GEMM.

    Implicit path (channels-last input whose channels fill whole 128-byte k-blocks: Cin % 64 == 0 in bf16, % 32 in tf32): the
    activation operand of fprop, wgrad and dgrad is gathered by TMA in im2col mode — no column matrix is ever written.  Explicit path (first layer from fp32 NCHW, small or
    odd channel counts, strided dgrad): im2col / col2im kernels around the same GEMM.  Tiny-M layers split K over the SMs
    and finish (bias, ReLU, 2x2 pool, bf16 cast) in one kernel.  Weight gradients accumulate straight into the fp32 arena
    on a side stream."""

    @staticmethod
    def forward(ctx, x, weight, bias, cfg):
        Cout, Cin, KH, KW = weight.shape
        B, _, H, W = x.shape
        SH, SW = cfg.stride
        PH, PW = cfg.padding
        Ho = (H + 2 * PH - KH) // SH + 1
        Wo = (W + 2 * PW - KW) // SW + 1
        M, Kd = B * Ho * Wo, Cin * KH * KW
        dt = cfg.dt
        bk = K.kblock(dt)
        Kpad = pad_for(dt, Kd)
        dev = x.device
        if cfg.smallc:
            return _ConvBlockFn._forward_smallc(ctx, x, weight, bias, cfg, (B, H, W, Cin, Cout, KH, KW, SH, SW, PH, PW, Ho, Wo))
        from_nchw_f32 = act4_pitch(x, dt) is None and x.dtype == torch.float32 and x.is_contiguous()
        direct = (not from_nchw_f32) and KH == 1 and KW == 1 and SH == 1 and SW == 1 and PH == 0 and PW == 0 and Cin % _vec(dt) == 0
        implicit = _conv_opts.implicit and (not from_nchw_f32) and (not direct) and Cin % bk == 0
        ldx, a_t, lda = 0, None, 0
        if from_nchw_f32:
            a_t = torch.empty((M, Kpad), dtype=dt, device=dev)
            K.im2col_nchw_f32(x, a_t, B, H, W, Cin, KH, KW, SH, SW, PH, PW, Kpad)
            lda = Kpad
        else:
            x, ldx = as_act4(x, dt)
            if direct:
                a_t, lda = x, ldx
            elif not implicit:
                a_t = torch.empty((M, Kpad), dtype=dt, device=dev)
                K.im2col_nhwc(x.data_ptr(), a_t, B, H, W, Cin, ldx, KH, KW, SH, SW, PH, PW, Kpad)
                lda = Kpad

        def fprop(out_t, ldo, **kw):
            if implicit:
                K.conv_implicit(0, x.data_ptr(), ldx, B, H, W, Cin, cfg.shadow.data_ptr(), Kpad, KH, KW, SH, SW, PH, PW, Cout, out_t, ldo, dt=dt, **kw)
            else:
                if _PAIR_CONV and dt == BF16 and direct and Cout >= 256 and M >= 8192 and "tile_n" not in kw and not kw.get("atomic"):
                    kw["tile_n"] = 256        # wide 1x1 convolutions: 256 x 256 tiles on CTA pairs (persistent, epilogue overlapped)
                K.gemm(a_t, lda, False, cfg.shadow, Kpad, False, M, Cout, Kd, out_t, ldo, **kw)

        y = new_act4(B, Cout, Ho, Wo, dev, dt)
        ldy = pad_for(dt, Cout)
        split = K.fprop_split(M, Cout, Kd, bk=bk)
        out = y
        fused_pool = False
        plan = K.cluster_plan(M, Cout, Kd, bk) if (split > 1 and implicit) else None
        if plan is not None:
            # tiny-M / long-K layer, cluster form: the K-slices of a tile reduce through distributed shared memory and the epilogue
            # (bias, ReLU, 2x2 pool) runs on the sum — no workspace, no atomics, no finishing kernel
            pool_w = 0
            if cfg.pool == (2, 2, 0) and Ho % 2 == 0 and Wo in (2, 4, 8, 16):
                out = new_act4(B, Cout, Ho // 2, Wo // 2, dev, dt)
                pool_w, fused_pool = Wo, True
            K.conv_implicit_cluster(0, x.data_ptr(), ldx, B, H, W, Cin, cfg.shadow.data_ptr(), Kpad, KH, KW, SH, SW, PH, PW, Cout, y.data_ptr(), ldy,
                                    plan[1], plan[0], bias=bias, relu=cfg.relu, pooled_ptr=out.data_ptr() if pool_w else 0, ldp=ldy, pool_w=pool_w,
                                    dt=dt)
        elif split > 1:
            # tiny-M / long-K layer: spread K over the SMs, reduce in an fp32 workspace, finish (bias, ReLU, bf16, 2x2 pool) in one
            # more kernel — or, opt-in (K.set_fused_finish(True); measured slower), in the GEMM's last-arriving CTA per tile
            ldw = (Cout + 3) // 4 * 4
            ws = K.splitk_workspace(dev, M * ldw)
            want_pool = cfg.pool == (2, 2, 0) and Ho % 2 == 0 and Wo % 2 == 0
            if dt == BF16 and K.fused_finish() and (not want_pool or K.fin_pool_supported(Ho, Wo)):
                if want_pool:
                    out = new_act4(B, Cout, Ho // 2, Wo // 2, dev)
                    fused_pool = True
                fin = dict(pool_hw=(Ho, Wo) if want_pool else None, pooled_ptr=out.data_ptr() if want_pool else 0, ldp=ldy if want_pool else 0)
                if implicit:
                    K.conv_implicit_fin(0, x.data_ptr(), ldx, B, H, W, Cin, cfg.shadow.data_ptr(), Kpad, KH, KW, SH, SW, PH, PW, Cout, ws, ldw, bias,
                                        cfg.relu, y.data_ptr(), ldy, split, **fin)
                else:
                    K.gemm_fin(a_t, lda, False, cfg.shadow, Kpad, False, M, Cout, Kd, ws, ldw, bias, cfg.relu, y.data_ptr(), ldy, split, **fin)
            else:
                fprop(ws, ldw, split_k=split, atomic=True)
                if want_pool:
                    out = new_act4(B, Cout, Ho // 2, Wo // 2, dev, dt)
                    K.splitk_finish(ws, ldw, bias, cfg.relu, y.data_ptr(), ldy, M, Cout, pool2=(B, Ho, Wo), pooled_ptr=out.data_ptr(), ldp=ldy,
                                    dt=dt)
                    fused_pool = True
                else:
                    K.splitk_finish(ws, ldw, bias, cfg.relu, y.data_ptr(), ldy, M, Cout, dt=dt)
        else:
            fprop(y, ldy, bias=bias, relu=cfg.relu)
        if cfg.pool is not None and not fused_pool:
            k, s, p = cfg.pool
            Hp, Wp = (Ho + 2 * p - k) // s + 1, (Wo + 2 * p - k) // s + 1
            out = new_act4(B, Cout, Hp, Wp, dev, dt)
            K.maxpool_fwd(y.data_ptr(), out.data_ptr(), B, Ho, Wo, Cout, ldy, ldy, k, s, p, dt=dt)
        ctx.cfg = cfg
        cfg.fwd = None
        if cfg.sole_consumer and _fuse.enabled:
            pool2 = cfg.pool == (2, 2, 0)
            if (cfg.pool is None or (pool2 and Ho % 2 == 0 and Wo % 2 == 0)) and Cout % 4 == 0:
                cfg.fwd = SimpleNamespace(y=y, B=B, Ho=Ho, Wo=Wo, C=Cout, ldy=ldy, dt=dt, pool2=pool2, relu=cfg.relu, out_ptr=out.data_ptr(),
                                          bias=bias, bsink=cfg.bsink if (bias is not None and _sink_ok(bias, cfg.bsink)) else None,
                                          needs_bias_grad=bias is not None and bias.requires_grad)
        ctx.x_ptr = x.data_ptr()
        ctx.geom = (B, H, W, Cin, Cout, KH, KW, SH, SW, PH, PW, Ho, Wo, Kd, Kpad, ldx, lda, direct, from_nchw_f32, implicit)
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x if implicit else a_t, y if (cfg.relu or cfg.pool is not None) else None, weight, bias)
        return out

    @staticmethod
    def _forward_smallc(ctx, x, weight, bias, cfg, geom):
        """First layers (Cin <= 4): the input is one 16-byte granule per pixel; eight filter taps form one k-block of the implicit GEMM."""
        (B, H, W, Cin, Cout, KH, KW, SH, SW, PH, PW, Ho, Wo) = geom
        dt, dev = cfg.dt, x.device
        g = K.granule(dt)
        rw = cfg.rowwin
        tag = getattr(x, "_dmlb_rowwin", None)
        if rw is not None and tag is not None and (tag.Hp, tag.Wp, tag.ph, tag.pw) == (rw[0], rw[1], PH, PW) and tag.buf.dtype == dt:
            xa = x                                               # already in the padded layout (TrainStep's static input)
        elif rw is None and act4_pitch(x, dt) == g and getattr(x, "_dmlb_padzero", False):
            xa = x
        else:
            xa = smallc_input(B, Cin, H, W, dt, dev, rw, (PH, PW))
            if x.dtype == torch.float32 and x.is_contiguous():
                fill_smallc_input(xa, x)                         # the user's NCHW batch (reference main.py:64-70): one small kernel
            else:
                _warn_once("smallc_in", "a small-channel convolution received an activation that is not an fp32 NCHW batch; re-laying it out "
                                        "with an ATen copy")
                xa.copy_(x)
        xbuf = xa._dmlb_rowwin.buf if rw is not None else xa
        hp, wp = (rw if rw is not None else (0, 0))
        Kg = cfg.shadow.shape[1]                                 # taps (kw padded to 8 in the row-window form) x granule
        M = B * Ho * Wo
        y = new_act4(B, Cout, Ho, Wo, dev, dt)
        ldy = pad_for(dt, Cout)
        out = y
        pool_w = 0
        want_pool2 = cfg.pool == (2, 2, 0) and Ho % 2 == 0 and Wo % 2 == 0
        split = K.fprop_split(M, Cout, Kg, bk=K.kblock(dt))
        done_pool = False
        if split > 1:
            # few row tiles, long reduction (AlexNet conv1: 32 tiles x 22 k-blocks): K over the SMs, then one finish (+ 2x2 pool) kernel
            ldw = (Cout + 3) // 4 * 4
            ws = K.splitk_workspace(dev, M * ldw)
            K.conv_smallc(0, xbuf.data_ptr(), B, H, W, Cin, cfg.shadow.data_ptr(), Kg, KH, KW, SH, SW, PH, PW, Cout, ws, ldw, split_k=split, dt=dt,
                          hp=hp, wp=wp)
            if want_pool2:
                out = new_act4(B, Cout, Ho // 2, Wo // 2, dev, dt)
                K.splitk_finish(ws, ldw, bias, cfg.relu, y.data_ptr(), ldy, M, Cout, pool2=(B, Ho, Wo), pooled_ptr=out.data_ptr(), ldp=ldy, dt=dt)
                done_pool = True
            else:
                K.splitk_finish(ws, ldw, bias, cfg.relu, y.data_ptr(), ldy, M, Cout, dt=dt)
        else:
            if want_pool2 and Wo in (2, 4, 8, 16):
                out = new_act4(B, Cout, Ho // 2, Wo // 2, dev, dt)   # 2x2 pool by warp shuffles in the GEMM epilogue
                pool_w = Wo
                done_pool = True
            K.conv_smallc(0, xbuf.data_ptr(), B, H, W, Cin, cfg.shadow.data_ptr(), Kg, KH, KW, SH, SW, PH, PW, Cout, y, ldy, bias=bias,
                          relu=cfg.relu, pooled_ptr=out.data_ptr() if pool_w else 0, ldp=ldy, pool_w=pool_w, dt=dt, hp=hp, wp=wp)
        if cfg.pool is not None and not done_pool:
            k, s, p = cfg.pool
            Hp_, Wp_ = (Ho + 2 * p - k) // s + 1, (Wo + 2 * p - k) // s + 1
            out = new_act4(B, Cout, Hp_, Wp_, dev, dt)
            K.maxpool_fwd(y.data_ptr(), out.data_ptr(), B, Ho, Wo, Cout, ldy, ldy, k, s, p, dt=dt)
        ctx.cfg = cfg
        cfg.fwd = None
        if cfg.sole_consumer and _fuse.enabled:
            pool2 = cfg.pool == (2, 2, 0)
            if (cfg.pool is None or (pool2 and Ho % 2 == 0 and Wo % 2 == 0)) and Cout % 4 == 0:
                cfg.fwd = SimpleNamespace(y=y, B=B, Ho=Ho, Wo=Wo, C=Cout, ldy=ldy, dt=dt, pool2=pool2, relu=cfg.relu, out_ptr=out.data_ptr(),
                                          bias=bias, bsink=cfg.bsink if (bias is not None and _sink_ok(bias, cfg.bsink)) else None,
                                          needs_bias_grad=bias is not None and bias.requires_grad)
        ctx.x_ptr = xa.data_ptr()
        ctx.geom = (B, H, W, Cin, Cout, KH, KW, SH, SW, PH, PW, Ho, Wo, Cin * KH * KW, Kg, g, 0, False, False, False)
        ctx.has_bias = bias is not None
        ctx.xa_view = xa if rw is None else None                 # granule activation (logical view) for the explicit tf32 wgrad
        ctx.save_for_backward(xbuf, y if (cfg.relu or cfg.pool is not None) else None, weight, bias)
        return out

    @staticmethod
    def backward(ctx, dout):
        cfg = ctx.cfg
        a_t, y, weight, bias = ctx.saved_tensors
        (B, H, W, Cin, Cout, KH, KW, SH, SW, PH, PW, Ho, Wo, Kd, Kpad, ldx, lda, direct, from_nchw_f32, implicit) = ctx.geom
        dev = dout.device
        dt = cfg.dt
        bk = K.kblock(dt)
        M = B * Ho * Wo
        ldy = pad_for(dt, Cout)
        entry = _fuse.pending.pop(dout.data_ptr(), None) if (_fuse.pending and dout.dim() == 4 and dout.stride() == (0, 0, 0, 0)) else None
        if entry is None:
            dout, ldd = as_act4(dout, dt)
        # ---- activation backward (+ pool backward, + bias gradient) ----
        gb = None
        bias_sink = cfg.bsink if (ctx.has_bias and _sink_ok(bias, cfg.bsink)) else None
        if entry is not None:
            # the consumer block already ran this block's activation backward inside its dgrad finish
            dy, lddy, gb = entry.dy, ldy, entry.gb
        elif ctx.has_bias and bias_sink is None and ctx.needs_input_grad[2]:
            gb = torch.zeros(Cout, dtype=torch.float32, device=dev)
        db = bias_sink if bias_sink is not None else gb
        if entry is not None:
            pass
        elif cfg.relu or cfg.pool is not None or db is not None:
            dy = new_act4(B, Cout, Ho, Wo, dev, dt)
            K.act_bwd(y.data_ptr() if y is not None else dy.data_ptr(), dout.data_ptr(), dy.data_ptr(), db, B, Ho, Wo, Cout, ldy, ldd,
                      cfg.pool, cfg.relu, dt=dt)
            lddy = ldy
            if bias_sink is not None:
                _notify_grad_ready(bias, dev)
        else:
            dy, lddy = dout, ldd
        # ---- weight gradient: dW[Cout, K] = dY^T[Cout, M] * Xcol[M, K]  (both operands MN-major; Xcol gathered when implicit) ----
        gw = None
        if ctx.needs_input_grad[1]:
            wsink = cfg.wsink if _sink_ok(weight, cfg.wsink) else None
            tiles = ((Cout + 127) // 128) * ((Kd + 127) // 128)
            split = K.pick_split_k(tiles, M, bk=bk)
            # a channels-last gradient (arena layout [Cout,kh,kw,Cin]) is exactly the GEMM's column order: contiguous
            # rows, vector red.add; a PyTorch-layout gradient needs the (kh,kw,cin)->(cin,kh,kw) column scatter
            if wsink is not None:
                target = wsink
                gemm_order = KH * KW == 1 or Cin == 1 or target.permute(0, 2, 3, 1).is_contiguous()
            elif implicit:
                # no arena: accumulate in GEMM order and hand autograd the permuted view (a strided gradient is legal)
                target = torch.zeros((Cout, KH, KW, Cin), dtype=torch.float32, device=dev)
                gw = target.permute(0, 3, 1, 2)
                gemm_order = True
            else:
                gw = torch.zeros_like(weight, dtype=torch.float32)
                target = gw
                gemm_order = KH * KW == 1 or Cin == 1 or target.permute(0, 2, 3, 1).is_contiguous()
            with _WgradFork(dev, dy, a_t, gw) as fork:
                if cfg.smallc and (cfg.rowwin is not None or dt == BF16):
                    # columns of the GEMM are (tap, padded channel); the epilogue drops the pad channels and writes either order
                    hp, wp = cfg.rowwin if cfg.rowwin is not None else (0, 0)
                    split = K.pick_split_k(((Cout + 127) // 128) * ((Kpad + 127) // 128), M, bk=bk)     # tiles of the PADDED column space
                    if _SMALLC_WGRAD_SPLIT > 0:
                        split = _SMALLC_WGRAD_SPLIT
                    K.conv_smallc(1, a_t.data_ptr(), B, H, W, Cin, dy.data_ptr(), lddy, KH, KW, SH, SW, PH, PW, Cout, target, Kd,
                                  split_k=split, pt_layout=not gemm_order, dt=dt, hp=hp, wp=wp)
                elif cfg.smallc:
                    # tf32 with flat taps: a 32-bit MN-major operand exists only in the 32-byte-granule swizzle, which granule boxes cannot
                    # produce -> column matrix of the real channels (on this side stream, off the critical path) + the ordinary GEMM
                    kp = pad_for(dt, Kd)
                    col = torch.empty((M, kp), dtype=dt, device=dev)
                    K.im2col_nhwc(a_t.data_ptr(), col, B, H, W, Cin, K.granule(dt), KH, KW, SH, SW, PH, PW, kp)
                    K.gemm(dy, lddy, True, col, kp, True, Cout, Kd, M, target, Kd, split_k=split, atomic=True,
                           perm_cin=0 if gemm_order else Cin, perm_khw=KH * KW)
                elif implicit and gemm_order:
                    K.conv_implicit(1, a_t.data_ptr(), ldx, B, H, W, Cin, dy.data_ptr(), lddy, KH, KW, SH, SW, PH, PW, Cout, target, Kd,
                                    split_k=split, atomic=True, dt=dt)
                else:
                    if implicit:   # PyTorch-layout sink with an implicit forward: rebuild the column matrix once
                        col = torch.empty((M, Kpad), dtype=dt, device=dev)
                        K.im2col_nhwc(a_t.data_ptr(), col, B, H, W, Cin, ldx, KH, KW, SH, SW, PH, PW, Kpad)
                        col.record_stream(torch.cuda.current_stream(dev))
                        b_t, ldb = col, Kpad
                    else:
                        b_t, ldb = a_t, lda
                    K.gemm(dy, lddy, True, b_t, ldb, True, Cout, Kd, M, target, Kd, split_k=split, atomic=True,
                           perm_cin=0 if gemm_order else Cin, perm_khw=KH * KW)
            if gw is not None and fork.active:
                join_side_streams()       # autograd consumes a returned gradient immediately
            if wsink is not None:
                _notify_grad_ready(weight, dev, fork.side if fork.active else None)
        # ---- data gradient ----
        dx = None
        if ctx.needs_input_grad[0]:
            if from_nchw_f32:
                raise RuntimeError("gradient w.r.t. an fp32 NCHW network input is not implemented on the CUDA path")
            dx = new_act4(B, Cin, H, W, dev, dt)
            ldxo = pad_for(dt, Cin)
            if cfg.smallc:
                # rare (a small-channel layer that is not the first): explicit transposed conv over the tap-padded weights
                g = K.granule(dt)
                dcol = torch.empty((M, Kpad), dtype=dt, device=dev)
                K.gemm(dy, lddy, False, cfg.shadow, Kpad, True, M, Kpad, Cout, dcol, Kpad)
                K.col2im_nhwc(dcol, dx.data_ptr(), B, H, W, g, ldxo, KH, KW, SH, SW, PH, PW, Kpad)
            elif direct:
                K.gemm(dy, lddy, False, cfg.shadow, Kpad, True, M, Kd, Cout, dx, ldxo)
            elif implicit and SH == 1 and SW == 1 and Cout % bk == 0 and Cin % _vec(dt) == 0 and Kpad == Kd:
                # transposed convolution as an implicit GEMM over dY: no dcol matrix, no col2im
                Min, Kdg = B * H * W, KH * KW * Cout
                dsplit = K.fprop_split(Min, Cin, Kdg, bk=bk)
                dplan = K.cluster_plan(Min, Cin, Kdg, bk, bns=(128, 64)) if dsplit > 1 else None
                pf = cfg.prod.fwd if (cfg.prod is not None and _fuse.enabled) else None
                if pf is not None and not (pf.out_ptr == ctx.x_ptr and pf.dt == dt and pf.C == Cin and pf.B == B and
                                           (pf.Ho, pf.Wo) == ((2 * H, 2 * W) if pf.pool2 else (H, W))):
                    pf = None
                if dplan is not None:
                    # cluster split-K dgrad; when the producer block allows it its pool / ReLU backward + bias gradient run in the epilogue
                    if pf is not None:
                        dyp = new_act4(B, Cin, pf.Ho, pf.Wo, dev, dt)
                        pgb = torch.zeros(Cin, dtype=torch.float32, device=dev) if (pf.bsink is None and pf.needs_bias_grad) else None
                        K.conv_implicit_cluster(2, dy.data_ptr(), lddy, B, Ho, Wo, Cout, cfg.shadow.data_ptr(), Kpad, KH, KW, 1, 1, PH, PW, Cin,
                                                dyp.data_ptr(), pf.ldy, dplan[1], dplan[0],
                                                act=(pf.y.data_ptr(), dyp.data_ptr(), pf.ldy, pf.bsink if pf.bsink is not None else pgb, pf.pool2,
                                                     pf.relu, H, W), dt=dt)
                        if pf.bsink is not None:
                            _notify_grad_ready(pf.bias, dev)
                        token = torch.empty(1, dtype=dt, device=dev)
                        if len(_fuse.pending) > 256:
                            _fuse.pending.clear()
                        _fuse.pending[token.data_ptr()] = SimpleNamespace(dy=dyp, gb=pgb, token=token)
                        dx = token.expand(B, Cin, H, W)
                    else:
                        K.conv_implicit_cluster(2, dy.data_ptr(), lddy, B, Ho, Wo, Cout, cfg.shadow.data_ptr(), Kpad, KH, KW, 1, 1, PH, PW, Cin,
                                                dx.data_ptr(), ldxo, dplan[1], dplan[0], dt=dt)
                elif dsplit > 1:
                    ldw = (Cin + 3) // 4 * 4
                    ws = K.splitk_workspace(dev, Min * ldw)
                    if dt == BF16 and K.fused_finish():
                        K.conv_implicit_fin(2, dy.data_ptr(), lddy, B, Ho, Wo, Cout, cfg.shadow.data_ptr(), Kpad, KH, KW, 1, 1, PH, PW, Cin, ws, ldw,
                                            None, False, dx.data_ptr(), ldxo, dsplit)
                    else:
                        K.conv_implicit(2, dy.data_ptr(), lddy, B, Ho, Wo, Cout, cfg.shadow.data_ptr(), Kpad, KH, KW, 1, 1, PH, PW, Cin, ws, ldw,
                                        split_k=dsplit, atomic=True, dt=dt)
                        pf = cfg.prod.fwd if (cfg.prod is not None and _fuse.enabled) else None
                        if pf is not None and pf.out_ptr == ctx.x_ptr and pf.dt == dt and pf.C == Cin and pf.B == B and \
                                (pf.Ho, pf.Wo) == ((2 * H, 2 * W) if pf.pool2 else (H, W)):
                            # one kernel: sum of the K-splits -> producer's pool / ReLU backward + bias gradient -> producer's dy
                            dyp = new_act4(B, Cin, pf.Ho, pf.Wo, dev, dt)
                            pgb = None
                            if pf.bsink is None and pf.needs_bias_grad:
                                pgb = torch.zeros(Cin, dtype=torch.float32, device=dev)
                            K.splitk_finish_actbwd(ws, ldw, pf.y.data_ptr(), dyp.data_ptr(), pf.ldy, pf.bsink if pf.bsink is not None else pgb, B, H, W,
                                                   Cin, pf.pool2, pf.relu, dt=dt)
                            if pf.bsink is not None:
                                _notify_grad_ready(pf.bias, dev)
                            token = torch.empty(1, dtype=dt, device=dev)
                            if len(_fuse.pending) > 256:
                                _fuse.pending.clear()
                            _fuse.pending[token.data_ptr()] = SimpleNamespace(dy=dyp, gb=pgb, token=token)
                            dx = token.expand(B, Cin, H, W)
                        else:
                            K.splitk_finish(ws, ldw, None, False, dx.data_ptr(), ldxo, Min, Cin, dt=dt)
                else:
                    K.conv_implicit(2, dy.data_ptr(), lddy, B, Ho, Wo, Cout, cfg.shadow.data_ptr(), Kpad, KH, KW, 1, 1, PH, PW, Cin, dx, ldxo,
                                    dt=dt)
            else:
                dcol = torch.empty((M, Kpad), dtype=dt, device=dev)
                K.gemm(dy, lddy, False, cfg.shadow, Kpad, True, M, Kd, Cout, dcol, Kpad)
                K.col2im_nhwc(dcol, dx.data_ptr(), B, H, W, Cin, ldxo, KH, KW, SH, SW, PH, PW, Kpad)
        return dx, gw, gb, None


def conv_block(x, weight, bias=None, stride=(1, 1), padding=(0, 0), relu=False, pool=None, shadow=None, wsink=None, bsink=None,
               sole_consumer=False, smallc=False, rowwin=None):
    """``max_pool(relu(conv2d(x, weight, bias)))`` with the optional parts fused (CUDA) or composed (CPU).

    ``sole_consumer=True`` promises that the result is read by exactly one op and that op is another ``conv_block``: its backward may
    then fold this block's pool / ReLU backward and bias gradient into its split-K dgrad finish (one launch instead of two).
    ``pool`` is ``None`` or ``(kernel, stride, padding)``; ``shadow`` is the GEMM-layout form of ``weight``
    (see :class:`~distributed_ml_pytorch_b200.nn.layers.Conv2d`) whose dtype selects the compute dtype: bf16, or
    fp32 for the tf32 path; ``wsink`` / ``bsink`` are views of the flat fp32 gradient arena that receive the
    parameter gradients directly.
    """
    if not x.is_cuda:
        y = F.conv2d(x, weight, bias, stride=stride, padding=padding)
        if relu:
            y = F.relu(y)
        if pool is not None:
            y = F.max_pool2d(y, pool[0], pool[1], pool[2])
        return y
    if shadow is None:
        raise RuntimeError("conv_block on CUDA needs the GEMM-layout weight shadow (use nn.Conv2d from this package)")
    cfg = SimpleNamespace(stride=tuple(stride), padding=tuple(padding), relu=bool(relu), pool=pool, shadow=shadow, wsink=wsink, bsink=bsink,
                          dt=shadow.dtype, sole_consumer=bool(sole_consumer), prod=getattr(x, "_dmlb_prod", None), fwd=None,
                          smallc=bool(smallc), rowwin=rowwin)
    out = _ConvBlockFn.apply(x, weight, bias, cfg)
    if cfg.fwd is not None:
        out._dmlb_prod = cfg          # "this activation's only reader is a conv block": lets that block fuse our activation backward
    return out


# ------------------------------------------------------------------------------------------------
# linear (+bias +ReLU)
# ------------------------------------------------------------------------------------------------
class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, cfg):
        O, I = weight.shape
        dt = cfg.dt
        x, ldx = as_act2(x, dt)
        B = x.shape[0]
        Kpad = pad_for(dt, I)
        odt = torch.float32 if cfg.out_fp32 else dt
        y = new_act2(B, O, x.device, odt)
        ldo = pad_for(odt, O)
        K.gemm(x, ldx, False, cfg.shadow, Kpad, False, B, O, I, y, ldo, bias=bias, relu=cfg.relu)
        ctx.cfg = cfg
        ctx.ldx = ldx
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x, y if cfg.relu else None, weight, bias)
        return y

    @staticmethod
    def backward(ctx, dout):
        cfg = ctx.cfg
        x, y, weight, bias = ctx.saved_tensors
        O, I = weight.shape
        B = x.shape[0]
        dev = dout.device
        dt = cfg.dt
        Kpad = pad_for(dt, I)
        dout, ldd = as_act2(dout, dt)
        gb = None
        bias_sink = cfg.bsink if (ctx.has_bias and _sink_ok(bias, cfg.bsink)) else None
        if ctx.has_bias and bias_sink is None and ctx.needs_input_grad[2]:
            gb = torch.zeros(O, dtype=torch.float32, device=dev)
        db = bias_sink if bias_sink is not None else gb
        side_bias = None
        if not cfg.relu and bias_sink is not None and _side.enabled and ctx.needs_input_grad[1]:
            # no mask to apply: dy IS dout; the bias gradient (column sums) joins the weight gradient on the side stream
            dy, lddy = dout, ldd
            side_bias = db
        elif cfg.relu or db is not None:
            dy = new_act2(B, O, dev, dt)
            lddy = pad_for(dt, O)
            ldyv = y.stride(0) if (y is not None and B > 1) else lddy
            if ldyv != lddy:          # fp32 logits of a bf16 layer never carry a ReLU, so y and dy share a pitch whenever y is read
                raise RuntimeError("linear backward: saved output and gradient pitches differ")
            K.act_bwd(y.data_ptr() if y is not None else dy.data_ptr(), dout.data_ptr(), dy.data_ptr(), db, B, 1, 1, O, lddy, ldd, None, cfg.relu,
                      dt=dt)
            if bias_sink is not None:
                _notify_grad_ready(bias, dev)
        else:
            dy, lddy = dout, ldd
        gw = None
        if ctx.needs_input_grad[1]:
            wsink = cfg.wsink if _sink_ok(weight, cfg.wsink) else None
            if wsink is None:
                gw = torch.zeros_like(weight, dtype=torch.float32)
            target = wsink if wsink is not None else gw
            tiles = ((O + 127) // 128) * ((I + 127) // 128)
            split = K.pick_split_k(tiles, B, bk=K.kblock(dt))
            pc, pk = (cfg.in_map[0], cfg.in_map[1] * cfg.in_map[2]) if cfg.in_map is not None else (0, 0)
            with _WgradFork(dev, dy, x, gw) as fork:
                if side_bias is not None:
                    scratch = new_act2(B, O, dev, dt)
                    K.act_bwd(scratch.data_ptr(), dy.data_ptr(), scratch.data_ptr(), side_bias, B, 1, 1, O, pad_for(dt, O), lddy, None, False, dt=dt)
                    _notify_grad_ready(bias, dev, fork.side if fork.active else None)
                K.gemm(dy, lddy, True, x, ctx.ldx, True, O, I, B, target, I, split_k=split, atomic=True, perm_cin=pc if pk > 1 else 0,
                       perm_khw=pk)
            if gw is not None and fork.active:
                join_side_streams()
            if wsink is not None:
                _notify_grad_ready(weight, dev, fork.side if fork.active else None)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = new_act2(B, I, dev, dt)
            K.gemm(dy, lddy, False, cfg.shadow, Kpad, True, B, I, O, dx, pad_for(dt, I))
        return dx, gw, gb, None


def linear(x, weight, bias=None, relu=False, out_fp32=False, in_map=None, shadow=None, wsink=None, bsink=None):
    """``relu(x @ weight.T + bias)`` (ReLU optional).  ``in_map=(C, H, W)`` declares that ``x`` is a flattened
    channels-last feature map while ``weight`` columns follow PyTorch's ``(c, h, w)`` order (LeNet's ``fc1``)."""
    if not x.is_cuda:
        y = F.linear(x, weight, bias)
        return F.relu(y) if relu else y
    if shadow is None:
        raise RuntimeError("linear on CUDA needs the GEMM-layout weight shadow (use nn.Linear from this package)")
    cfg = SimpleNamespace(relu=bool(relu), out_fp32=bool(out_fp32), in_map=in_map, shadow=shadow, wsink=wsink, bsink=bsink, dt=shadow.dtype)
    return _LinearFn.apply(x, weight, bias, cfg)


# ------------------------------------------------------------------------------------------------
# standalone pool / relu / flatten / dropout
# ------------------------------------------------------------------------------------------------
class _MaxPoolFn(torch.autograd.Function):
    """Stand-alone max-pool.  C % 8 == 0: forward stores one arg-max byte per output, backward gathers with it; otherwise the
    backward recomputes the arg-max from the saved input (generic scalar kernel)."""

    @staticmethod
    def forward(ctx, x, k, s, p):
        x, ldx = as_act4(x)
        dt = x.dtype
        B, C, H, W = x.shape
        Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
        out = new_act4(B, C, Ho, Wo, x.device, dt)
        ctx.args = (k, s, p, ldx, (B, C, H, W))
        if C % 8 == 0 and dt == BF16:
            idx = torch.empty((B, Ho, Wo, C), dtype=torch.uint8, device=x.device)
            K.maxpool_fwd_idx(x.data_ptr(), out.data_ptr(), idx, B, H, W, C, ldx, pad8(C), k, s, p)
            ctx.indexed = True
            ctx.save_for_backward(idx)
        else:
            K.maxpool_fwd(x.data_ptr(), out.data_ptr(), B, H, W, C, ldx, pad_for(dt, C), k, s, p, dt=dt)
            ctx.indexed = False
            ctx.save_for_backward(x)
        return out

    @staticmethod
    def backward(ctx, dout):
        (saved,) = ctx.saved_tensors
        k, s, p, ldx, (B, C, H, W) = ctx.args
        dt = BF16 if ctx.indexed else saved.dtype
        dout, ldd = as_act4(dout, dt)
        dx = new_act4(B, C, H, W, dout.device, dt)
        if ctx.indexed:
            K.maxpool_bwd_idx(dout.data_ptr(), saved, dx.data_ptr(), B, H, W, C, pad8(C), ldd, k, s, p)
        else:
            x = saved
            if pad_for(dt, C) != ldx:
                x, ldx = as_act4(x.contiguous(), dt)
            K.act_bwd(x.data_ptr(), dout.data_ptr(), dx.data_ptr(), None, B, H, W, C, ldx, ldd, (k, s, p), False, dt=dt)
        return dx, None, None, None


def max_pool2d(x, kernel_size, stride=None, padding=0):
    stride = kernel_size if stride is None else stride
    if not x.is_cuda:
        return F.max_pool2d(x, kernel_size, stride, padding)
    return _MaxPoolFn.apply(x, int(kernel_size), int(stride), int(padding))


class _ReLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        dt = x.dtype
        if x.dim() == 4:
            x, ld = as_act4(x, dt)
            B, C, H, W = x.shape
            y = new_act4(B, C, H, W, x.device, dt)
            rows = B * H * W
        else:
            x, ld = as_act2(x, dt)
            rows, C = x.shape
            y = new_act2(rows, C, x.device, dt)
        K.relu_fwd(x.data_ptr(), y.data_ptr(), rows, C, ld, pad_for(dt, C), dt=dt)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dt = y.dtype
        if y.dim() == 4:
            dy, ldd = as_act4(dy, dt)
            B, C, H, W = y.shape
            dx = new_act4(B, C, H, W, y.device, dt)
            rows = B * H * W
        else:
            dy, ldd = as_act2(dy, dt)
            rows, C = y.shape
            dx = new_act2(rows, C, y.device, dt)
        K.relu_bwd(y.data_ptr(), dy.data_ptr(), dx.data_ptr(), rows, C, pad_for(dt, C), ldd, pad_for(dt, C), dt=dt)
        return dx


def relu(x):
    if not x.is_cuda:
        return F.relu(x)
    if x.dtype not in (BF16, F32) or x.dim() not in (2, 4):
        _warn_once("relu", f"relu on a {x.dtype} {x.dim()}-D CUDA tensor runs the ATen kernel (the sm_100a kernels take bf16 / fp32, 2-D or 4-D)")
        return F.relu(x)
    return _ReLUFn.apply(x)


class _FlattenFn(torch.autograd.Function):
    """[B, C, H, W] channels-last -> [B, H*W*C] (features ordered (h, w, c)); a pure re-view when the channel rows carry no padding."""

    @staticmethod
    def forward(ctx, x):
        x, ld = as_act4(x)
        B, C, H, W = x.shape
        ctx.shape = (B, C, H, W)
        if ld == C:
            return torch.as_strided(x.detach(), (B, H * W * C), (H * W * C, 1), x.storage_offset())
        return x.detach().permute(0, 2, 3, 1).reshape(B, H * W * C)

    @staticmethod
    def backward(ctx, g):
        B, C, H, W = ctx.shape
        g, ldg = as_act2(g)
        if C % _vec(g.dtype) == 0 and ldg == H * W * C:
            return torch.as_strided(g, (B, C, H, W), (H * W * C, 1, W * C, C), g.storage_offset())
        out = new_act4(B, C, H, W, g.device, g.dtype)
        out.copy_(g[:, :H * W * C].reshape(B, H, W, C).permute(0, 3, 1, 2))
        return out


def flatten(x):
    """Flatten a feature map for a following :func:`linear`.

    CPU: PyTorch's ``x.view(B, -1)`` (``(c, h, w)`` order, reference models.py:18,47).  CUDA: channels-last
    ``(h, w, c)`` order — pair it with ``linear(..., in_map=(C, H, W))`` so the weights keep the reference layout.
    """
    if not x.is_cuda:
        return x.reshape(x.size(0), -1)
    if x.dim() == 2:
        return x
    return _FlattenFn.apply(x)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, channel_mode, seed, counter):
        dt = x.dtype
        if x.dim() == 4:
            x, ld = as_act4(x, dt)
            B, C, H, W = x.shape
            rows, rps = B * H * W, H * W
            y = new_act4(B, C, H, W, x.device, dt)
        else:
            x, ld = as_act2(x, dt)
            rows, C = x.shape
            rps = 1
            y = new_act2(rows, C, x.device, dt)
        mask = torch.empty(rows * C, dtype=torch.uint8, device=x.device)
        K.dropout_fwd(x.data_ptr(), y.data_ptr(), mask, rows, C, ld, pad_for(dt, C), rps, channel_mode, p, seed, counter, dt=dt)
        K.counter_bump(counter)
        ctx.args = (p, rows, C, x.dim() == 4, tuple(x.shape), dt)
        ctx.save_for_backward(mask)
        return y

    @staticmethod
    def backward(ctx, dy):
        (mask,) = ctx.saved_tensors
        p, rows, C, is4, shape, dt = ctx.args
        if is4:
            dy, ldd = as_act4(dy, dt)
            dx = new_act4(*shape, dy.device, dt)
        else:
            dy, ldd = as_act2(dy, dt)
            dx = new_act2(rows, C, dy.device, dt)
        K.dropout_bwd(dy.data_ptr(), mask, dx.data_ptr(), rows, C, ldd, pad_for(dt, C), p, dt=dt)
        return dx, None, None, None, None


_dropout_counters = {}


def _counter_for(device) -> torch.Tensor:
    key = (device.type, device.index)
    if key not in _dropout_counters:
        _dropout_counters[key] = torch.zeros(1, dtype=torch.int64, device=device)
    return _dropout_counters[key]


def dropout(x, p=0.5, training=True, channel_mode=False, seed=0x5EED):
    """Element dropout (``F.dropout``) or, with ``channel_mode``, ``Dropout2d`` (whole channels per sample)."""
    if not training or p == 0.0:
        return x
    if not x.is_cuda:
        return F.dropout2d(x, p, True) if channel_mode else F.dropout(x, p, True)
    if x.dtype not in (BF16, F32) or x.dim() not in (2, 4):
        _warn_once("dropout", f"dropout on a {x.dtype} {x.dim()}-D CUDA tensor runs the ATen kernel")
        return F.dropout2d(x, p, True) if channel_mode else F.dropout(x, p, True)
    return _DropoutFn.apply(x, float(p), bool(channel_mode), int(seed), _counter_for(x.device))


# ------------------------------------------------------------------------------------------------
# batch norm (+ residual add + ReLU) and global average pooling  (NORTH-STAR ResNets)
# ------------------------------------------------------------------------------------------------
class _BatchNormFn(torch.autograd.Function):
    """Training-mode BatchNorm2d with the residual add and ReLU of a ResNet block folded into the same two passes."""

    @staticmethod
    def forward(ctx, x, residual, weight, bias, cfg):
        x, ldx = as_act4(x, BF16)
        B, C, H, W = x.shape
        M = B * H * W
        dev = x.device
        res_ptr, ldr = 0, 0
        if residual is not None:
            residual, ldr = as_act4(residual, BF16)
            res_ptr = residual.data_ptr()
        y = new_act4(B, C, H, W, dev)
        # Persistent per-layer scratch [sums | dsum]: each half is re-zeroed by the kernel that runs after its consumer
        # (bn_apply cleans dsum, bn_bwd_apply cleans sums), so a training step launches no memsets for batch norm.
        scratch = cfg.state.get("scratch")
        if scratch is None or scratch.device != dev or scratch.numel() != 4 * C:
            scratch = torch.zeros(4 * C, dtype=torch.float32, device=dev)
            cfg.state["scratch"] = scratch
            cfg.state["sums_clean"] = True
        sums, dsum = scratch[: 2 * C], scratch[2 * C:]
        if not cfg.state.get("sums_clean", False):
            sums.zero_()                 # a forward without a backward (e.g. train-mode evaluation) left statistics behind
        save = torch.empty(2 * C, dtype=torch.float32, device=dev)
        K.bn_stats(x.data_ptr(), ldx, sums, M, C)
        K.bn_apply(x.data_ptr(), ldx, y.data_ptr(), pad8(C), res_ptr, ldr, sums, None, weight, bias, save, cfg.running_mean, cfg.running_var, M, C,
                   cfg.eps, cfg.momentum, cfg.relu, zero=dsum)
        cfg.state["sums_clean"] = False
        ctx.scratch = scratch
        ctx.cfg = cfg
        ctx.dims = (B, C, H, W, ldx)
        ctx.has_res = residual is not None
        ctx.save_for_backward(x, y if cfg.relu else None, weight, bias, save)
        return y

    @staticmethod
    def backward(ctx, dout):
        cfg = ctx.cfg
        x, y, weight, bias, save = ctx.saved_tensors
        B, C, H, W, ldx = ctx.dims
        M = B * H * W
        dev = dout.device
        dout, ldd = as_act4(dout, BF16)
        ldc = pad8(C)
        sums, dsum = ctx.scratch[: 2 * C], ctx.scratch[2 * C:]
        need_g = cfg.relu or ctx.has_res
        g = new_act4(B, C, H, W, dev) if need_g else dout
        K.bn_bwd_reduce(x.data_ptr(), ldx, dout.data_ptr(), ldd, y.data_ptr() if y is not None else 0, ldc, g.data_ptr() if need_g else 0, ldc,
                        save, dsum, M, C, cfg.relu)
        ldg = ldc if need_g else ldd
        wsink = cfg.wsink if (weight is not None and _sink_ok(weight, cfg.wsink)) else None
        bsink = cfg.bsink if (bias is not None and _sink_ok(bias, cfg.bsink)) else None
        gw = torch.zeros(C, dtype=torch.float32, device=dev) if (weight is not None and wsink is None) else None
        gb = torch.zeros(C, dtype=torch.float32, device=dev) if (bias is not None and bsink is None) else None
        dx = new_act4(B, C, H, W, dev)
        K.bn_bwd_apply(x.data_ptr(), ldx, g.data_ptr(), ldg, dx.data_ptr(), ldc, save, weight, dsum, wsink if wsink is not None else gw,
                       bsink if bsink is not None else gb, M, C, zero=sums)
        cfg.state["sums_clean"] = True
        if wsink is not None:
            _notify_grad_ready(weight, dev)
        if bsink is not None:
            _notify_grad_ready(bias, dev)
        return dx, (g if ctx.has_res else None), gw, gb, None


class _BatchNormEvalFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, cfg):
        x, ldx = as_act4(x, BF16)
        B, C, H, W = x.shape
        res_ptr, ldr = 0, 0
        if residual is not None:
            residual, ldr = as_act4(residual, BF16)
            res_ptr = residual.data_ptr()
        y = new_act4(B, C, H, W, x.device)
        stats = torch.cat([cfg.running_mean, torch.rsqrt(cfg.running_var + cfg.eps)]).float()
        K.bn_apply(x.data_ptr(), ldx, y.data_ptr(), pad8(C), res_ptr, ldr, None, stats, weight, bias, None, None, None, B * H * W, C, cfg.eps, 0.0,
                   cfg.relu)
        return y

    @staticmethod
    def backward(ctx, dout):
        raise RuntimeError("backward through inference-mode batch norm is not implemented on the CUDA path")


def batch_norm(x, weight, bias, running_mean, running_var, training=True, momentum=0.1, eps=1e-5, residual=None, relu=False, wsink=None,
               bsink=None, state=None):
    """``relu(batch_norm(x) + residual)`` (residual / ReLU optional) in two kernels forward, two backward."""
    if x.is_cuda and x.dtype == F32 and act4_pitch(x, F32) is not None:
        raise NotImplementedError("batch norm has no fp32-activation (tf32) kernels: the tf32 compute dtype covers the reference's models "
                                  "(LeNet, AlexNet); the ResNets run in bf16")
    if x.is_cuda and (x.shape[1] % 8 != 0 or x.shape[1] > 2048):
        _warn_once("batch_norm", f"batch_norm with C={x.shape[1]} (needs C % 8 == 0 and C <= 2048) runs the ATen kernel in fp32")
    if not x.is_cuda or x.shape[1] % 8 != 0 or x.shape[1] > 2048:
        y = F.batch_norm(x.float() if x.is_cuda else x, running_mean, running_var, weight, bias, training, momentum, eps)
        if residual is not None:
            y = y + residual
        y = F.relu(y) if relu else y
        return y.to(x.dtype) if x.is_cuda else y
    cfg = SimpleNamespace(running_mean=running_mean, running_var=running_var, eps=eps, momentum=momentum, relu=bool(relu), wsink=wsink, bsink=bsink,
                          state=state if state is not None else {})
    if training:
        return _BatchNormFn.apply(x, residual, weight, bias, cfg)
    return _BatchNormEvalFn.apply(x, residual, weight, bias, cfg)


class _AvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x, ldx = as_act4(x, BF16)
        B, C, H, W = x.shape
        y = new_act2(B, C, x.device)
        K.avgpool_fwd(x.data_ptr(), ldx, y.data_ptr(), pad8(C), B, H * W, C)
        ctx.dims = (B, C, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        B, C, H, W = ctx.dims
        dy, ldd = as_act2(dy, BF16)
        dx = new_act4(B, C, H, W, dy.device)
        K.avgpool_bwd(dy.data_ptr(), ldd, dx.data_ptr(), pad8(C), B, H * W, C)
        return dx


def global_avg_pool(x):
    """``adaptive_avg_pool2d(x, 1).flatten(1)``: [B, C, H, W] -> [B, C]."""
    if x.is_cuda and (x.shape[1] % 8 != 0 or x.dtype != BF16):
        _warn_once("global_avg_pool", f"global_avg_pool on a {x.dtype} tensor with C={x.shape[1]} (needs bf16, C % 8 == 0) runs the ATen kernel")
    if not x.is_cuda or x.shape[1] % 8 != 0 or x.dtype != BF16:
        return x.float().mean(dim=(2, 3)).to(x.dtype)
    return _AvgPoolFn.apply(x)


# ------------------------------------------------------------------------------------------------
# softmax cross-entropy (forward + backward in one kernel)
# ------------------------------------------------------------------------------------------------
class _XentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, metrics, ring=None):
        B, C = logits.shape
        if logits.dtype == torch.bfloat16:
            logits, ldl = as_act2(logits)
            dlogits = new_act2(B, C, logits.device)
            ldd = pad8(C)
        else:
            logits = logits.float().contiguous() if (logits.dtype != torch.float32 or logits.stride(1) != 1) else logits
            ldl = logits.stride(0) if B > 1 else C
            dlogits = new_act2(B, C, logits.device, F32)       # padded rows: the gradient is a legal GEMM operand of the tf32 path as is
            ldd = pad_for(F32, C)
        need_grad = ctx.needs_input_grad[0]
        K.softmax_xent(logits, labels, metrics, dlogits if need_grad else None, B, C, ldl, ldd, ring=ring)
        ctx.save_for_backward(dlogits if need_grad else None)
        ctx.dims = (B, C, ldd)
        return metrics[0]

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        B, C, ldd = ctx.dims
        if not _state.unit_loss_grad:
            K.scale_rows(dlogits, g.reshape(1).float(), B, C, ldd)
        return dlogits, None, None, None


def cross_entropy(logits, labels, metrics: Optional[torch.Tensor] = None, ring=None):
    """Mean softmax cross-entropy (reference main.py:71).  On CUDA the kernel also writes
    ``metrics[1]`` = number of correct arg-max predictions (the reference recomputes it with
    ``torch.max(outputs, 1)``, main.py:75) and the returned loss is a view of ``metrics[0]``.  ``ring = (pinned float32 [slots, 4],
    device int64 counter)`` makes the kernel post ``{loss, hits, launch number + 1}`` to host memory as well (see ``engine.TrainStep``)."""
    if not logits.is_cuda:
        return F.cross_entropy(logits.float(), labels)
    if metrics is None:
        metrics = torch.empty(2, dtype=torch.float32, device=logits.device)
    if labels.dtype != torch.int64:
        labels = labels.long()
    loss = _XentFn.apply(logits, labels, metrics, ring)
    return loss
