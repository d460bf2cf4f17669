"""``torch.nn.Module`` layers backed by the sm_100a kernels.

Parameters keep PyTorch's names, shapes and fp32 dtype (``weight [Cout, Cin, kh, kw]``, ``bias``), so
``state_dict()`` checkpoints are interchangeable with the reference's ``nn.Conv2d`` / ``nn.Linear`` models
(reference example/models.py:8-13,28-43).  For the tensor cores each layer additionally
owns a *shadow* of its weight in GEMM layout ``[Cout, Kpad]`` with K ordered ``(kh, kw, cin)``;
shadows are refreshed by one table-driven kernel for the whole model (:class:`ShadowSet`) whenever the
fp32 masters change (optimizer step, parameter pull, ``load_state_dict``).

The shadow's dtype is the layer's *compute dtype* (:func:`set_compute_dtype`): bf16 (default — bf16 operands and
activations), or fp32 for the tf32 path, where the tensor cores read fp32 words as tf32 and activations stay fp32
like the reference's (example/main.py:41,53-54,70-73).  In tf32 mode a weight that already sits in GEMM order with
16-byte rows (every arena-backed AlexNet layer but conv1) needs no copy at all: its shadow *is* its master storage.
"""
from __future__ import annotations

import math
import struct
from typing import List, Optional, Tuple

import torch
import torch.nn as nn

from ..ops import functional as OF
from ..ops import kernels as K


def _pair(v) -> Tuple[int, int]:
    return (v, v) if isinstance(v, int) else (int(v[0]), int(v[1]))


_DTYPES = {"bf16": torch.bfloat16, "bfloat16": torch.bfloat16, "tf32": torch.float32, "fp32": torch.float32, "float32": torch.float32}


def resolve_dtype(dtype) -> torch.dtype:
    """'bf16' | 'tf32' (or a torch dtype) -> the storage dtype of activations / GEMM operands."""
    if isinstance(dtype, torch.dtype):
        if dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("compute dtype must be bf16 or tf32 (fp32 storage)")
        return dtype
    try:
        return _DTYPES[str(dtype).lower()]
    except KeyError:
        raise ValueError(f"unknown compute dtype {dtype!r}: choose 'bf16' or 'tf32'") from None


def set_compute_dtype(model: nn.Module, dtype) -> nn.Module:
    """Select the CUDA compute dtype of every kernel-backed layer of ``model``: ``'bf16'`` or ``'tf32'``.
    Call it before the optimizer (the flat arena keeps a bf16 mirror only for bf16 models)."""
    dt = resolve_dtype(dtype)
    if dt == torch.float32 and any(isinstance(m, BatchNorm2d) for m in model.modules()):
        raise NotImplementedError("the tf32 compute dtype covers the reference's models (LeNet, AlexNet); batch-norm models run in bf16")
    for m in model.modules():
        if isinstance(m, _ShadowedModule):
            if m.compute_dtype != dt:
                m.compute_dtype = dt
                m._shadow, m._shadow_key, m._flat = None, None, False
    model._dmlb_compute_dtype = dt
    ss = getattr(model, "_dmlb_shadows", None)
    if ss is not None:
        ss.invalidate_table()
    return model


def compute_dtype_name(model: nn.Module) -> str:
    return "tf32" if getattr(model, "_dmlb_compute_dtype", torch.bfloat16) == torch.float32 else "bf16"


def _gemm_order(w: torch.Tensor) -> bool:
    """True when a 4-D weight is physically stored as [Cout, kh, kw, Cin] (channels-last), i.e. already in GEMM K order."""
    return w.dim() == 4 and w.permute(0, 2, 3, 1).is_contiguous()


class _ShadowedModule(nn.Module):
    """Mixin for layers with a ``weight`` that is mirrored into a GEMM-layout shadow of the compute dtype.

    Two kinds of shadow storage:
    * *flat*  — bf16: the weight lives in a CUDA :class:`FlatArena` in GEMM order with ``K % 8 == 0``: the shadow is the slice of the
      arena's flat bf16 mirror at the same offset, written by the optimizer kernel itself (no extra launch);
      tf32: the weight's own fp32 storage is in GEMM order with 16-byte rows: the shadow aliases it, nothing is ever copied;
    * *table* — anything else (PyTorch-layout weights, unaligned ``K``, ``in_map`` linears): own storage with padded rows,
      filled by the table-driven ``refresh_shadows`` kernel.
    """

    def _init_shadow_state(self) -> None:
        self.compute_dtype = torch.bfloat16
        self._shadow: Optional[torch.Tensor] = None
        self._shadow_key = None           # (data_ptr, _version) of the weight the shadow was built from
        self._shadow_set: Optional["ShadowSet"] = None
        self._wsink: Optional[torch.Tensor] = None
        self._bsink: Optional[torch.Tensor] = None
        self._arena_slot = None           # (arena, element offset) when the weight is arena-backed on CUDA
        self._flat = False

    # (rows, cin, khw) of the weight viewed as [rows, cin, khw] in PyTorch's logical order
    def _shadow_dims(self) -> Tuple[int, int, int]:
        raise NotImplementedError

    def _needs_permute(self) -> bool:
        """Does the fp32 weight need a (cin,khw)->(khw,cin) re-order on its way into the shadow?"""
        rows, cin, khw = self._shadow_dims()
        if khw == 1 or cin == 1:
            return False
        return not _gemm_order(self.weight)

    def _smallc(self) -> bool:
        """Small-channel layout (first layers, <= 4 input channels): the shadow holds one 16-byte granule per filter tap."""
        return False

    def _kw_pad(self):
        """(kw, kwpad) of the row-window weight layout, or (0, 0): taps in plain (kh, kw) order."""
        return 0, 0

    def _shadow_ld(self) -> int:
        rows, cin, khw = self._shadow_dims()
        if self._smallc():
            kw, kwpad = self._kw_pad()
            taps = khw // kw * kwpad if kwpad else khw
            return taps * K.granule(self.compute_dtype)
        return K.pad_for(self.compute_dtype, cin * khw)

    def _shadow_entry(self) -> bytes:
        rows, cin, khw = self._shadow_dims()
        ld = self._shadow_ld()
        dst = self._shadow.data_ptr() // self._shadow.element_size()
        if self._smallc():
            in_gemm_order = _gemm_order(self.weight) or cin == 1
            kw, kwpad = self._kw_pad()
            return struct.pack("<qqiiiiiiii", self.weight.data_ptr() // 4, dst, rows, cin, khw, ld, K.granule(self.compute_dtype), int(in_gemm_order),
                               kw, kwpad)
        if not self._needs_permute():
            cin, khw = cin * khw, 1       # source already in GEMM order: plain cast with a row pitch
        return struct.pack("<qqiiiiiiii", self.weight.data_ptr() // 4, dst, rows, cin, khw, ld, 0, 0, 0, 0)

    def _ensure_shadow_storage(self) -> None:
        rows, cin, khw = self._shadow_dims()
        kd = cin * khw
        w = self.weight
        slot = self._arena_slot
        flat_ok = False
        if self._smallc():
            ld = self._shadow_ld()
            if self._flat or self._shadow is None or self._shadow.device != w.device or tuple(self._shadow.shape) != (rows, ld) \
                    or self._shadow.dtype != self.compute_dtype:
                self._shadow = torch.zeros((rows, ld), dtype=self.compute_dtype, device=w.device)     # pad channels stay zero for ever
                self._flat, self._shadow_key = False, None
                if self._shadow_set is not None:
                    self._shadow_set.invalidate_table()
            return
        if self.compute_dtype == torch.float32:
            # tf32: fp32 storage already in GEMM order with 16-byte rows is read by TMA as it is
            if w.dim() == 4:
                in_order = khw == 1 and w.is_contiguous() or cin == 1 and w.is_contiguous() or _gemm_order(w)
            else:
                in_order = w.is_contiguous() and khw == 1
            if in_order and kd % 4 == 0 and (w.data_ptr() & 15) == 0 and not self._needs_permute():
                view = torch.as_strided(w.detach(), (rows, kd), (kd, 1), w.storage_offset())
                if self._shadow is None or self._shadow.data_ptr() != view.data_ptr() or not self._flat or self._shadow.dtype != torch.float32:
                    self._shadow, self._flat, self._shadow_key = view, True, None
                    if self._shadow_set is not None:
                        self._shadow_set.invalidate_table()
                return
            ld = K.pad_for(torch.float32, kd)
            if self._flat or self._shadow is None or self._shadow.device != w.device or tuple(self._shadow.shape) != (rows, ld) \
                    or self._shadow.dtype != torch.float32:
                self._shadow = torch.zeros((rows, ld), dtype=torch.float32, device=w.device)
                self._flat, self._shadow_key = False, None
                if self._shadow_set is not None:
                    self._shadow_set.invalidate_table()
            return
        if slot is not None and slot[0].shadow is not None and slot[0].params.device == w.device and not self._needs_permute():
            arena, off = slot
            flat_ok = kd % 8 == 0 and off % 8 == 0 and w.data_ptr() == arena.params.data_ptr() + 4 * off
        if flat_ok:
            arena, off = slot
            view = arena.shadow[off:off + rows * kd].view(rows, kd)
            if self._shadow is None or self._shadow.data_ptr() != view.data_ptr() or not self._flat:
                self._shadow, self._flat, self._shadow_key = view, True, None
                if self._shadow_set is not None:
                    self._shadow_set.invalidate_table()
            return
        ld = K.pad8(kd)
        if self._flat or self._shadow is None or self._shadow.device != w.device or tuple(self._shadow.shape) != (rows, ld) \
                or self._shadow.dtype != torch.bfloat16:
            self._shadow = torch.zeros((rows, ld), dtype=torch.bfloat16, device=w.device)
            self._flat, self._shadow_key = False, None
            if self._shadow_set is not None:
                self._shadow_set.invalidate_table()

    def shadow(self) -> torch.Tensor:
        """Up-to-date bf16 shadow of ``weight`` (refreshing it, or the whole model's set, when stale)."""
        w = self.weight
        key = (w.data_ptr(), w._version)
        if self._shadow is None or self._shadow_key != key or self._shadow.device != w.device:
            if self._shadow_set is not None:
                self._shadow_set.refresh()
            else:
                self._ensure_shadow_storage()
                if self._flat:
                    if self.compute_dtype == torch.bfloat16:
                        arena = self._arena_slot[0]
                        K.cast_f32_bf16(arena.params, arena.shadow.data_ptr(), 1, arena.padded, arena.padded, arena.padded)
                else:
                    table = torch.frombuffer(bytearray(self._shadow_entry()), dtype=torch.uint8).to(w.device)
                    K.refresh_shadows(table, 1, dt=self.compute_dtype)
                self._shadow_key = key
        return self._shadow

    def _mark_shadow_fresh(self) -> None:
        self._shadow_key = (self.weight.data_ptr(), self.weight._version)


class ShadowSet:
    """All shadowed layers of one model, refreshed with at most two kernel launches.

    ``refresh()`` is called by the framework's optimizers right after they update the fp32 masters (the kernels write
    through raw pointers, so ``Tensor._version`` does not change) and lazily by :meth:`_ShadowedModule.shadow` when a
    foreign optimizer or ``load_state_dict`` bumped ``_version``.  ``flat_done=True`` means the caller's kernel already
    wrote the arena's flat bf16 mirror (``sgd_flat`` does), so only table-backed shadows need work.
    """

    def __init__(self, model: nn.Module):
        self.model = model
        self.modules: List[_ShadowedModule] = [m for m in model.modules() if isinstance(m, _ShadowedModule)]
        for m in self.modules:
            m._shadow_set = self
        self._table: Optional[torch.Tensor] = None
        self._n_table = 0
        self._ptrs = None
        self._has_flat = False
        model._dmlb_shadows = self

    def invalidate_table(self) -> None:
        self._table = None

    def _signature(self):
        return [(m.weight.data_ptr(), m._arena_slot[1] if m._arena_slot is not None else -1) for m in self.modules]

    def _build_table(self) -> None:
        for m in self.modules:
            m._ensure_shadow_storage()
        tabled = [m for m in self.modules if not m._flat]
        self._has_flat = any(m._flat for m in self.modules)
        self._n_table = len(tabled)
        dev = self.modules[0].weight.device
        raw = b"".join(m._shadow_entry() for m in tabled) or b"\0" * 48
        self._table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        self._ptrs = self._signature()

    def fused_table(self, arena):
        """``(table, n_entries, shadow dtype, lo, hi)`` for an optimizer kernel that refreshes the table-backed shadows itself
        (``K.sgd_flat_ex``), or ``None`` when there is nothing to refresh or a table-backed weight lives outside ``arena``."""
        if not self.modules or not self.modules[0].weight.is_cuda:
            return None
        if self._table is None or self._ptrs != self._signature():
            self._build_table()
        if not self._n_table:
            return None
        base, n = arena.params.data_ptr(), arena.padded
        lo, hi = n, 0
        for m in self.modules:
            if m._flat:
                continue
            off = (m.weight.data_ptr() - base) // 4
            if off < 0 or off + m.weight.numel() > n:
                return None
            lo, hi = min(lo, off), max(hi, off + m.weight.numel())
        return self._table, self._n_table, self.modules[0].compute_dtype, lo, hi

    def refresh(self, flat_done: bool = False, table_done: bool = False) -> None:
        if not self.modules or not self.modules[0].weight.is_cuda:
            return
        if self._table is None or self._ptrs != self._signature():
            self._build_table()
            flat_done = False
        dt = self.modules[0].compute_dtype
        if self._has_flat and not flat_done and dt == torch.bfloat16:
            arena = getattr(self.model, "_dmlb_arena", None)
            K.cast_f32_bf16(arena.params, arena.shadow.data_ptr(), 1, arena.padded, arena.padded, arena.padded)
        if self._n_table and not table_done:
            K.refresh_shadows(self._table, self._n_table, dt=dt)
        for m in self.modules:
            m._mark_shadow_fresh()

    def prepare(self) -> None:
        """Allocate storage + table now (call before CUDA-graph capture so capture sees no allocation)."""
        if self.modules and self.modules[0].weight.is_cuda:
            self._build_table()
            self.refresh()


def attach_shadows(model: nn.Module) -> ShadowSet:
    """Create (or return) the model-wide :class:`ShadowSet`."""
    s = getattr(model, "_dmlb_shadows", None)
    if s is None:
        s = ShadowSet(model)
    return s


class Conv2d(_ShadowedModule):
    """2-D convolution with optional fused ReLU and max-pool (``activation='relu'``, ``pool=(k, s, p)``).

    Same parameters / initialisation as ``torch.nn.Conv2d`` (reference models.py:8-9,29-39).
    """

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, bias=True, activation=None, pool=None,
                 sole_consumer_is_conv=False):
        super().__init__()
        # the model author's promise that this layer's output is read by exactly one op, another Conv2d of this package: that layer's
        # backward may then run this layer's pool / ReLU backward inside its own split-K dgrad finish (ops/functional.py, K8)
        self.sole_consumer_is_conv = bool(sole_consumer_is_conv)
        self._rowwin = None               # (Hp, Wp) of the padded input buffer when the row-window first-layer form is in use
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.padding = _pair(kernel_size), _pair(stride), _pair(padding)
        self.activation = activation
        self.pool = None if pool is None else (int(pool[0]), int(pool[1]), int(pool[2]) if len(pool) > 2 else 0)
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, *self.kernel_size))
        self.bias = nn.Parameter(torch.empty(out_channels)) if bias else None
        self._init_shadow_state()
        self.reset_parameters()

    def reset_parameters(self) -> None:  # identical to torch.nn.Conv2d
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            fan_in = self.in_channels * self.kernel_size[0] * self.kernel_size[1]
            bound = 1 / math.sqrt(fan_in) if fan_in > 0 else 0
            nn.init.uniform_(self.bias, -bound, bound)

    def _shadow_dims(self):
        return self.out_channels, self.in_channels, self.kernel_size[0] * self.kernel_size[1]

    def _smallc(self) -> bool:
        # RGB / grey-scale first layers: <= 4 channels fit one 16-byte granule in either compute dtype
        return self.in_channels <= 4 and self.kernel_size[0] * self.kernel_size[1] > 1 and OF.smallc_enabled()

    def _kw_pad(self):
        return (self.kernel_size[1], (self.kernel_size[1] + 7) // 8 * 8) if self._rowwin is not None else (0, 0)

    def forward(self, x):
        relu = self.activation == "relu"
        if not x.is_cuda:
            return OF.conv_block(x, self.weight, self.bias, self.stride, self.padding, relu, self.pool)
        smallc = self._smallc()
        if smallc:
            # the weight layout depends on the variant the input geometry allows: row windows (power-of-two outputs) or flat taps
            rw = K.rowwin_geometry(x.shape[2], x.shape[3], *self.kernel_size, *self.stride, *self.padding, self.compute_dtype)
            if rw != self._rowwin:
                self._rowwin = rw
                self._shadow, self._shadow_key, self._flat = None, None, False
                if self._shadow_set is not None:
                    self._shadow_set.invalidate_table()
        return OF.conv_block(x, self.weight, self.bias, self.stride, self.padding, relu, self.pool, shadow=self.shadow(),
                             wsink=self._wsink, bsink=self._bsink, sole_consumer=self.sole_consumer_is_conv, smallc=smallc,
                             rowwin=self._rowwin if smallc else None)

    def extra_repr(self):
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, stride={self.stride}, padding={self.padding}, "
                f"activation={self.activation}, pool={self.pool}")


class Linear(_ShadowedModule):
    """Fully connected layer with optional fused ReLU.  ``in_map=(C, H, W)`` marks the input as a flattened
    channels-last feature map (see :func:`ops.functional.flatten`); the weight keeps PyTorch's column order."""

    def __init__(self, in_features, out_features, bias=True, activation=None, out_fp32=False, in_map=None):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.activation, self.out_fp32, self.in_map = activation, out_fp32, in_map
        if in_map is not None and in_map[0] * in_map[1] * in_map[2] != in_features:
            raise ValueError("in_map does not multiply to in_features")
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.empty(out_features)) if bias else None
        self._init_shadow_state()
        self.reset_parameters()

    def reset_parameters(self) -> None:  # identical to torch.nn.Linear
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            bound = 1 / math.sqrt(self.in_features) if self.in_features > 0 else 0
            nn.init.uniform_(self.bias, -bound, bound)

    def _shadow_dims(self):
        if self.in_map is not None:
            c, h, w = self.in_map
            return self.out_features, c, h * w
        return self.out_features, self.in_features, 1

    def forward(self, x):
        relu = self.activation == "relu"
        if not x.is_cuda:
            return OF.linear(x, self.weight, self.bias, relu)
        return OF.linear(x, self.weight, self.bias, relu, self.out_fp32, self.in_map, shadow=self.shadow(), wsink=self._wsink,
                         bsink=self._bsink)

    def extra_repr(self):
        return f"in_features={self.in_features}, out_features={self.out_features}, activation={self.activation}"


class ReLU(nn.Module):
    """ReLU.  ``fused=True`` makes it a placeholder (identity) because the preceding layer already applied
    it in its GEMM epilogue — kept so ``nn.Sequential`` indices match the reference's ``state_dict`` keys."""

    def __init__(self, inplace: bool = False, fused: bool = False):
        super().__init__()
        self.fused = fused

    def forward(self, x):
        return x if self.fused else OF.relu(x)

    def extra_repr(self):
        return "fused_into_previous" if self.fused else ""


class MaxPool2d(nn.Module):
    """Max pooling; ``fused=True`` = placeholder for a pool already done by the preceding :class:`Conv2d`."""

    def __init__(self, kernel_size, stride=None, padding=0, fused: bool = False):
        super().__init__()
        self.kernel_size, self.stride, self.padding, self.fused = kernel_size, stride or kernel_size, padding, fused

    def forward(self, x):
        return x if self.fused else OF.max_pool2d(x, self.kernel_size, self.stride, self.padding)

    def extra_repr(self):
        return f"kernel_size={self.kernel_size}, stride={self.stride}, padding={self.padding}" + (", fused_into_previous" if self.fused else "")


class Dropout(nn.Module):
    def __init__(self, p: float = 0.5):
        super().__init__()
        self.p = p

    def forward(self, x):
        return OF.dropout(x, self.p, self.training, channel_mode=False)


class Dropout2d(nn.Module):
    def __init__(self, p: float = 0.5):
        super().__init__()
        self.p = p

    def forward(self, x):
        return OF.dropout(x, self.p, self.training, channel_mode=True)


class Flatten(nn.Module):
    def forward(self, x):
        return OF.flatten(x)


class BatchNorm2d(nn.Module):
    """``torch.nn.BatchNorm2d``-compatible parameters/buffers (``weight``, ``bias``, ``running_mean``, ``running_var``,
    ``num_batches_tracked``) with optional fused residual add and ReLU: ``forward(x, residual=None)``."""

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: float = 0.1, activation=None):
        super().__init__()
        self.num_features, self.eps, self.momentum, self.activation = num_features, eps, momentum, activation
        self.weight = nn.Parameter(torch.ones(num_features))
        self.bias = nn.Parameter(torch.zeros(num_features))
        self.register_buffer("running_mean", torch.zeros(num_features))
        self.register_buffer("running_var", torch.ones(num_features))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))
        self._wsink = None
        self._bsink = None
        self._arena_slot = None
        self._state = {}                  # persistent statistics scratch of the CUDA kernels

    def forward(self, x, residual=None):
        # num_batches_tracked is bookkeeping only (momentum is fixed); it is not bumped on the CUDA-graph path
        return OF.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var, self.training, self.momentum, self.eps,
                             residual=residual, relu=self.activation == "relu", wsink=self._wsink, bsink=self._bsink, state=self._state)

    def extra_repr(self):
        return f"{self.num_features}, eps={self.eps}, momentum={self.momentum}, activation={self.activation}"


class GlobalAvgPool(nn.Module):
    def forward(self, x):
        return OF.global_avg_pool(x)
