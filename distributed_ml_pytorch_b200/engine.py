"""Whole-step CUDA-graph training engine.

The reference's hot loop (example/main.py:57-91) issues, per iteration: an H2D copy of the batch,
``zero_grad``, forward, ``F.cross_entropy``, backward, ``optimizer.step()`` (one launch per tensor) and a
blocking ``loss.item()`` — on a B200 that loop is bound by launch latency and host synchronisation, not by
math (AlexNet at batch 64 is ~6 GFLOP per step).  :class:`TrainStep` keeps the same semantics but

* captures ``zero_grad -> forward -> loss -> backward -> local optimizer update -> bf16 shadow refresh`` once
  into a CUDA graph over static buffers and replays it per step (no Python, no allocator, no per-op launch
  overhead on the hot path);
* feeds it from pinned host memory through a copy stream and a double-buffered device staging area, so the PCIe
  transfer of batch i+1 overlaps the replay of step i (set ``DMLB_NVTX=1`` to wrap every step in an NVTX range);
* returns the loss / accuracy counters through a pinned, double-buffered D2H copy that is read one step
  late, so the host never stalls the GPU (the reference blocks on ``loss.item()`` every step, main.py:79);
* leaves the communication part of a distributed optimizer (``communicate()``: gradient push / parameter
  pull / synchronous reduce-scatter+all-gather kernels) outside the graph, launched on the same stream
  right after the replay, because its schedule (every ``n_push`` / ``n_pull`` steps) is host-side state.
"""
from __future__ import annotations

import os
from typing import Callable, Optional, Tuple

import torch

from .ops import functional as OF

_NVTX = os.environ.get("DMLB_NVTX", "0") == "1"


class StepResult:
    """Loss and number of correct predictions of one training step (host floats)."""
    __slots__ = ("loss", "correct", "step")

    def __init__(self, loss: float, correct: float, step: int):
        self.loss, self.correct, self.step = loss, correct, step

    def __repr__(self):
        return f"StepResult(step={self.step}, loss={self.loss:.4f}, correct={self.correct:.0f})"


class TrainStep:
    """Graph-captured training step for ``model`` + ``optimizer`` on a fixed batch shape.

    ``optimizer`` is any optimizer of this package (``FusedSGD``, ``Asynchronous``, ``Synchronous``); if it
    exposes ``step_local()`` / ``communicate()`` the former is captured and the latter runs eagerly after each
    replay, otherwise ``step()`` is captured whole.

    The step runs on its own stream (``self.stream``).  Anything else that touches the model — an evaluation pass on the
    default stream, a checkpoint — must be ordered against it: call ``synchronize()`` first and, afterwards,
    ``self.stream.wait_stream(torch.cuda.current_stream())`` (``example/main.py`` does this around ``evaluate``).
    """

    def __init__(self, model: torch.nn.Module, optimizer, batch_size: int, input_shape: Tuple[int, int, int], device=None,
                 use_graph: bool = True, loss_fn: Optional[Callable] = None, warmup_iters: int = 3):
        self.model, self.optimizer = model, optimizer
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        if self.device.type != "cuda":
            use_graph = False
        self.use_graph = use_graph
        self.batch_size = batch_size
        self.loss_fn = loss_fn or OF.cross_entropy
        self.static_x = torch.zeros((batch_size, *input_shape), dtype=torch.float32, device=self.device)
        # A model whose first layer is a small-channel convolution (RGB input) reads a channels-last, one-16-byte-granule-per-pixel
        # activation: keep the static input in that layout, so the per-step device copy of the batch IS the layout change (one kernel,
        # outside the graph, in place of a memcpy) and the captured step starts directly with the implicit-GEMM convolution.
        self._smallc_in = False
        if self.device.type == "cuda":
            from . import nn as dnn
            from .ops import kernels as K
            first = next((m for m in model.modules() if isinstance(m, dnn.Conv2d)), None)
            if first is not None and first._smallc() and first.in_channels == input_shape[0]:
                C, H, W = input_shape
                rw = K.rowwin_geometry(H, W, *first.kernel_size, *first.stride, *first.padding, first.compute_dtype)
                self.static_x = OF.smallc_input(batch_size, C, H, W, first.compute_dtype, self.device, rw, first.padding)
                self._smallc_in = True
        self.static_y = torch.zeros((batch_size,), dtype=torch.int64, device=self.device)
        self.metrics = torch.zeros(2, dtype=torch.float32, device=self.device)
        self._pinned = None
        self._events = None
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.step_count = 0
        self.kernels_per_step = 0       # launches of our kernels inside one captured step
        self._pending: Optional[int] = None
        self._staging = None
        self._local = getattr(optimizer, "step_local", None) or optimizer.step
        self._communicate = getattr(optimizer, "communicate", None)
        if self.device.type == "cuda":
            self._pinned = [torch.zeros(2, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._events = [torch.cuda.Event() for _ in range(2)]
            self.stream = torch.cuda.Stream(device=self.device)
            shadows = getattr(model, "_dmlb_shadows", None)
            if shadows is not None:
                shadows.prepare()
            if hasattr(optimizer, "fused_zero_grad") and type(optimizer).__name__ != "Synchronous":
                optimizer.zero_grad()                 # start from clean gradients, then let the update kernel keep them clean
                optimizer.fused_zero_grad = True
            if use_graph:
                self._capture(warmup_iters)

    # ------------------------------------------------------------------------------------------
    def _body(self) -> None:
        self.optimizer.zero_grad()
        out = self.model(self.static_x)
        loss = self.loss_fn(out, self.static_y, self.metrics) if self.loss_fn is OF.cross_entropy else self.loss_fn(out, self.static_y)
        loss.backward()
        self._local()

    def _capture(self, warmup_iters: int) -> None:
        from . import _native
        # Warm-up on the side stream: sets kernel attributes, builds shadow tables, sizes the caching
        # allocator.  The parameters are restored afterwards so capture does not change the trajectory.
        arena = getattr(self.model, "_dmlb_arena", None)
        saved = arena.params.clone() if arena is not None else [p.detach().clone() for p in self.model.parameters()]
        extra = self.optimizer.capture_state() if hasattr(self.optimizer, "capture_state") else None
        old_flag = OF.set_unit_loss_grad(True)
        self.stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.stream):
            for _ in range(warmup_iters):
                self._body()
            self.stream.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            before = _native.launch_count()
            with torch.cuda.graph(self.graph, stream=self.stream):
                self._body()
            self.kernels_per_step = _native.launch_count() - before
        OF.set_unit_loss_grad(old_flag)
        with torch.no_grad():
            if arena is not None:
                arena.params.copy_(saved)
                arena.mark_params_changed()
            else:
                for p, s in zip(self.model.parameters(), saved):
                    p.copy_(s)
        if extra is not None:
            self.optimizer.restore_state(extra)
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------------------------------
    def __call__(self, x: torch.Tensor, y: torch.Tensor, sync: bool = False) -> Optional[StepResult]:
        """Run one step on batch ``(x, y)`` (pinned host tensors or device tensors).

        Returns the :class:`StepResult` of the *previous* step (``None`` on the first call) unless
        ``sync=True``, in which case it waits for and returns this step's result.
        """
        if self.device.type != "cuda":
            self.static_x.copy_(x)
            self.static_y.copy_(y)
            self.optimizer.zero_grad()
            out = self.model(self.static_x)
            loss = self.loss_fn(out, self.static_y)
            loss.backward()
            self.optimizer.step()
            self.step_count += 1
            correct = float((out.argmax(1) == self.static_y).sum())
            return StepResult(float(loss), correct, self.step_count - 1)
        if hasattr(self.optimizer, "_sync_lr"):
            self.optimizer._sync_lr()
        if _NVTX:
            torch.cuda.nvtx.range_push(f"dmlb.step{self.step_count}")
        slot = self.step_count & 1
        if not x.is_cuda:
            # Host batch: the PCIe copy runs on its own stream into a double-buffered staging area, so it overlaps the
            # previous step's graph replay; the step itself only does a device-to-device move into the static input.
            if self._staging is None:
                self._staging = [(torch.empty_like(self.static_x), torch.empty_like(self.static_y)) for _ in range(2)]
                self._copy_stream = torch.cuda.Stream(device=self.device)
                self._staged = [torch.cuda.Event() for _ in range(2)]
                self._consumed = [torch.cuda.Event() for _ in range(2)]
                for ev in self._consumed:
                    ev.record(self.stream)
            sx, sy = self._staging[slot]
            self._copy_stream.wait_event(self._consumed[slot])     # the step that last read this slot is past its D2D move
            with torch.cuda.stream(self._copy_stream):
                sx.copy_(x, non_blocking=True)
                sy.copy_(y, non_blocking=True)
                self._staged[slot].record(self._copy_stream)
            self.stream.wait_event(self._staged[slot])
            x, y = sx, sy
        else:
            # Device batch (e.g. utils.data.GpuResidentLoader): it was produced on the caller's current stream.  Order the step's
            # D2D move behind that work, and keep the caching allocator from recycling the batch while this stream still reads it.
            cur = torch.cuda.current_stream(self.device)
            if cur.cuda_stream != self.stream.cuda_stream:
                self.stream.wait_stream(cur)
                x.record_stream(self.stream)
                y.record_stream(self.stream)
        with torch.cuda.stream(self.stream):
            if self._smallc_in:
                OF.fill_smallc_input(self.static_x, x)
            else:
                self.static_x.copy_(x, non_blocking=True)
            self.static_y.copy_(y, non_blocking=True)
            if self._staging is not None and x is self._staging[slot][0]:
                self._consumed[slot].record(self.stream)
            if self.graph is not None:
                self.graph.replay()
            else:
                old_flag = OF.set_unit_loss_grad(True)
                self._body()
                OF.set_unit_loss_grad(old_flag)
            self._pinned[slot].copy_(self.metrics, non_blocking=True)
            self._events[slot].record(self.stream)
            if self._communicate is not None:
                self._communicate()
        if _NVTX:
            torch.cuda.nvtx.range_pop()
        prev = self._pending
        self._pending = slot
        self.step_count += 1
        if sync:
            return self.flush()
        if prev is None:
            return None
        self._events[prev].synchronize()
        return StepResult(float(self._pinned[prev][0]), float(self._pinned[prev][1]), self.step_count - 2)

    def flush(self) -> Optional[StepResult]:
        """Wait for the most recent step and return its result."""
        if self._pending is None:
            return None
        slot, self._pending = self._pending, None
        self._events[slot].synchronize()
        return StepResult(float(self._pinned[slot][0]), float(self._pinned[slot][1]), self.step_count - 1)

    def synchronize(self) -> None:
        if self.device.type == "cuda":
            self.stream.synchronize()
