"""Whole-step CUDA-graph training engine.

The reference's hot loop (example/main.py:57-91) issues, per iteration: an H2D copy of the batch,
``zero_grad``, forward, ``F.cross_entropy``, backward, ``optimizer.step()`` (one launch per tensor) and a
blocking ``loss.item()`` — on a B200 that loop is bound by launch latency and host synchronisation, not by
math (AlexNet at batch 64 is ~6 GFLOP per step).  :class:`TrainStep` keeps the same semantics but

* captures ``zero_grad -> forward -> loss -> backward -> local optimizer update -> bf16 shadow refresh`` once
  into a CUDA graph over static buffers and replays it per step (no Python, no allocator, no per-op launch
  overhead on the hot path);
* feeds it through a copy stream and a double-buffered device staging area (one captured graph per staging slot, the
  first kernel of each reads its slot), so the PCIe transfer of batch i+1 overlaps the replay of step i and the step
  stream carries nothing but graph launches back to back — measured on B200: 189 us/step for replays alone, 207 us
  with a layout kernel, a label copy, a D2H copy and an event between them (set ``DMLB_NVTX=1`` for NVTX ranges);
* returns the loss / accuracy counters through a ring in pinned host memory that the loss kernel itself writes
  (value, then sequence number); the host reads it one step late, so nothing ever stalls the GPU and no D2H copy or
  event sits between two replays (the reference blocks on ``loss.item()`` every step, main.py:79);
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

    RING = 64        # slots of the pinned result ring (the host reads one step late: two would do)

    def __init__(self, model: torch.nn.Module, optimizer, batch_size: int, input_shape: Tuple[int, int, int], device=None,
                 use_graph: bool = True, loss_fn: Optional[Callable] = None, warmup_iters: int = 3):
        self.model, self.optimizer = model, optimizer
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        if self.device.type != "cuda":
            use_graph = False
        self.use_graph = use_graph
        self.batch_size = batch_size
        self.loss_fn = loss_fn or OF.cross_entropy
        self._own_loss = self.loss_fn is OF.cross_entropy
        cuda = self.device.type == "cuda"
        # Double-buffered staging of the batch as the user supplies it (fp32 NCHW, int64 labels): written by the copy stream, read
        # by the step.  On CPU there is one slot and plain copies.
        n_slots = 2 if cuda else 1
        self._staging = [(torch.zeros((batch_size, *input_shape), dtype=torch.float32, device=self.device),
                          torch.zeros((batch_size,), dtype=torch.int64, device=self.device)) for _ in range(n_slots)]
        self.static_x, self.static_y = self._staging[0]
        # A model whose first layer is a small-channel convolution (RGB input) reads a channels-last, one-16-byte-granule-per-pixel
        # activation (zero-bordered for the row-window form).  The layout kernel runs on the COPY stream right behind the staging copy,
        # one buffer per slot: it overlaps the previous step instead of being the first link of this step's kernel chain.
        self._smallc_in = False
        self._smallc_x = None
        if cuda:
            from . import nn as dnn
            from .ops import kernels as K
            first = next((m for m in model.modules() if isinstance(m, dnn.Conv2d)), None)
            if first is not None and first._smallc() and first.in_channels == input_shape[0]:
                C, H, W = input_shape
                rw = K.rowwin_geometry(H, W, *first.kernel_size, *first.stride, *first.padding, first.compute_dtype)
                self._smallc_x = [OF.smallc_input(batch_size, C, H, W, first.compute_dtype, self.device, rw, first.padding) for _ in range(2)]
                self.static_x = self._smallc_x[0]
                self._smallc_in = True
        self.metrics = torch.zeros(2, dtype=torch.float32, device=self.device)
        self.graphs = []
        self.step_count = 0
        self.kernels_per_step = 0       # launches of our kernels inside one captured step
        self._pending: Optional[int] = None
        self._local = getattr(optimizer, "step_local", None) or optimizer.step
        self._communicate = getattr(optimizer, "communicate", None)
        self._ring = None
        if cuda:
            self.stream = torch.cuda.Stream(device=self.device)
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._staged = [torch.cuda.Event() for _ in range(2)]
            self._consumed = [torch.cuda.Event() for _ in range(2)]
            for ev in self._consumed:
                ev.record(self.stream)
            self._done = [torch.cuda.Event() for _ in range(2)]        # only used with a foreign loss function
            self._pinned = [torch.zeros(2, dtype=torch.float32).pin_memory() for _ in range(2)]
            if self._own_loss:
                ring_host = torch.zeros((self.RING, 4), dtype=torch.float32).pin_memory()
                self._ring = (ring_host, torch.zeros(1, dtype=torch.int64, device=self.device))
                self._ring_base = 0          # launch number of step 0 (set after capture / warm-up)
            shadows = getattr(model, "_dmlb_shadows", None)
            if shadows is not None:
                shadows.prepare()
            if hasattr(optimizer, "fused_zero_grad") and type(optimizer).__name__ != "Synchronous":
                optimizer.zero_grad()                 # start from clean gradients, then let the update kernel keep them clean
                optimizer.fused_zero_grad = True
            if use_graph:
                self._capture(warmup_iters)

    @property
    def graph(self):
        """The captured step (slot 0), ``None`` when running eagerly."""
        return self.graphs[0] if self.graphs else None

    # ------------------------------------------------------------------------------------------
    def _body(self, slot: int = 0) -> None:
        sx, sy = self._staging[slot]
        self.optimizer.zero_grad()
        if self._smallc_in:
            out = self.model(self._smallc_x[slot])           # laid out (and cast) by __call__ on the copy stream
        else:
            out = self.model(sx)
        if self._own_loss:
            loss = self.loss_fn(out, sy, self.metrics, ring=self._ring)
        else:
            loss = self.loss_fn(out, sy)
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
                self._body(0)
            self.stream.synchronize()
            pool = None
            for slot in range(2):                      # one graph per staging slot; they share one memory pool (never run concurrently)
                g = torch.cuda.CUDAGraph()
                before = _native.launch_count()
                with torch.cuda.graph(g, stream=self.stream, pool=pool):
                    self._body(slot)
                self.kernels_per_step = _native.launch_count() - before
                pool = g.pool()
                self.graphs.append(g)
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
        if self._ring is not None:
            self._ring_base = int(self._ring[1].item())      # warm-up launches so far; the graphs continue from here

    # ------------------------------------------------------------------------------------------
    def _result(self, step: int, wait: bool) -> Optional[StepResult]:
        """Result of step ``step`` from the pinned ring (``wait``: spin until the loss kernel of that step has posted it)."""
        if self._ring is None:
            slot = step & 1
            self._done[slot].synchronize()
            return StepResult(float(self._pinned[slot][0]), float(self._pinned[slot][1]), step)
        n = self._ring_base + step
        row = self._ring[0][n % self.RING]
        want = float((n + 1) & 0xFFFFFF)
        if float(row[2]) != want:
            if not wait:
                return None
            import time
            t0 = time.monotonic()
            while float(row[2]) != want:
                if time.monotonic() - t0 > 120.0:
                    raise TimeoutError(f"TrainStep: the result of step {step} never arrived")
        return StepResult(float(row[0]), float(row[1]), step)

    def __call__(self, x: torch.Tensor, y: torch.Tensor, sync: bool = False) -> Optional[StepResult]:
        """Run one step on batch ``(x, y)`` (pinned host tensors or device tensors).

        Returns the :class:`StepResult` of the *previous* step (``None`` on the first call) unless
        ``sync=True``, in which case it waits for and returns this step's result.
        """
        if self.device.type != "cuda":
            sx, sy = self._staging[0]
            sx.copy_(x)
            sy.copy_(y)
            self.optimizer.zero_grad()
            out = self.model(sx)
            loss = self.loss_fn(out, sy)
            loss.backward()
            self.optimizer.step()
            self.step_count += 1
            correct = float((out.argmax(1) == sy).sum())
            return StepResult(float(loss), correct, self.step_count - 1)
        if hasattr(self.optimizer, "_sync_lr"):
            self.optimizer._sync_lr()
        if _NVTX:
            torch.cuda.nvtx.range_push(f"dmlb.step{self.step_count}")
        slot = self.step_count & 1
        sx, sy = self._staging[slot]
        # The batch travels to the staging slot on the copy stream (PCIe for a pinned host batch, a device-to-device move otherwise):
        # it overlaps the previous step, and the step stream sees nothing between two graph launches.
        self._copy_stream.wait_event(self._consumed[slot])         # the step that last read this slot has finished
        if x.is_cuda:
            # produced on the caller's current stream (e.g. utils.data.GpuResidentLoader): order the copy behind that work and keep
            # the caching allocator from recycling the batch while the copy stream still reads it
            cur = torch.cuda.current_stream(self.device)
            self._copy_stream.wait_stream(cur)
            x.record_stream(self._copy_stream)
            y.record_stream(self._copy_stream)
        with torch.cuda.stream(self._copy_stream):
            sx.copy_(x, non_blocking=True)
            sy.copy_(y, non_blocking=True)
            if self._smallc_in:
                OF.fill_smallc_input(self._smallc_x[slot], sx)
            self._staged[slot].record(self._copy_stream)
        self.stream.wait_event(self._staged[slot])
        with torch.cuda.stream(self.stream):
            if self.graphs:
                self.graphs[slot].replay()
            else:
                old_flag = OF.set_unit_loss_grad(True)
                self._body(slot)
                OF.set_unit_loss_grad(old_flag)
            self._consumed[slot].record(self.stream)
            if self._ring is None:
                self._pinned[slot].copy_(self.metrics, non_blocking=True)
                self._done[slot].record(self.stream)
            if self._communicate is not None:
                self._communicate()
        if _NVTX:
            torch.cuda.nvtx.range_pop()
        prev = self._pending
        self._pending = self.step_count
        self.step_count += 1
        if sync:
            return self.flush()
        if prev is None:
            return None
        return self._result(prev, wait=True)

    def flush(self) -> Optional[StepResult]:
        """Wait for the most recent step and return its result."""
        if self._pending is None:
            return None
        step, self._pending = self._pending, None
        return self._result(step, wait=True)

    def synchronize(self) -> None:
        if self.device.type == "cuda":
            self.stream.synchronize()
