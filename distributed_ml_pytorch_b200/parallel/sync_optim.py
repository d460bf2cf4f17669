"""Synchronous data-parallel SGD (NORTH-STAR; the reference has no synchronous mode).

Baseline formulation: ``all_reduce(grads) / world`` followed by ``optim.SGD.step()``.  Here, on GPUs, one kernel
(``sync_rs_update_ag``, csrc/comm.cu K4) does barrier -> reduce-scatter of the gradient arena over NVLink peer loads ->
SGD / momentum on the owner's fp32 slice -> all-gather of the new weights by peer stores into every replica -> barrier.
The parameter and gradient arenas of every replica live in CUDA-IPC-shared memory for that purpose.

Overlap with the backward pass: the flat vector is cut into ``n_buckets`` contiguous buckets; the ops' backward functions
announce when the kernel that finalises a parameter's gradient has been launched (``ops.functional.add_grad_ready_hook``),
and as soon as every parameter of a bucket is announced the fused kernel for that bucket is queued on a communication
stream behind those kernels — backward runs last layer first, so the last bucket's reduce-scatter / update / all-gather
travels over NVLink while the earlier layers are still differentiating.  Inside one bucket every GPU owns 1/world of it.
All of this is capturable: in :class:`~distributed_ml_pytorch_b200.engine.TrainStep` the bucket kernels are parallel
branches of the step's CUDA graph.

NVLS: when the arenas can be mapped behind an NVSwitch multicast address (``nvls=None`` = use it where it was measured faster: :func:`nvls_pays`), the
reduce-scatter is one ``multimem.ld_reduce`` per 16 bytes (summed inside the switch) and the all-gather one ``multimem.st``.

On CPU tensors (gloo) the same optimizer falls back to ``dist.all_reduce`` + an in-place update.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .arena import FlatArena
from .fused_sgd import ArenaOptimizer


class _Bucket:
    __slots__ = ("begin", "end", "params", "pending", "events", "launched")

    def __init__(self, begin: int, end: int):
        self.begin, self.end = begin, end
        self.params: List[int] = []
        self.pending = 0
        self.events = []
        self.launched = False


class Synchronous(ArenaOptimizer):
    def __init__(self, params, lr: float, model: torch.nn.Module, momentum: float = 0.0, weight_decay: float = 0.0, group=None,
                 overlap: bool = True, n_buckets: int = 4, comm_blocks: int = 48, nvls: Optional[bool] = None):
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        device = next(model.parameters()).device
        self.group = group
        self.is_peer = device.type == "cuda"
        self.overlap = bool(overlap) and self.is_peer
        self.comm_blocks = comm_blocks
        arena = None
        if self.is_peer:
            from . import peer
            self.peer = peer.PeerGroup(group=group, device=device)
            n = sum(p.numel() for p in model.parameters())
            world = self.peer.world
            shard = ((n + world - 1) // world + 255) // 256 * 256
            padded = shard * world
            # Symmetric arenas.  With more than one rank they come from torch's symmetric-memory allocator (plumbing only), which
            # also maps them behind an NVLS multicast address when the NVSwitch fabric offers one; then the reduce-scatter and
            # the all-gather run inside the switch (multimem.ld_reduce / multimem.st).  Otherwise: CUDA-IPC buffers, peer loads/stores.
            self.nvls = False
            self.params_sym = self.grads_sym = self.flags_sym = None
            if nvls is None and not nvls_pays(world, padded * 4):
                nvls = False                     # measured: the peer-load form is faster here (see nvls_pays)
            if world > 1 and nvls is not False:
                try:
                    self.params_sym = peer.SymmMemBuffer(padded * 4, device, group)
                    self.grads_sym = peer.SymmMemBuffer(padded * 4, device, group)
                    self.flags_sym = peer.SymmMemBuffer(peer.flag_words() * 4, device, group)
                    self.nvls = bool(self.params_sym.mc_ptr and self.grads_sym.mc_ptr)
                except Exception as exc:          # allocator or fabric without symmetric-memory support
                    if nvls:
                        raise
                    self.params_sym = self.grads_sym = self.flags_sym = None
            if nvls and not self.nvls:
                raise RuntimeError("NVLS multicast was requested but is not available on this fabric")
            if self.params_sym is None:
                self.params_sym = self.peer.symmetric(padded * 4)
                self.grads_sym = self.peer.symmetric(padded * 4)
                self.flags_sym = self.peer.symmetric(peer.flag_words() * 4)
            arena = FlatArena(model, n_shards=world, params_storage=self.params_sym.view(torch.float32, padded),
                              grads_storage=self.grads_sym.view(torch.float32, padded))
            assert arena.shards.shard_size == shard
            self.world = world
        else:
            self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay), model, arena=arena)
        a = self.arena
        self.momentum_buf = torch.zeros_like(a.params) if momentum else None      # indexed like the arena
        self._buckets: List[_Bucket] = []
        self._by_param: Dict[int, List[_Bucket]] = {}
        self._armed = False
        if self.is_peer:
            self.comm_stream = torch.cuda.Stream(device=device)
            self._make_buckets(max(1, int(n_buckets)) if self.overlap else 1)
            if self.overlap:
                from ..ops import functional as OF
                OF.add_grad_ready_hook(self._on_grad_ready)
            self.peer.barrier()
        self.broadcast_parameters()

    # ------------------------------------------------------------------------------------------
    def _make_buckets(self, k: int) -> None:
        a = self.arena
        unit = 256 * self.world
        units = a.padded // unit
        k = min(k, units)
        bounds = [round(i * units / k) * unit for i in range(k + 1)]
        self._buckets = [_Bucket(bounds[i], bounds[i + 1]) for i in range(k) if bounds[i + 1] > bounds[i]]
        self._by_param = {}
        for name, p in self.model.named_parameters():
            off, n = a.offsets[name]
            for b in self._buckets:
                if off < b.end and off + n > b.begin:
                    b.params.append(id(p))
                    self._by_param.setdefault(id(p), []).append(b)

    def _arm(self) -> None:
        for b in self._buckets:
            b.pending = len(b.params)
            b.events = []
            b.launched = False
        self._armed = True

    def _on_grad_ready(self, param, stream) -> None:
        if not self._armed:
            return
        bs = self._by_param.get(id(param))
        if not bs:
            return
        ev = torch.cuda.Event()
        ev.record(stream)
        for b in bs:
            if b.launched:
                continue
            b.events.append(ev)
            b.pending -= 1
            if b.pending == 0:
                self._launch_bucket(b)

    def _launch_bucket(self, b: _Bucket, after_stream: Optional[torch.cuda.Stream] = None) -> None:
        from . import peer
        g = self.param_groups[0]
        for ev in b.events:
            self.comm_stream.wait_event(ev)
        if after_stream is not None:
            self.comm_stream.wait_stream(after_stream)
        n = (b.end - b.begin) // self.world
        base = b.begin + self.peer.rank * n
        mom_ptr = self.momentum_buf.data_ptr() + 4 * base if self.momentum_buf is not None else 0
        blocks = self.comm_blocks if self.overlap else 0
        with torch.cuda.stream(self.comm_stream):
            if self.nvls:
                peer.sync_nvls_range(self.grads_sym.mc_ptr, self.params_sym.mc_ptr, self.arena.params, self.flags_sym.table, mom_ptr,
                                     self.peer.rank, self.world, base, n, self.lr_dev, g["momentum"], g["weight_decay"], 1.0 / self.world,
                                     max_blocks=blocks)
            else:
                peer.sync_rs_update_ag_range(self.grads_sym.table, self.params_sym.table, self.flags_sym.table, mom_ptr, self.peer.rank,
                                             self.world, base, n, self.lr_dev, g["momentum"], g["weight_decay"], 1.0 / self.world,
                                             max_blocks=blocks)
        b.launched = True

    # ------------------------------------------------------------------------------------------
    def broadcast_parameters(self, src: int = 0) -> None:
        """Make every replica start from rank ``src``'s parameters."""
        if self.world == 1 or not (dist.is_available() and dist.is_initialized()):
            return
        dist.broadcast(self.arena.params, src=src, group=self.group)
        self.arena.mark_params_changed()
        if self.is_peer:
            self.peer.barrier()

    def zero_grad(self, set_to_none: bool = False) -> None:
        super().zero_grad(set_to_none)
        if self.overlap:
            self._arm()

    def capture_state(self):
        return self.momentum_buf.clone() if self.momentum_buf is not None else None

    def restore_state(self, st) -> None:
        if st is not None:
            self.momentum_buf.copy_(st)

    @torch.no_grad()
    def step_local(self) -> None:
        """Finish the step: queue the buckets that were not announced during backward, join the communication stream and
        refresh the bf16 weight mirrors.  (Everything here is stream work, so TrainStep captures it in the step's graph.)"""
        if not self.is_peer:
            return
        self._sync_lr()
        main = torch.cuda.current_stream(self.arena.params.device)
        from ..ops import functional as OF
        OF.join_side_streams()
        if not self._armed:
            self._arm()
        for b in reversed(self._buckets):
            if not b.launched:
                self._launch_bucket(b, after_stream=main)
        self._armed = False
        main.wait_stream(self.comm_stream)
        self.arena.mark_params_changed()

    @torch.no_grad()
    def communicate(self) -> None:
        """CPU / gloo transport only (the GPU work is in :meth:`step_local`)."""
        if self.is_peer:
            return
        g = self.param_groups[0]
        a = self.arena
        if self.world > 1:
            dist.all_reduce(a.grads, group=self.group)
            a.grads.div_(self.world)
        d = a.grads
        if g["weight_decay"]:
            d = d.add(a.params, alpha=g["weight_decay"])
        if self.momentum_buf is not None:
            self.momentum_buf.mul_(g["momentum"]).add_(d)
            d = self.momentum_buf
        a.params.add_(d, alpha=-g["lr"])

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self.step_local()
        self.communicate()
        return loss

    def finish(self) -> None:
        if self.is_peer and self.peer is not None:
            if self.overlap:
                from ..ops import functional as OF
                OF.remove_grad_ready_hook(self._on_grad_ready)
            self.peer.barrier()

    def close(self) -> None:
        """Release the shared arenas.  The model's parameters live IN them, so they are first moved to a private arena (values kept);
        then the CUDA-IPC / symmetric-memory buffers are unmapped and freed.  Collective, call after :meth:`finish`."""
        if not self.is_peer or self.peer is None:
            return
        torch.cuda.synchronize(self.arena.params.device)
        self.arena = FlatArena(self.model, n_shards=1)          # copies the parameters out and re-points them
        self.params_sym = self.grads_sym = self.flags_sym = None
        self.peer.close()
        self.peer = None


def nvls_pays(world: int, nbytes: int) -> bool:
    """Should the synchronous step reduce / broadcast inside the NVSwitch (``multimem``) or with peer loads / stores?  From the measured
    sweeps (profiles/comm_sweep_2gpu_nvls.json, r2_comm_sweep_4gpu.json, comm_sweep_8gpu_v3_nvls.json): at 2 GPUs the peer form is never
    slower and wins from 4 MB up (64 MB: 120 vs 182 us); at 4 GPUs the switch is 5-10 % ahead up to 4 MB and behind at 16-64 MB
    (56 vs 65 us); at 8 GPUs it wins from 1 KB (18 vs 23 us) to 256 MB (604 vs 776 us) and loses at 1 GB (3.43 vs 3.08 ms).
    ``Synchronous(nvls=True / False)`` overrides."""
    if world >= 8:
        return nbytes <= (512 << 20)
    return world >= 4 and nbytes <= (8 << 20)
