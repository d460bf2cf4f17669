"""Asynchronous (Downpour-style) SGD — the worker-side optimizer of the parameter-server scheme.

API kept from the reference (asgd/optim/Asynchronous.py:20-71): ``Asynchronous(params, lr, n_push, n_pull, model)``
is a ``torch.optim.Optimizer``; every ``step()`` applies plain SGD locally, accumulates ``-lr * grad`` and, every
``n_push`` / ``n_pull`` steps, pushes the accumulated update to the server and pulls fresh parameters
(protocol: SURVEY.md §3.7).  ``DownpourSGD`` is the same class (the name the reference's ``__init__`` exports).

Three transports, chosen from where the model lives:

``peer``  (CUDA model)  — every GPU is a worker *and* owns one shard of the parameter server.  Push =
          ``red.global.add.v4.f32`` straight into the owners' master shards over NVLink (K1), pull = peer loads into the
          local arena (K3); with ``momentum`` / ``staleness`` the push goes through per-worker mailboxes and the owner
          applies ``v = mu v + s(tau) delta; theta += v`` (K2).  No server process, no request messages, no host copies.
``gloo``  (CPU model, process group initialised) — the reference's topology: rank 0 runs
          :class:`~distributed_ml_pytorch_b200.parallel.server.ParameterServer`, workers exchange messages
          (``utils/messaging.py``).  This is BASELINE.json config 1 and the semantic oracle.
``local`` (no process group) — a loop-back server inside the process; with one worker and ``n_push = n_pull = 1``
          the trajectory equals plain SGD.

Differences from the reference, all deliberate: ``Optimizer.__init__`` runs first; the push
precedes the pull inside a step so pulled parameters already contain the worker's own update; pulled
parameters are applied at a step boundary instead of racing with forward/backward (``torn_pulls=True`` restores
the reference's behaviour on the gloo transport); a pull request carries no payload; only ``init_rank``
uploads the initial parameters.

Overlap on the peer transport (``overlap=True``, the default): pushes, owner-side applies and pulls run on a communication stream
while the worker keeps training, and the worker keeps NO accumulator.  What a worker has not pushed yet is, by definition, the
difference between its replica and the replica as of its last push, so the worker keeps a *snapshot* instead: a push sends
``params - snapshot`` (exact in fp32: the same variable at two times) and advances the snapshot to the value it read — element by
element, so it may run while training keeps writing ``params`` and every local update is still sent exactly once.  The training step
is then the plain fused SGD kernel (no accumulator read + write: 2 x 4 bytes per parameter per step saved, and nothing to switch inside
a captured CUDA graph).  A pull is fetched into a staging buffer; ``pull_lag`` steps later one kernel on the step stream lands it,
``params = pulled + (params - snapshot); snapshot = pulled`` — the server's parameters plus exactly the local updates the server has
not seen.  ``accumulated_gradients`` (the reference's attribute) is computed on demand in this mode.
"""
from __future__ import annotations

import logging
import threading
from typing import Optional

import torch
import torch.distributed as dist

from ..utils import messaging as msg
from ..utils import trace
from ..utils.messaging import MessageCode, MessageListener, send_message
from ..utils.serialization import ravel_model_params, unravel_model_params
from .arena import FlatArena
from .fused_sgd import ArenaOptimizer

_LOGGER = logging.getLogger(__name__)
_STALENESS = {None: 0, "none": 0, "inverse": 1, "exp": 2}


class Listener(MessageListener):
    """Receives ``ParameterUpdate`` replies from the server (reference Asynchronous.py:9-18)."""

    def __init__(self, model, owner: "Asynchronous", src: Optional[int] = msg.SERVER_RANK, done_expected: int = 1):
        super().__init__(model, src=src, done_expected=done_expected)
        self.owner = owner

    def receive(self, sender, message_code, parameter):
        _LOGGER.info("Processing message: %s", message_code.name)
        trace.instant("reply " + message_code.name, "comm", sender=sender)
        if message_code == MessageCode.ParameterUpdate:
            o = self.owner
            shard = o.server_ranks.index(sender) if sender in o.server_ranks else 0
            if o.torn_pulls:
                o._write_shard(shard, parameter)                  # reference behaviour: overwrite immediately
            with o._lock:
                if not o.torn_pulls:
                    o._pending_slices[shard] = parameter.clone()
                o._replies += 1
                complete = o._replies >= len(o.server_ranks)
            if complete:
                o._reply_event.set()


class Asynchronous(ArenaOptimizer):
    def __init__(self, params, lr, n_push: int, n_pull: int, model: torch.nn.Module, momentum: float = 0.0,
                 staleness: Optional[str] = None, staleness_param: float = 0.9, transport: str = "auto", group=None,
                 init_rank: Optional[int] = None, torn_pulls: bool = False, server_rank: int = msg.SERVER_RANK,
                 blocking_pull: bool = False, pull_timeout_s: float = 60.0, server_ranks=None, overlap: bool = True, pull_lag: int = 2,
                 comm_blocks: int = 24, server_gain=None):
        """``params, lr, n_push, n_pull, model`` are the reference's arguments (Asynchronous.py:20-31).  Additions:

        ``momentum``, ``staleness`` (``None | "inverse" | "exp"``), ``staleness_param``: server-side update ``v = mu v + gain s(tau) delta``.
        ``server_gain``: ``None`` (1), ``"auto"`` ((1 - momentum) / workers) or a number — see :func:`resolve_server_gain`.
        ``transport``: ``"auto" | "peer" | "gloo" | "local"``; ``group``: process group of the workers.
        ``overlap`` (peer): run pushes / pulls on a communication stream, keep a snapshot instead of an accumulator (default).
        ``pull_lag`` (peer, overlap): steps between issuing a pull and landing it on the replica; ``comm_blocks``: CTAs of an overlapped
        push / pull kernel.  ``init_rank``, ``server_rank(s)``, ``torn_pulls``, ``blocking_pull``, ``pull_timeout_s``: gloo transport."""
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if n_push < 1 or n_pull < 1:
            raise ValueError("n_push and n_pull must be >= 1")
        if staleness not in _STALENESS:
            raise ValueError(f"staleness must be one of {list(_STALENESS)}")
        device = next(model.parameters()).device
        if transport == "auto":
            if device.type == "cuda":
                transport = "peer"
            elif dist.is_available() and dist.is_initialized():
                transport = "gloo"
            else:
                transport = "local"
        if transport == "gloo" and device.type == "cuda":
            # the CUDA arena keeps conv weights channels-last, the message protocol speaks model.parameters() order
            raise ValueError("a CUDA model uses the NVLink 'peer' transport; the gloo message transport is for CPU models")
        self.transport = transport
        self.n_push, self.n_pull = int(n_push), int(n_pull)
        self.server_momentum = float(momentum)
        self.staleness, self.staleness_param = staleness, float(staleness_param)
        self._server_gain_arg = server_gain        # resolved once the number of workers is known (peer / local transports)
        self.server_gain = 1.0
        self.torn_pulls = torn_pulls
        # gloo transport only: wait for the server's reply to every pull request and apply it before the next step.  The reference
        # never waits (replies land whenever the listener thread gets them); this switch makes a run reproducible for debugging.
        self.blocking_pull, self.pull_timeout_s = bool(blocking_pull), float(pull_timeout_s)
        self._reply_event = threading.Event()
        self.server_rank = server_rank
        # gloo transport, NORTH-STAR sharded server: several server ranks, each owning one contiguous slice of the flat vector
        self.server_ranks = [int(r) for r in server_ranks] if server_ranks else [int(server_rank)]
        self._pending_slices = {}
        self._replies = 0
        self.group = group
        self.idx = 0
        self._lock = threading.Lock()
        self._listener: Optional[Listener] = None
        self.pushes = self.pulls = 0
        self.overlap = bool(overlap) and transport == "peer"
        self.pull_lag = max(0, int(pull_lag))
        self.comm_blocks = int(comm_blocks)    # CTAs of an overlapped push / pull kernel (a background trickle next to the training step)

        arena = None
        if transport == "peer":
            from . import peer
            self.peer = peer.PeerGroup(group=group, device=device)
            self._chunk = peer.chunk_elems()
            arena = FlatArena(model, n_shards=self.peer.world, align=self._chunk)   # shards are whole mailbox chunks
        super().__init__(params, dict(lr=lr), model, arena=arena)
        a = self.arena
        self._snapshot = None
        self._acc_buf = None if self.overlap else torch.zeros(a.padded, dtype=torch.float32, device=a.params.device)

        n_workers = self.peer.world if transport == "peer" else 1
        self.server_gain = resolve_server_gain(server_gain, self.server_momentum, n_workers)
        if transport == "peer":
            self._init_peer()
        elif transport == "gloo":
            self.rank = dist.get_rank()
            from .arena import ShardMap
            self._srv_shards = ShardMap(a.numel, len(self.server_ranks), align=1)
            if init_rank is None:
                init_rank = min(r for r in range(dist.get_world_size(group)) if r not in self.server_ranks)
            if self.rank == init_rank:
                # this sets the initial model parameters on the server (reference Asynchronous.py:33-34)
                flat = ravel_model_params(model)
                for k, r in enumerate(self.server_ranks):
                    b, e = self._srv_shards.valid_bounds(k)
                    send_message(MessageCode.ParameterUpdate, flat[b:e], dst=r)
            single = len(self.server_ranks) == 1
            self._listener = Listener(model, self, src=self.server_ranks[0] if single else None, done_expected=len(self.server_ranks))
            self._listener.start()
        else:
            self._master = a.params.clone()
            self._master_mom = torch.zeros_like(self._master) if self.server_momentum else None
            self._master_version = 0
            self._pulled_version = 0

    # ------------------------------------------------------------------------------------------
    # peer transport set-up
    # ------------------------------------------------------------------------------------------
    def _init_peer(self) -> None:
        from . import peer
        pg, a = self.peer, self.arena
        S = a.shards.shard_size
        cps = S // self._chunk
        self._cps = cps
        self._use_mailbox = bool(self.server_momentum) or _STALENESS[self.staleness] != 0
        self.master_sym = pg.symmetric(S * 4)
        self.version_sym = pg.symmetric(cps * 4)
        self.master = self.master_sym.view(torch.float32, S)
        self.version = self.version_sym.view(torch.int32, cps)
        b, e = a.shards.bounds(pg.rank)
        # One initial model for everybody (the reference: one worker's vector sets the server, Asynchronous.py:33-34): rank 0's
        # replica is broadcast before the owners seed their shards, so differently seeded ranks do not start from a patchwork.
        _broadcast_flat(a.params, pg.group, pg.world)
        self.master.copy_(a.params[b:e])                       # owner's (now common) replica defines its shard
        self.pulled_ver = torch.zeros(pg.world * cps, dtype=torch.int32, device=a.params.device)
        self.stats = torch.zeros(4, dtype=torch.int32, device=a.params.device)
        if self._use_mailbox:
            self.mailbox_sym = pg.symmetric(pg.world * S * 4)
            self.full_sym = pg.symmetric(pg.world * cps * 4)
            self.busy_sym = pg.symmetric(pg.world * cps * 4)
            self.mailbox = self.mailbox_sym.view(torch.float32, pg.world * S)
            self.full = self.full_sym.view(torch.int32, pg.world * cps)
            self.busy = self.busy_sym.view(torch.int32, pg.world * cps)
            self.master_mom = torch.zeros(S, dtype=torch.float32, device=a.params.device) if self.server_momentum else None
        if self.overlap:
            dev = a.params.device
            self._pulled = torch.empty_like(a.params)
            self._pending_pull = None
            self.comm_stream = torch.cuda.Stream(device=dev)
        pg.barrier()
        self._pull()                                           # every replica starts from the server's parameters
        if self.overlap:
            self._snapshot = a.params.clone()                  # nothing unpushed yet
        torch.cuda.synchronize(a.params.device)
        pg.barrier()

    @property
    def accumulated_gradients(self) -> torch.Tensor:
        """The update this worker has not pushed yet, ``sum(-lr * grad)`` since the last push (reference Asynchronous.py:27,55).  In
        overlap mode it is not stored: it is the replica minus its snapshot (a fresh tensor, ordered behind outstanding pushes)."""
        if self._snapshot is None:
            return self._acc_buf
        a = self.arena
        torch.cuda.current_stream(a.params.device).wait_stream(self.comm_stream)
        return a.params - self._snapshot

    def _push(self) -> None:
        from . import peer
        pg, a = self.peer, self.arena
        S = a.shards.shard_size
        if self.overlap and self._use_mailbox:
            peer.push_mailbox_delta(a.params, self._snapshot, self.busy, self.pulled_ver, self.mailbox_sym.table, self.full_sym.table, pg.rank,
                                    pg.world, S)
        elif self.overlap:
            peer.push_delta(a.params, self._snapshot, self.master_sym.table, pg.world, S, a.padded)
        elif self._use_mailbox:
            peer.push_mailbox(self._acc_buf, self.busy, self.pulled_ver, self.mailbox_sym.table, self.full_sym.table, pg.rank, pg.world, S)
        else:
            peer.push_add(self._acc_buf, self.master_sym.table, pg.world, S, a.padded)
        self.pushes += 1

    def _owner_apply(self) -> None:
        from . import peer
        pg, a = self.peer, self.arena
        peer.owner_apply(self.master, self.master_mom, self.version, self.full, self.mailbox, self.busy_sym.table, self.stats, pg.rank, pg.world,
                         a.shards.shard_size, self.server_momentum, _STALENESS[self.staleness], self.staleness_param, self.server_gain)

    def _pull(self) -> None:
        from . import peer
        pg, a = self.peer, self.arena
        if self._snapshot is not None:      # delta scheme: fetch, then re-base the replica (keeps what has not been pushed)
            from ..ops import kernels as K
            peer.pull(self._pulled, self.master_sym.table, pg.world, a.shards.shard_size, a.padded, self.version_sym.table, self.pulled_ver)
            K.land_pull(a.params, self._pulled, self._snapshot, a.shadow, a.padded)
        else:
            peer.pull(a.params, self.master_sym.table, pg.world, a.shards.shard_size, a.padded, self.version_sym.table, self.pulled_ver,
                      shadow=a.shadow)
        a.mark_params_changed(flat_done=a.shadow is not None)
        self.pulls += 1

    def set_unpushed(self, update: torch.Tensor) -> None:
        """Restore the not-yet-pushed update (``accumulated_gradients``), e.g. from a checkpoint."""
        if self._snapshot is None:
            if update is not self._acc_buf:
                self._acc_buf.copy_(update)
        else:
            # the model was just loaded with the SERVER's parameters (that is what a checkpoint of this transport holds): the replica
            # the run was stopped with is that plus the update it had not pushed yet
            a = self.arena
            torch.cuda.current_stream(a.params.device).wait_stream(self.comm_stream)
            self._snapshot.copy_(a.params)
            a.params.add_(update)
            a.mark_params_changed()

    def _land_pull(self, force: bool = False) -> None:
        """Overlap mode: make a fetched pull the model's parameters (on the current stream, at a step boundary)."""
        pp = self._pending_pull
        if pp is None or (not force and self.idx - pp[0] < self.pull_lag):
            return
        from ..ops import kernels as K
        a = self.arena
        cur = torch.cuda.current_stream(a.params.device)
        # behind the fetch AND behind every push issued since (with n_push < pull_lag one may still be advancing the snapshot this kernel
        # rewrites); both were queued steps ago, so in practice nothing waits
        cur.wait_stream(self.comm_stream)
        # the server's parameters + the local updates it has not seen (replica - snapshot): what a pull that had landed the moment it
        # was issued, followed by the local steps since, would have produced
        K.land_pull(a.params, self._pulled, self._snapshot, a.shadow, a.padded)
        a.mark_params_changed(flat_done=a.shadow is not None)
        self._pending_pull = None

    def _communicate_overlapped(self, push_now: bool, pull_now: bool) -> None:
        from . import peer
        pg, a = self.peer, self.arena
        cur = torch.cuda.current_stream(a.params.device)
        self._land_pull()
        if not (push_now or pull_now or self._use_mailbox):
            return
        if pull_now:
            self._land_pull(force=True)                        # the staging buffer is about to be overwritten
        ready = torch.cuda.Event()
        ready.record(cur)
        self.comm_stream.wait_event(ready)                     # behind this step's update and behind a landing (both touch the snapshot)
        peer.set_comm_max_blocks(self.comm_blocks)             # background transfer: leave the SMs to the training step
        with torch.cuda.stream(self.comm_stream):
            if push_now:
                self._push()
            if self._use_mailbox:
                self._owner_apply()
            if pull_now:
                peer.pull(self._pulled, self.master_sym.table, pg.world, a.shards.shard_size, a.padded, self.version_sym.table, self.pulled_ver)
                fetched = torch.cuda.Event()
                fetched.record(self.comm_stream)
                self._pending_pull = (self.idx, fetched)
                self.pulls += 1
        peer.set_comm_max_blocks(0)
        if pull_now and self.pull_lag == 0:
            self._land_pull(force=True)

    # ------------------------------------------------------------------------------------------
    # the optimizer step
    # ------------------------------------------------------------------------------------------
    def capture_state(self):
        """State a dry run of ``step_local`` changes besides the parameters (``engine.TrainStep`` warms up and captures with it)."""
        return (self._acc_buf.clone() if self._acc_buf is not None else None, self.idx)

    def restore_state(self, st) -> None:
        if st[0] is not None:
            self._acc_buf.copy_(st[0])        # overlap mode: the unpushed update is params - snapshot; restoring params restores it
        self.idx = st[1]

    def drain(self) -> None:
        """Overlap mode: wait (stream-wise) for outstanding pushes and land a fetched pull, e.g. before evaluating or checkpointing."""
        if not self.overlap:
            return
        torch.cuda.current_stream(self.arena.params.device).wait_stream(self.comm_stream)
        self._land_pull(force=True)

    @torch.no_grad()
    def step_local(self) -> None:
        """Local SGD on the replica + accumulate ``-lr * grad`` (reference Asynchronous.py:52-55,63-68): one kernel."""
        a = self.arena
        self._sync_lr()
        if a.params.is_cuda:
            from ..ops import kernels as K
            table = a.fused_shadow_table()
            K.sgd_flat_ex(a.params, a.grads, self._acc_buf, None, a.padded, self.lr_dev, 0.0, 0.0, shadow=a.shadow,
                          zero_grads=self.fused_zero_grad, table=table)
            a.mark_params_changed(flat_done=a.shadow is not None, table_done=table is not None)
        else:
            lr = self.lr
            self._acc_buf.add_(a.grads, alpha=-lr)
            a.params.add_(a.grads, alpha=-lr)

    @torch.no_grad()
    def communicate(self) -> None:
        """Push every ``n_push`` steps, pull every ``n_pull`` steps (both fire at step 0, like the reference)."""
        push_now = self.idx % self.n_push == 0
        pull_now = self.idx % self.n_pull == 0
        if self.transport == "peer" and self.overlap:
            self._communicate_overlapped(push_now, pull_now)
        elif self.transport == "peer":
            if push_now:
                self._push()
            if self._use_mailbox:
                self._owner_apply()
            if pull_now:
                self._pull()
        elif self.transport == "gloo":
            if push_now:
                with trace.span("push", "comm", step=self.idx, floats=self.arena.numel):
                    for k, r in enumerate(self.server_ranks):
                        b, e = self._srv_shards.valid_bounds(k)
                        send_message(MessageCode.GradientUpdate, self._acc_buf[b:e], dst=r)
                    self._acc_buf.zero_()
                self.pushes += 1
            if pull_now:
                with trace.span("pull request", "comm", step=self.idx, blocking=self.blocking_pull):
                    with self._lock:
                        self._replies = 0
                    self._reply_event.clear()
                    for r in self.server_ranks:
                        send_message(MessageCode.ParameterRequest, None, dst=r)
                    self.pulls += 1
                    if self.blocking_pull:
                        if not self._reply_event.wait(self.pull_timeout_s):
                            raise TimeoutError(f"no ParameterUpdate from the server within {self.pull_timeout_s} s")
                        self._apply_pending()
        else:
            if push_now:
                self._local_server_apply(self._acc_buf)
                self._acc_buf.zero_()
                self.pushes += 1
            if pull_now:
                self.arena.params.copy_(self._master)
                self._pulled_version = self._master_version
                self.arena.mark_params_changed()
                self.pulls += 1
        self.idx += 1

    def _local_server_apply(self, delta: torch.Tensor) -> None:
        tau = self._master_version - self._pulled_version
        s = self.server_gain * staleness_scale(self.staleness, self.staleness_param, tau)
        if self._master_mom is not None:
            self._master_mom.mul_(self.server_momentum).add_(delta, alpha=s)
            self._master.add_(self._master_mom)
        else:
            self._master.add_(delta, alpha=s)
        self._master_version += 1

    def _write_shard(self, shard: int, values: torch.Tensor) -> None:
        """Overwrite the slice of the model owned by server ``shard`` with ``values`` (one server: the whole model)."""
        if len(self.server_ranks) == 1:
            unravel_model_params(self.model, values)
            return
        b, e = self._srv_shards.valid_bounds(shard)
        flat = ravel_model_params(self.model).detach().clone()
        flat[b:e] = values[: e - b]
        unravel_model_params(self.model, flat)

    def _apply_pending(self) -> None:
        if not self._pending_slices:
            return
        with self._lock:
            slices, self._pending_slices = self._pending_slices, {}
        for shard, p in sorted(slices.items()):
            self._write_shard(shard, p)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self._apply_pending()
        self.step_local()
        self.communicate()
        return loss

    # ------------------------------------------------------------------------------------------
    def finish(self) -> None:
        """End of training: flush outstanding traffic, tell the server, release peer memory."""
        if self.transport == "gloo":
            for r in self.server_ranks:
                send_message(MessageCode.WorkerDone, None, dst=r)
            msg.flush_sends()
            if self._listener is not None:
                self._listener.join(timeout=30)
            self._apply_pending()        # the server answers every request before it echoes WorkerDone: end on its latest parameters
        elif self.transport == "peer":
            if self.peer is None:
                return
            self.drain()
            torch.cuda.synchronize(self.arena.params.device)
            self.peer.barrier()
            if self._use_mailbox:
                self._owner_apply()
                torch.cuda.synchronize(self.arena.params.device)
                self.peer.barrier()

    def close(self) -> None:
        """Release the NVLink-shared buffers (masters, mailboxes, flags) and their CUDA-IPC mappings.  Collective: every rank of the
        group must call it, after :meth:`finish`.  The model and its arena stay usable; this optimizer does not."""
        if self.transport != "peer" or self.peer is None:
            return
        for name in ("master", "version", "mailbox", "full", "busy"):
            if hasattr(self, name):
                setattr(self, name, None)
        for name in ("master_sym", "version_sym", "mailbox_sym", "full_sym", "busy_sym"):
            if hasattr(self, name):
                setattr(self, name, None)
        self.peer.close()
        self.peer = None

    def server_parameters(self) -> torch.Tensor:
        """The server's current flat parameter vector (length N) as seen from this worker."""
        a = self.arena
        if self.transport == "peer":
            if self.overlap:      # pushes still in flight on the communication stream belong to the answer
                torch.cuda.current_stream(a.params.device).wait_stream(self.comm_stream)
            out = torch.empty_like(a.params)
            from . import peer
            peer.pull(out, self.master_sym.table, self.peer.world, a.shards.shard_size, a.padded)
            return out[: a.numel]
        if self.transport == "local":
            return self._master[: a.numel].clone()
        raise RuntimeError("on the gloo transport the parameters live on the server rank")

    def staleness_stats(self) -> dict:
        if self.transport != "peer":
            return {}
        s = self.stats.tolist()
        return {"chunks_applied": s[0], "mean_staleness": (s[1] / s[0] if s[0] else 0.0), "max_staleness": s[2]}


def _broadcast_flat(t: torch.Tensor, group, world: int, src: int = 0) -> None:
    """Broadcast a flat device tensor from group rank ``src`` over the control-plane process group (one-off, at set-up)."""
    if world <= 1 or not (dist.is_available() and dist.is_initialized()):
        return
    gsrc = dist.get_global_rank(group, src) if group is not None else src
    try:
        dist.broadcast(t, src=gsrc, group=group)
    except RuntimeError:                      # a control plane without CUDA support (gloo): stage through the host
        h = t.detach().cpu()
        dist.broadcast(h, src=gsrc, group=group)
        t.copy_(h)
    torch.cuda.synchronize(t.device)


def resolve_server_gain(gain, momentum: float, n_workers: int) -> float:
    """Step-size gain of the server update ``v = mu v + gain * s(tau) * delta; theta += v``.

    ``None`` -> 1 (the classic momentum update).  ``"auto"`` -> ``(1 - mu) / n_workers``: with ``W`` workers feeding one momentum
    buffer the steady-state step per unit of gradient is ``W * lr / (1 - mu)`` — 80x the nominal learning rate for 8 workers and
    ``mu = 0.9``, which is what made ResNet-50 with server momentum climb to loss 10 in round 1; ``auto`` brings it back to ``lr``.
    (Without momentum the gain stays 1: summing the workers' updates is the reference's own semantics, Asynchronous.py:58-60.)"""
    if gain is None:
        return 1.0
    if gain == "auto":
        return (1.0 - float(momentum)) / max(1, int(n_workers)) if momentum else 1.0
    return float(gain)


def staleness_scale(mode: Optional[str], param: float, tau: int) -> float:
    """s(tau): 'inverse' -> 1/(1+tau), 'exp' -> param**tau, None/'none' -> 1 (the reference: no scaling)."""
    if mode == "inverse":
        return 1.0 / (1.0 + tau)
    if mode == "exp":
        return float(param) ** tau
    return 1.0


DownpourSGD = Asynchronous
