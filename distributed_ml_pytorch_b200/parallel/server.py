"""Parameter server role.

Re-specifies the reference's ABSENT ``asgd.server.ParameterServer`` from its call sites (example/main.py:21,135-138;
contract in SURVEY.md §0.2 A3, §3.2): it holds the global flat parameter vector; ``ParameterUpdate`` sets it,
``GradientUpdate`` adds the (already ``-lr``-scaled) payload, ``ParameterRequest`` is answered with a
``ParameterUpdate`` to the sender.  ``run()`` is the blocking serve loop of the server rank.

Additions over the reference: a termination path (``WorkerDone`` from every worker ends ``run()`` — the reference's
server never exits), optional server-side momentum and staleness-scaled updates (NORTH-STAR), and
``snapshot()`` / ``save()`` to checkpoint the server's parameters as a model ``state_dict``.

On GPUs there is no server *process*: the server is memory — the flat vector is sharded over all GPUs of the box
and updated in place by NVLink reductions (``parallel/async_optim.py``, transport ``peer``).  A rank started with
``--server`` in that setting (:meth:`run_peer`) owns its shard and serves it without training.
"""
from __future__ import annotations

import logging
import time
from typing import Dict, Optional

import torch
import torch.distributed as dist

from ..utils import messaging as msg
from ..utils import trace
from ..utils.messaging import MessageCode, recv_message, send_message
from ..utils.serialization import ravel_model_params, unravel_model_params
from .arena import ShardMap
from .async_optim import staleness_scale

_LOGGER = logging.getLogger(__name__)


class WorkerLost(RuntimeError):
    """The server's idle timeout expired before every worker reported ``WorkerDone``."""


class ParameterServer:
    def __init__(self, model: torch.nn.Module, n_workers: Optional[int] = None, momentum: float = 0.0, staleness: Optional[str] = None,
                 staleness_param: float = 0.9, group=None, shard_index: int = 0, n_shards: int = 1, idle_timeout_s: Optional[float] = None,
                 gain=None):
        """``n_shards > 1``: the flat parameter vector is cut into ``n_shards`` contiguous pieces (``ShardMap``, the same cut the
        workers use) and this server owns piece ``shard_index``; every message payload is then that slice (NORTH-STAR sharded
        server, SURVEY.md §2.4 — the reference has exactly one server rank)."""
        self.model = model
        self.idle_timeout_s = idle_timeout_s       # run() raises WorkerLost when no message arrives for this long (None: wait forever)
        full = ravel_model_params(model).detach().clone().float().cpu()
        self.shards = ShardMap(full.numel(), n_shards, align=1)
        self.shard_index, self.n_shards = int(shard_index), int(n_shards)
        self.bounds = self.shards.valid_bounds(self.shard_index)
        self.parameter_shard = full[self.bounds[0]:self.bounds[1]].clone()
        self.momentum = float(momentum)
        self.velocity = torch.zeros_like(self.parameter_shard) if self.momentum else None
        self.staleness, self.staleness_param = staleness, staleness_param
        self.version = 0
        self.pulled_version: Dict[int, int] = {}
        self.group = group
        if n_workers is None and dist.is_available() and dist.is_initialized():
            n_workers = dist.get_world_size(group) - self.n_shards
        self.n_workers = n_workers
        from .async_optim import resolve_server_gain
        self.gain = resolve_server_gain(gain, self.momentum, n_workers or 1)     # "auto": (1 - mu) / workers, see resolve_server_gain
        self.done_workers = set()
        self.running = False
        self.counters = {"ParameterUpdate": 0, "GradientUpdate": 0, "ParameterRequest": 0}
        self.staleness_hist: Dict[int, int] = {}

    # ------------------------------------------------------------------------------------------
    def receive(self, sender: int, message_code: MessageCode, parameter: torch.Tensor) -> None:
        _LOGGER.info("Processing message: %s from %d", message_code.name, sender)
        if message_code == MessageCode.ParameterUpdate:
            self.parameter_shard.copy_(parameter[: self.parameter_shard.numel()])
            self.counters["ParameterUpdate"] += 1
        elif message_code == MessageCode.GradientUpdate:
            tau = self.version - self.pulled_version.get(sender, self.version)
            self.staleness_hist[tau] = self.staleness_hist.get(tau, 0) + 1
            s = self.gain * staleness_scale(self.staleness, self.staleness_param, tau)
            delta = parameter[: self.parameter_shard.numel()]
            if self.velocity is not None:
                self.velocity.mul_(self.momentum).add_(delta, alpha=s)
                self.parameter_shard.add_(self.velocity)
            else:
                self.parameter_shard.add_(delta, alpha=s)
            self.version += 1
            self.counters["GradientUpdate"] += 1
        elif message_code == MessageCode.ParameterRequest:
            self.pulled_version[sender] = self.version
            send_message(MessageCode.ParameterUpdate, self.parameter_shard, dst=sender, group=self.group)
            self.counters["ParameterRequest"] += 1
        elif message_code == MessageCode.WorkerDone:
            self.done_workers.add(sender)
            send_message(MessageCode.WorkerDone, None, dst=sender, group=self.group)   # lets the worker's listener exit

    def run(self) -> None:
        """Blocking serve loop (reference main.py:138); returns once every worker has sent ``WorkerDone``."""
        self.running = True
        seen = 0
        while self.running:
            try:
                # the idle clock starts with the first message: workers may need long to start (imports, data, CUDA context)
                sender, code, payload = recv_message(src=None, group=self.group, timeout_s=self.idle_timeout_s if seen else None)
                seen += 1
            except (TimeoutError, RuntimeError) as exc:
                # idle timeout, or the transport reported that a worker's connection went away
                self.running = False
                why = f"no message for {self.idle_timeout_s} s" if isinstance(exc, TimeoutError) else f"transport error: {str(exc)[:120]}"
                raise WorkerLost(f"parameter server: {why}; {len(self.done_workers)} of {self.n_workers} "
                                 f"workers had finished (done: {sorted(self.done_workers)})") from None
            with trace.span(code.name, "server", sender=sender, floats=int(payload.numel())):
                self.receive(sender, code, payload)
            if self.n_workers is not None and len(self.done_workers) >= self.n_workers:
                self.running = False
        msg.flush_sends()

    def stop(self) -> None:
        self.running = False

    # ------------------------------------------------------------------------------------------
    def snapshot(self) -> dict:
        """The server's parameters as a ``state_dict`` of ``model`` (keys identical to the reference's models).  A sharded server
        only knows its own slice: the rest of the vector keeps the values ``model`` was built with (see :meth:`merge_shards`)."""
        flat = ravel_model_params(self.model).detach().clone().float().cpu()
        flat[self.bounds[0]:self.bounds[1]] = self.parameter_shard
        unravel_model_params(self.model, flat)
        return {k: v.detach().cpu().clone() for k, v in self.model.state_dict().items()}

    def save(self, path: str) -> None:
        """One server: a plain ``state_dict``.  Sharded: this rank's slice, to ``<path>.shard<i>`` (combine with :meth:`merge_shards`)."""
        if self.n_shards == 1:
            torch.save(self.snapshot(), path)
        else:
            torch.save({"format": "dmlb-server-shard-v1", "shard_index": self.shard_index, "n_shards": self.n_shards, "bounds": self.bounds,
                        "parameter_shard": self.parameter_shard.clone(), "version": self.version}, f"{path}.shard{self.shard_index}")

    @staticmethod
    def merge_shards(paths, model: torch.nn.Module) -> dict:
        """Combine the ``.shard<i>`` files written by every server rank into one ``state_dict`` of ``model``."""
        flat = ravel_model_params(model).detach().clone().float().cpu()
        seen = set()
        for p in paths:
            part = torch.load(p, weights_only=False)
            b, e = part["bounds"]
            flat[b:e] = part["parameter_shard"]
            seen.add(int(part["shard_index"]))
            n = int(part["n_shards"])
        if seen != set(range(n)):
            raise ValueError(f"need the shards 0..{n - 1}, got {sorted(seen)}")
        unravel_model_params(model, flat)
        return {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}

    # ------------------------------------------------------------------------------------------
    @staticmethod
    def run_peer(model: torch.nn.Module, lr: float = 0.0, momentum: float = 0.0, staleness: Optional[str] = None, group=None,
                 poll_s: float = 0.0005, gain=None):
        """``--server`` rank on the NVLink transport: join the peer group as a shard owner that does not train.

        The rank allocates the same symmetric buffers as the workers (collective), then — for the mailbox protocol —
        keeps applying incoming updates to its shard until every worker has reached ``finish()``.
        """
        from .async_optim import Asynchronous
        opt = Asynchronous(model.parameters(), lr=lr, n_push=1, n_pull=1, model=model, momentum=momentum, staleness=staleness, group=group,
                           server_gain=gain)
        dev = opt.arena.params.device
        work = dist.barrier(group=group, async_op=True) if opt.peer.world > 1 else None
        while work is not None and not work.is_completed():
            if opt._use_mailbox:
                opt._owner_apply()
                torch.cuda.synchronize(dev)
            time.sleep(poll_s)
        if opt._use_mailbox:
            opt._owner_apply()
            torch.cuda.synchronize(dev)
            opt.peer.barrier()
        return opt
