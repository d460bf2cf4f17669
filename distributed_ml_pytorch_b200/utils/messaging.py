"""Parameter-server messaging for the CPU / ``gloo`` transport.

Re-specifies the reference's ABSENT ``asgd.utils.messaging`` from its call sites
(reference asgd/optim/Asynchronous.py:5,9-18,34,37-38,49,59; protocol in SURVEY.md §3.7):

* :class:`MessageCode` — ``ParameterRequest``, ``GradientUpdate``, ``ParameterUpdate`` (+ ``WorkerDone``,
  which gives the server the termination path the reference lacks);
* :func:`send_message` ``(code, payload, dst=server)`` — non-blocking from the caller's view;
* :class:`MessageListener` — a thread calling ``receive(sender, code, payload)`` for every message.

Wire format: a 3-element int64 header ``[sender, code, payload_len]`` followed (when ``payload_len > 0``)
by the flat fp32 payload.  Unlike the reference, a ``ParameterRequest`` carries **no** payload
(the reference uploads N dummy floats per pull, Asynchronous.py:49).

On GPUs none of this is used for data: pushes and pulls are fused kernels over NVLink peer memory
(:mod:`distributed_ml_pytorch_b200.parallel.peer`).  This module remains the control plane and the
semantic oracle.
"""
from __future__ import annotations

import enum
import logging
import threading
import time
from datetime import timedelta
from typing import Optional

import torch
import torch.distributed as dist

_LOGGER = logging.getLogger(__name__)

SERVER_RANK = 0
_HEADER_TAG = 0
_PAYLOAD_TAG = 1
_send_lock = threading.Lock()
_pending = []  # keeps (work, tensors) alive until the isend completes


class MessageCode(enum.Enum):
    """Message kinds of the Downpour protocol."""
    ParameterRequest = 0
    GradientUpdate = 1
    ParameterUpdate = 2
    WorkerDone = 3          # not in the reference: lets ParameterServer.run() return
    EvaluateParams = 4      # kept for wire compatibility with the upstream lineage (unused)


_MAX_INFLIGHT = 16     # outstanding isend work items (header + payload per message) before the sender waits for the oldest


def _reap_pending() -> None:
    """Bound the send queue.  gloo only marks a send ``is_completed()`` once somebody has waited on it, so polling the flag never
    frees anything; waiting on the oldest entries (delivered long ago, the peer receives in order) returns at once and releases
    their payload snapshots, and acts as back-pressure if the peer has really fallen behind."""
    while len(_pending) >= _MAX_INFLIGHT:
        work, _ = _pending.pop(0)
        work.wait()


def send_message(message_code: MessageCode, payload: Optional[torch.Tensor] = None, dst: int = SERVER_RANK,
                 group=None, blocking: bool = False) -> None:
    """Send ``(code, payload)`` to ``dst`` (default: the server, rank 0 — reference main.py:151).

    The payload is snapshotted before the call returns, so the caller may immediately overwrite it —
    the reference zeroes its accumulator right after sending (Asynchronous.py:59-60), which would race
    with a true asynchronous send of the same storage.
    """
    if payload is None:
        payload = torch.zeros(0)
    flat = payload.detach().reshape(-1).to(device="cpu", dtype=torch.float32, copy=True)
    header = torch.tensor([dist.get_rank(), message_code.value, flat.numel()], dtype=torch.int64)
    with _send_lock:
        _reap_pending()
        if blocking:
            dist.send(header, dst=dst, group=group, tag=_HEADER_TAG)
            if flat.numel():
                dist.send(flat, dst=dst, group=group, tag=_PAYLOAD_TAG)
            return
        w = dist.isend(header, dst=dst, group=group, tag=_HEADER_TAG)
        _pending.append((w, header))
        if flat.numel():
            w2 = dist.isend(flat, dst=dst, group=group, tag=_PAYLOAD_TAG)
            _pending.append((w2, flat))


def flush_sends() -> None:
    """Block until every outstanding :func:`send_message` has left this process."""
    with _send_lock:
        for work, _ in _pending:
            work.wait()
        _pending.clear()


def recv_message(src: Optional[int] = None, group=None, timeout_s: Optional[float] = None):
    """Blocking receive of one message. Returns ``(sender, MessageCode, payload)``.

    ``timeout_s``: give up with ``TimeoutError`` when no message header arrives in that time (failure detection: the
    reference's server blocks forever when its workers die)."""
    header = torch.zeros(3, dtype=torch.int64)
    if timeout_s is None:
        dist.recv(header, src=src, group=group, tag=_HEADER_TAG)
    else:
        # gloo closes the pair when a timed wait expires, so after the TimeoutError the group is only good for shutting down
        work = dist.irecv(header, src=src, group=group, tag=_HEADER_TAG)
        try:
            work.wait(timedelta(seconds=float(timeout_s)))
        except RuntimeError as exc:
            if "Timed out" in str(exc) or "timeout" in str(exc).lower():
                raise TimeoutError(f"no message within {timeout_s} s") from None
            raise
    sender, code, n = int(header[0]), int(header[1]), int(header[2])
    payload = torch.zeros(n, dtype=torch.float32)
    if n:
        dist.recv(payload, src=sender, group=group, tag=_PAYLOAD_TAG)
    return sender, MessageCode(code), payload


class MessageListener(threading.Thread):
    """Thread that receives messages from ``src`` and hands them to :meth:`receive`.

    Mirrors the base class the reference's ``Listener`` derives from (Asynchronous.py:9-18): the
    constructor takes the model, ``start()`` spawns the thread, subclasses override ``receive``.
    ``stop()`` ends the loop once the peer sends ``WorkerDone`` (or any message after the flag is set).
    """

    def __init__(self, model, src: Optional[int] = SERVER_RANK, group=None, done_expected: int = 1):
        super().__init__(daemon=True)
        self.model = model
        self.src = src                        # None = any source (several server ranks)
        self.group = group
        self.running = True
        self.done_expected = int(done_expected)   # WorkerDone echoes to see before the loop ends (one per server rank)
        self._done_seen = 0

    def receive(self, sender: int, message_code: MessageCode, parameter: torch.Tensor) -> None:
        raise NotImplementedError

    def stop(self) -> None:
        self.running = False

    def run(self) -> None:
        while self.running:
            try:
                sender, code, payload = recv_message(src=self.src, group=self.group)
            except Exception as exc:  # process group torn down while blocked in recv
                if self.running:
                    _LOGGER.debug("listener exiting: %s", exc)
                return
            if code == MessageCode.WorkerDone:
                self._done_seen += 1
                if self._done_seen >= self.done_expected:
                    self.running = False
                    return
                continue
            self.receive(sender, code, payload)
