"""Host-side tracing in the Chrome / Perfetto trace-event format (the reference records one ``datetime.now()`` per
iteration, example/main.py:77, and nothing else).

    from distributed_ml_pytorch_b200.utils import trace
    trace.start()                                  # or: example/main.py --trace out.json
    with trace.span("step", "train", iteration=i): ...
    trace.save("out.json")                         # open in chrome://tracing or ui.perfetto.dev

Spans are wall-clock intervals of the calling thread (``perf_counter_ns``).  On the CUDA path they therefore show when work was
*enqueued* (the step is one graph launch); device time comes from CUDA events (``utils/timing.py``, ``bench.py``) and NVTX ranges
(``DMLB_NVTX=1``).  On the CPU / gloo path they are the real thing: forward/backward, push, pull, server-side message handling.
When no tracer is active ``span()`` is a no-op context manager, so library code can call it unconditionally.
"""
from __future__ import annotations

import contextlib
import json
import os
import threading
import time
from typing import Any, Dict, List, Optional


class Tracer:
    def __init__(self, process_name: Optional[str] = None):
        self._events: List[Dict[str, Any]] = []
        self._lock = threading.Lock()
        self._t0 = time.perf_counter_ns()
        self.pid = os.getpid()
        self.process_name = process_name

    def _now_us(self) -> float:
        return (time.perf_counter_ns() - self._t0) / 1e3

    @contextlib.contextmanager
    def span(self, name: str, cat: str = "", **args):
        ts = self._now_us()
        try:
            yield
        finally:
            ev = {"name": name, "cat": cat or "dmlb", "ph": "X", "ts": ts, "dur": self._now_us() - ts, "pid": self.pid,
                  "tid": threading.get_ident() & 0x7FFFFFFF}
            if args:
                ev["args"] = args
            with self._lock:
                self._events.append(ev)

    def instant(self, name: str, cat: str = "", **args) -> None:
        ev = {"name": name, "cat": cat or "dmlb", "ph": "i", "s": "t", "ts": self._now_us(), "pid": self.pid,
              "tid": threading.get_ident() & 0x7FFFFFFF}
        if args:
            ev["args"] = args
        with self._lock:
            self._events.append(ev)

    def events(self) -> List[Dict[str, Any]]:
        with self._lock:
            return list(self._events)

    def save(self, path: str) -> str:
        meta = []
        if self.process_name:
            meta.append({"name": "process_name", "ph": "M", "pid": self.pid, "args": {"name": self.process_name}})
        with open(path, "w") as f:
            json.dump({"traceEvents": meta + self.events(), "displayTimeUnit": "ms"}, f)
        return path


_current: Optional[Tracer] = None


def start(process_name: Optional[str] = None) -> Tracer:
    """Install a process-wide tracer (library spans start recording)."""
    global _current
    _current = Tracer(process_name)
    return _current


def stop() -> Optional[Tracer]:
    global _current
    t, _current = _current, None
    return t


def current() -> Optional[Tracer]:
    return _current


@contextlib.contextmanager
def _noop():
    yield


def span(name: str, cat: str = "", **args):
    t = _current
    return t.span(name, cat, **args) if t is not None else _noop()


def instant(name: str, cat: str = "", **args) -> None:
    t = _current
    if t is not None:
        t.instant(name, cat, **args)


def save(path: str) -> Optional[str]:
    t = _current
    return t.save(path) if t is not None else None
