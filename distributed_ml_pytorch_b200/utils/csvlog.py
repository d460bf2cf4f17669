"""Per-iteration training log with the reference's CSV schema.

The reference appends a dict per iteration (``timestamp``, ``iteration``, ``training_loss`` and, every ``log_interval``
iterations, ``test_loss`` / ``test_accuracy``) and dumps it through pandas to ``log/{single,gpu,node<rank>}.csv`` with
``index_label='index'`` (example/main.py:76-91,97-105).  Same columns and file names here (plus ``images_per_sec``),
without requiring pandas and without requiring ``log/`` to pre-exist.
"""
from __future__ import annotations

import csv
import os
from datetime import datetime
from typing import Dict, List, Optional

COLUMNS = ["index", "timestamp", "iteration", "training_loss", "test_loss", "test_accuracy", "images_per_sec"]


class CSVLogger:
    def __init__(self, path: str):
        self.path = path
        self.rows: List[Dict] = []

    def append(self, iteration: int, training_loss: float, test_loss: Optional[float] = None, test_accuracy: Optional[float] = None,
               images_per_sec: Optional[float] = None, timestamp: Optional[datetime] = None) -> Dict:
        row = {"timestamp": timestamp or datetime.now(), "iteration": iteration, "training_loss": training_loss}
        if test_loss is not None:
            row["test_loss"] = test_loss
            row["test_accuracy"] = test_accuracy
        if images_per_sec is not None:
            row["images_per_sec"] = images_per_sec
        self.rows.append(row)
        return row

    def to_csv(self) -> str:
        os.makedirs(os.path.dirname(os.path.abspath(self.path)) or ".", exist_ok=True)
        with open(self.path, "w", newline="") as f:
            w = csv.DictWriter(f, fieldnames=COLUMNS)
            w.writeheader()
            for i, r in enumerate(self.rows):
                out = {"index": i}
                out.update({k: ("" if v is None else v) for k, v in r.items()})
                w.writerow(out)
        return self.path


def log_path(no_distributed: bool, cuda: bool, rank: Optional[int], log_dir: str = "log") -> str:
    """File naming of the reference: log/single.csv, log/gpu.csv, log/node<rank>.csv (main.py:99-105)."""
    if no_distributed:
        return os.path.join(log_dir, "gpu.csv" if cuda else "single.csv")
    return os.path.join(log_dir, f"node{rank}.csv")


def read_csv(path: str) -> List[Dict[str, str]]:
    with open(path, newline="") as f:
        return list(csv.DictReader(f))
