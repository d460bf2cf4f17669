"""``make graph``: CSV logs -> ``train_time.png`` / ``test_time.png`` (reference Makefile:9-11 runs the ABSENT
``example/graph.py``).  Uses matplotlib when it is installed and a tiny built-in PNG line plotter
otherwise (this image has no matplotlib and no network)."""
from __future__ import annotations

import struct
import zlib
from datetime import datetime
from typing import Dict, List, Sequence, Tuple

import numpy as np

from .csvlog import read_csv

_PALETTE = [(31, 119, 180), (255, 127, 14), (44, 160, 44), (214, 39, 40), (148, 103, 189), (140, 86, 75), (227, 119, 194), (127, 127, 127)]


def _png(path: str, img: np.ndarray) -> None:
    h, w, _ = img.shape
    raw = b"".join(b"\x00" + img[y].tobytes() for y in range(h))

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw, 6)) +
                chunk(b"IEND", b""))


def _line(img: np.ndarray, x0: int, y0: int, x1: int, y1: int, color) -> None:
    n = max(abs(x1 - x0), abs(y1 - y0), 1)
    xs = np.linspace(x0, x1, n + 1).round().astype(int)
    ys = np.linspace(y0, y1, n + 1).round().astype(int)
    ok = (xs >= 0) & (xs < img.shape[1]) & (ys >= 0) & (ys < img.shape[0])
    img[ys[ok], xs[ok]] = color


def plot_series(path: str, series: Dict[str, Tuple[Sequence[float], Sequence[float]]], width: int = 800, height: int = 480) -> str:
    """Line plot of ``{label: (xs, ys)}``; returns ``path``."""
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        fig, ax = plt.subplots(figsize=(width / 100, height / 100))
        for label, (xs, ys) in series.items():
            ax.plot(xs, ys, label=label)
        ax.set_xlabel("seconds")
        ax.legend()
        fig.savefig(path, dpi=100)
        plt.close(fig)
        return path
    except ImportError:
        pass
    img = np.full((height, width, 3), 255, dtype=np.uint8)
    m = 40
    allx = [x for xs, _ in series.values() for x in xs] or [0.0, 1.0]
    ally = [y for _, ys in series.values() for y in ys] or [0.0, 1.0]
    x_lo, x_hi, y_lo, y_hi = min(allx), max(allx), min(ally), max(ally)
    x_hi = x_hi if x_hi > x_lo else x_lo + 1.0
    y_hi = y_hi if y_hi > y_lo else y_lo + 1.0
    _line(img, m, height - m, width - m, height - m, (0, 0, 0))
    _line(img, m, m, m, height - m, (0, 0, 0))
    for i, (label, (xs, ys)) in enumerate(series.items()):
        color = _PALETTE[i % len(_PALETTE)]
        px = [m + int((x - x_lo) / (x_hi - x_lo) * (width - 2 * m)) for x in xs]
        py = [height - m - int((y - y_lo) / (y_hi - y_lo) * (height - 2 * m)) for y in ys]
        for j in range(1, len(px)):
            _line(img, px[j - 1], py[j - 1], px[j], py[j], color)
        img[m + 12 * i: m + 12 * i + 8, width - m - 60: width - m - 20] = color   # legend swatch, in series order
    _png(path, img)
    return path


def _seconds(rows: List[Dict[str, str]]) -> List[float]:
    ts = [datetime.fromisoformat(r["timestamp"]) for r in rows]
    return [(t - ts[0]).total_seconds() for t in ts]


def graph(csv_paths: Sequence[str], out_dir: str = ".") -> Tuple[str, str]:
    """Training loss vs. time and test accuracy vs. time for every log file (one line per file)."""
    import os
    train, test = {}, {}
    for p in csv_paths:
        rows = read_csv(p)
        if not rows:
            continue
        secs = _seconds(rows)
        label = os.path.splitext(os.path.basename(p))[0]
        train[label] = (secs, [float(r["training_loss"]) for r in rows])
        pts = [(s, float(r["test_accuracy"])) for s, r in zip(secs, rows) if r.get("test_accuracy") not in (None, "")]
        if pts:
            test[label] = ([a for a, _ in pts], [b for _, b in pts])
    os.makedirs(out_dir, exist_ok=True)
    return (plot_series(os.path.join(out_dir, "train_time.png"), train), plot_series(os.path.join(out_dir, "test_time.png"), test))
