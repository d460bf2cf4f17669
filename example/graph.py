#!/usr/bin/env python
"""``make graph`` (reference Makefile:9-11; the script itself is ABSENT from the reference): turn log/*.csv into
train_time.png and test_time.png."""
import glob
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from distributed_ml_pytorch_b200.utils.plot import graph  # noqa: E402

if __name__ == "__main__":
    paths = sys.argv[1:] or sorted(glob.glob("log/*.csv"))
    if not paths:
        raise SystemExit("no CSV logs found (run example/main.py first; it writes log/{single,gpu,node<rank>}.csv)")
    train_png, test_png = graph(paths, out_dir=".")
    print("wrote", train_png, test_png)
