#!/usr/bin/env python
"""Build libdmlb.so (all sm_100a kernels + native runtime) in-tree with nvcc.

    python build_native.py            # -> distributed_ml_pytorch_b200/_native/libdmlb.so
    python build_native.py --harness  # also builds tests/native/* standalone harnesses
    python build_native.py --sass     # also dumps artifacts/sass/*.sass (cuobjdump) for the record

nvcc cross-compiles for sm_100a without a GPU; objects are cached by source mtime so a rebuild after
touching one .cu takes seconds.  The .so is git-ignored but travels to the GPU box with the snapshot.
"""
from __future__ import annotations

import argparse
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
OUT_DIR = ROOT / "distributed_ml_pytorch_b200" / "_native"
BUILD_DIR = ROOT / "build" / "native"
LIB = OUT_DIR / "libdmlb.so"

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "--use_fast_math", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
          "-Xcompiler", "-Wno-unused-function", "-I", str(ROOT / "include"), "-I", str(CSRC),
          # every translation unit sees the public header first: a definition that drifts from include/dmlb.h is a build error
          "--pre-include", str(ROOT / "include" / "dmlb.h")]


def _sources():
    return sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cpp"))


def _stamp(src: Path) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for hdr in sorted(CSRC.glob("*.cuh")) + sorted((ROOT / "include").glob("*.h")):
        h.update(hdr.read_bytes())
    h.update(" ".join(ARCH_FLAGS + COMMON).encode())
    return h.hexdigest()


def _compile(src: Path, verbose: bool) -> Path:
    BUILD_DIR.mkdir(parents=True, exist_ok=True)
    obj = BUILD_DIR / (src.stem + ".o")
    stamp_file = BUILD_DIR / (src.stem + ".stamp")
    stamp = _stamp(src)
    if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return obj
    cmd = [NVCC, *ARCH_FLAGS, *COMMON, "-Xptxas", "-v" if verbose else "-warn-spills", "-x", "cu", "-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"nvcc failed for {src.name}")
    if verbose or "spill" in res.stderr and "0 bytes spill" not in res.stderr:
        (BUILD_DIR / (src.stem + ".ptxas.log")).write_text(res.stderr)
    if verbose:
        sys.stderr.write(res.stderr)
    stamp_file.write_text(stamp)
    return obj


def build(verbose: bool = False, harness: bool = False, sass: bool = False) -> Path:
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    srcs = _sources()
    if not srcs:
        raise RuntimeError("no sources under csrc/")
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if not LIB.exists() or LIB.stat().st_mtime < newest:
        cmd = [NVCC, *ARCH_FLAGS, "-shared", "-o", str(LIB), *map(str, objs), "-lcudart", "-lpthread", "-ldl"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("link failed")
    if harness:
        for h in sorted((ROOT / "tests" / "native").glob("*.cu")):
            exe = h.with_suffix("")
            cmd = [NVCC, *ARCH_FLAGS, "-O2", "-std=c++17", "-I", str(ROOT / "include"), str(h), "-o", str(exe),
                   "-L", str(OUT_DIR), "-ldmlb", "-Xlinker", f"-rpath={OUT_DIR}", "-Xlinker", "-rpath=$ORIGIN/../../distributed_ml_pytorch_b200/_native"]
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                sys.stderr.write(res.stdout + res.stderr)
                raise RuntimeError(f"harness build failed: {h.name}")
    if sass:
        dump_sass()
    return LIB


def dump_sass() -> None:
    """cuobjdump -sass per object -> artifacts/sass/<name>.sass (compact listings, capped in size), CENSUS.txt (mnemonic counts per
    object) and FUNCTIONS.txt (mnemonic counts per kernel: the proof of tcgen05 / TMA / TMA-im2col / NVLS multimem use)."""
    import re
    out = ROOT / "artifacts" / "sass"
    out.mkdir(parents=True, exist_ok=True)
    keys = ("UTCHMMA", "UTCHMMA.2CTA", "UTCBAR", "UTCBAR.2CTA.MULTICAST", "LDTM", "UTMALDG.2D", "UTMALDG.2D.2CTA", "UCGABAR_ARV", "UTMALDG.3D", "UTMALDG.4D.IM2COL", "UBLKCP.S.G", "UBLKCP.G.S", "UBLKRED.G.S.ADD.F32", "SYNCS", "REDG", "LDGMC", "STG.E.128", "LDGSTS")
    census, per_fn = [], []
    cap = 1_200_000
    for obj in sorted(BUILD_DIR.glob("*.o")):
        res = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True)
        if res.returncode != 0:
            continue
        text = res.stdout
        compact, size, skipped = [], 0, 0
        fn_name, fn_lines = None, []

        def flush():
            nonlocal size, skipped
            if fn_name is None:
                return
            body = "\n".join(fn_lines)
            per_fn.append(f"{obj.stem:14s} " + " ".join(f"{k}={body.count(k)}" for k in keys if body.count(k)) + f"  {fn_name}")
            if size + len(body) <= cap:
                compact.append(fn_name)
                compact.append(body)
                size += len(body)
            else:
                skipped += 1

        for ln in text.splitlines():
            if "Function :" in ln:
                flush()
                fn_name, fn_lines = ln.strip(), []
            elif "/*" in ln and fn_name is not None:
                parts = ln.split("*/")
                if len(parts) >= 2 and parts[1].strip():
                    fn_lines.append("    " + parts[1].split("/*")[0].strip())
        flush()
        if skipped:
            compact.append(f"... {skipped} further kernels of this object omitted from the listing (see FUNCTIONS.txt for their mnemonic census)")
        (out / (obj.stem + ".sass")).write_text("\n".join(compact) + "\n")
        row = {k: text.count(k) for k in keys}
        row["HMMA(legacy mma.sync)"] = len(re.findall(r"(?<![A-Z])HMMA", text))
        census.append(f"{obj.stem:16s} " + " ".join(f"{k}={v}" for k, v in row.items()))
    (out / "CENSUS.txt").write_text("\n".join(census) + "\n")
    (out / "FUNCTIONS.txt").write_text("\n".join(per_fn) + "\n")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--verbose", "-v", action="store_true")
    ap.add_argument("--harness", action="store_true")
    ap.add_argument("--sass", action="store_true")
    a = ap.parse_args()
    lib = build(a.verbose, a.harness, a.sass)
    print(lib)
