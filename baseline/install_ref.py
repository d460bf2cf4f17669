#!/usr/bin/env python
"""Install the UNMODIFIED reference into baseline/_ref (git-ignored) for ``bench.py --impl reference``.

1. Tries the prescribed offline ``pip install --target baseline/_ref`` from a /tmp copy of /root/reference.
   The reference has no setup.py / pyproject.toml, so pip cannot build it; the outcome is
   logged to baseline/_ref/INSTALL_LOG.txt (and quoted in DESIGN.md).
2. Falls back to a byte-identical copy of the reference tree and adds ONLY the modules that are absent from
   it (baseline/ref_completion/: asgd/__init__.py, asgd/utils/*, asgd/server.py, asgd/optim/downpour_sgd.py).
   Every shipped file keeps its sha256 (verified and logged).
"""
import hashlib
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
DEST = HERE / "_ref"
SRC = Path("/root/reference")
COMPLETION = HERE / "ref_completion"


def sha(p: Path) -> str:
    return hashlib.sha256(p.read_bytes()).hexdigest()


def main() -> int:
    if not SRC.exists():
        print("reference tree not present on this machine; keeping existing baseline/_ref", file=sys.stderr)
        return 0 if DEST.exists() else 1
    if DEST.exists():
        shutil.rmtree(DEST)
    DEST.mkdir(parents=True)
    log = []
    with tempfile.TemporaryDirectory() as tmp:
        work = Path(tmp) / "reference"
        shutil.copytree(SRC, work)
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links", "/opt/wheelhouse",
               "--target", str(DEST), str(work)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        log.append("$ " + " ".join(cmd))
        log.append(f"exit code {res.returncode}")
        log.append((res.stdout + res.stderr).strip()[-1500:])
        if res.returncode != 0:
            res2 = subprocess.run(cmd[:-1] + ["--no-deps", str(work)], capture_output=True, text=True)
            log.append("retry with --no-deps: exit code %d" % res2.returncode)
            log.append((res2.stdout + res2.stderr).strip()[-600:])
    log.append("")
    log.append("pip cannot install the reference (no setup.py/pyproject.toml): copying the tree byte-for-byte instead.")
    for f in sorted(SRC.rglob("*")):
        if f.is_file():
            out = DEST / f.relative_to(SRC)
            out.parent.mkdir(parents=True, exist_ok=True)
            shutil.copy2(f, out)
    added = []
    for f in sorted(COMPLETION.rglob("*.py")):
        out = DEST / f.relative_to(COMPLETION)
        if out.exists():
            raise SystemExit(f"completion would overwrite a shipped reference file: {out}")
        out.parent.mkdir(parents=True, exist_ok=True)
        shutil.copy2(f, out)
        added.append(str(f.relative_to(COMPLETION)))
    log.append("shipped files (sha256 identical to /root/reference):")
    for f in sorted(SRC.rglob("*")):
        if f.is_file():
            a, b = sha(f), sha(DEST / f.relative_to(SRC))
            assert a == b, f
            log.append(f"  {a[:16]}  {f.relative_to(SRC)}")
    log.append("added (ABSENT from the reference, SURVEY.md §0.2): " + ", ".join(added))
    (DEST / "INSTALL_LOG.txt").write_text("\n".join(log) + "\n")
    print("\n".join(log))
    return 0


if __name__ == "__main__":
    sys.exit(main())
