"""CPU unit tests of the auxiliary subsystems: CSV log schema, plotting, clock sampling, timing helpers, messaging codes."""
import os

import torch

from distributed_ml_pytorch_b200.utils import csvlog, plot
from distributed_ml_pytorch_b200.utils.messaging import MessageCode
from distributed_ml_pytorch_b200.utils.timing import ClockSampler


def test_csv_schema_and_paths(tmp_path):
    assert csvlog.log_path(True, False, None) == os.path.join("log", "single.csv")
    assert csvlog.log_path(True, True, None) == os.path.join("log", "gpu.csv")
    assert csvlog.log_path(False, True, 2) == os.path.join("log", "node2.csv")
    lg = csvlog.CSVLogger(str(tmp_path / "log" / "node1.csv"))       # the directory does not exist yet (reference needs it to)
    lg.append(0, 2.30)
    lg.append(100, 2.10, test_loss=2.2, test_accuracy=0.31, images_per_sec=1234.5)
    rows = csvlog.read_csv(lg.to_csv())
    assert list(rows[0].keys()) == csvlog.COLUMNS
    assert rows[0]["test_loss"] == "" and rows[1]["test_accuracy"] == "0.31" and rows[1]["index"] == "1"


def test_plot_writes_valid_png_without_matplotlib(tmp_path):
    out = plot.plot_series(str(tmp_path / "p.png"), {"a": ([0, 1, 2], [3.0, 2.0, 1.0]), "b": ([0, 1, 2], [1.0, 1.5, 2.5])})
    data = open(out, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n" and len(data) > 200
    lg = csvlog.CSVLogger(str(tmp_path / "log" / "single.csv"))
    for i in range(5):
        lg.append(i, 2.0 - 0.1 * i, test_loss=2.0, test_accuracy=0.1 * i)
    lg.to_csv()
    a, b = plot.graph([lg.path], out_dir=str(tmp_path))
    assert os.path.basename(a) == "train_time.png" and os.path.basename(b) == "test_time.png"      # reference Makefile:9-11


def test_clock_sampler_summary_parsing():
    s = ClockSampler()
    s.lines = ["0, 1965, 1965, 310.5, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active",
               "0, 1830, 1965, 990.1, 0x0000000000000004, Not Active, Not Active, Not Active, Active",
               "0, 1845, 1965, 1001.7, 0x0000000000000004, Not Active, Not Active, Not Active, Active", "garbage"]
    out = s.summary()
    assert out["sm_max_mhz"] == 1965.0 and out["reasons"] == ["sw_power_cap"] and out["samples"] == 3
    assert out["sm_mhz"] in (1830.0, 1837.5, 1845.0)          # median of the samples taken under load
    assert ClockSampler().summary()["samples"] == 0


def test_message_codes_cover_the_reference_protocol():
    names = {c.name for c in MessageCode}
    assert {"ParameterRequest", "GradientUpdate", "ParameterUpdate"} <= names       # reference Asynchronous.py:17,34,49,59
    assert MessageCode(MessageCode.GradientUpdate.value) is MessageCode.GradientUpdate


def test_resident_loader_matches_totensor_normalize_on_cpu():
    from distributed_ml_pytorch_b200.utils.data import GpuResidentLoader
    g = torch.Generator().manual_seed(0)
    imgs = torch.randint(0, 256, (50, 32, 32, 3), generator=g, dtype=torch.uint8)
    labels = torch.randint(0, 10, (50,), generator=g)
    ld = GpuResidentLoader(imgs.numpy(), labels.numpy(), batch_size=16, device="cpu", shuffle=False)
    assert len(ld) == 3
    x, y = next(iter(ld))
    ref = (imgs[:16].permute(0, 3, 1, 2).float() / 255 - 0.5) / 0.5       # Normalize((.5,.5,.5),(.5,.5,.5))(ToTensor(img))
    assert torch.allclose(x, ref, atol=1e-6) and torch.equal(y, labels[:16])
    seen = torch.cat([yy for _, yy in GpuResidentLoader(imgs, labels, 10, "cpu", shuffle=True, seed=3)])
    assert seen.numel() == 50


def _fake_cifar(root, n=40):
    import os, pickle
    import numpy as np
    base = os.path.join(root, "cifar-10-batches-py")
    os.makedirs(base)
    rng = np.random.default_rng(0)
    for name in [f"data_batch_{i}" for i in range(1, 6)] + ["test_batch"]:
        with open(os.path.join(base, name), "wb") as f:
            pickle.dump({"data": rng.integers(0, 256, (n, 3072), dtype=np.uint8), "labels": rng.integers(0, 10, n).tolist()}, f)
    return base


def test_cifar10_arrays_reads_torchvision_layout(tmp_path):
    from distributed_ml_pytorch_b200.utils.data import cifar10_arrays
    _fake_cifar(str(tmp_path))
    x, y = cifar10_arrays(str(tmp_path), train=True)
    assert x.shape == (200, 32, 32, 3) and x.dtype.name == "uint8" and y.shape == (200,)
    xt, yt = cifar10_arrays(str(tmp_path), train=False)
    assert xt.shape == (40, 32, 32, 3) and yt.shape == (40,)


def test_reference_import_paths_resolve():
    """Every import the reference's own files use (example/main.py:19-21, asgd/optim/__init__.py:1, Asynchronous.py:4-5) works here."""
    from asgd.optim import Asynchronous, DownpourSGD
    from asgd.optim.Asynchronous import Asynchronous as A2, Listener
    from asgd.optim.downpour_sgd import DownpourSGD as D2
    from asgd.server import ParameterServer
    from asgd.utils.messaging import MessageCode, MessageListener, send_message
    from asgd.utils.serialization import ravel_model_params, unravel_model_params
    import asgd.optim
    assert A2 is Asynchronous and D2 is DownpourSGD is Asynchronous
    assert asgd.optim.Asynchronous is Asynchronous            # the class, not a shadowing submodule
    assert issubclass(Listener, MessageListener)
    assert {c.name for c in MessageCode} >= {"ParameterRequest", "GradientUpdate", "ParameterUpdate"}
    assert callable(send_message) and callable(ravel_model_params) and callable(unravel_model_params) and ParameterServer is not None


def test_c_abi_header_matches_the_built_library():
    """include/dmlb.h declares exactly the dmlb_* functions libdmlb.so exports, and every entry point named in INTEGRATION.md exists."""
    import os
    import re
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = os.path.join(root, "distributed_ml_pytorch_b200", "_native", "libdmlb.so")
    if not os.path.exists(lib) or shutil.which("nm") is None:
        import pytest
        pytest.skip("library not built here")
    out = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T dmlb_" in ln}
    header = open(os.path.join(root, "include", "dmlb.h")).read()
    declared = set(re.findall(r"\b(dmlb_[a-z0-9_]+)\s*\(", header))
    assert exported == declared, (sorted(exported - declared), sorted(declared - exported))
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    named = set(re.findall(r"`(dmlb_[a-z0-9_]+)`", doc))             # plain names only; brace patterns like dmlb_relu_{fwd,bwd} are skipped
    assert named <= exported, sorted(named - exported)


def test_ctypes_signatures_match_the_header():
    """Every ctypes prototype in _native._SIGNATURES has the arity and the argument classes (pointer / int / long long / float) of the
    declaration in include/dmlb.h — a c_int where the C side reads a long long would silently corrupt the call."""
    import ctypes
    import os
    import re
    from distributed_ml_pytorch_b200 import _native
    from distributed_ml_pytorch_b200.parallel import peer  # noqa: F401  (registers the peer-memory / collective entry points)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = re.sub(r"/\*.*?\*/", "", open(os.path.join(root, "include", "dmlb.h")).read(), flags=re.S)

    def c_class(text):
        if "*" in text:
            return "ptr"
        for key, name in (("long long", "i64"), ("float", "float"), ("double", "double"), ("size_t", "i64"), ("int", "int")):
            if key in text:
                return name
        return text

    def ctypes_class(c):
        if c in (ctypes.c_void_p, ctypes.c_char_p) or (isinstance(c, type) and issubclass(c, ctypes._Pointer)):
            return "ptr"
        if c in (ctypes.c_float, ctypes.c_double):
            return "float" if c is ctypes.c_float else "double"
        return {4: "int", 8: "i64"}.get(ctypes.sizeof(c), str(c))      # integer classes by width (c_size_t aliases c_ulong here)

    declared = {}
    for name, args in re.findall(r"\b(dmlb_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", header, flags=re.S):
        params = [a for a in (x.strip() for x in args.split(",")) if a and a != "void"]
        declared[name] = [c_class(p) for p in params]
    checked = 0
    for name, proto in _native._SIGNATURES.items():
        assert name in declared, f"{name} is not declared in include/dmlb.h"
        assert [ctypes_class(c) for c in proto] == declared[name], name
        checked += 1
    assert checked >= 45 and "dmlb_sync_nvls_range" in _native._SIGNATURES
