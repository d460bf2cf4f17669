"""The reference's entry points keep their names, flags and roles (SURVEY.md §2.1 C4-C12, §5 flag table)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.timeout(300)


def _run(cmd, cwd, timeout=240):
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.run(cmd, cwd=cwd, env=env, capture_output=True, text=True, timeout=timeout)


def test_reference_flags_and_defaults_are_kept():
    sys.path.insert(0, os.path.join(ROOT, "example"))
    import importlib
    main = importlib.import_module("main")
    a = main.build_parser().parse_args([])
    assert (a.batch_size, a.test_batch_size, a.epochs, a.lr, a.num_pull, a.num_push) == (64, 10000, 20, 0.008, 10, 10)
    assert (a.cuda, a.log_interval, a.no_distributed, a.rank, a.world_size, a.server, a.master, a.port) == \
        (False, 100, False, None, 3, False, "localhost", "29500")
    assert a.model == "alexnet"       # the reference hard-codes AlexNet (main.py:41)


def test_single_process_run_writes_reference_csv(tmp_path):
    r = _run([sys.executable, os.path.join(ROOT, "example", "main.py"), "--no-distributed", "--model", "lenet", "--epochs", "1",
              "--steps-per-epoch", "12", "--log-interval", "5", "--test-batch-size", "50", "--synthetic"], cwd=tmp_path)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Finished Training" in r.stdout
    from distributed_ml_pytorch_b200.utils.csvlog import read_csv
    rows = read_csv(os.path.join(tmp_path, "log", "single.csv"))
    assert len(rows) == 12
    assert set(["index", "timestamp", "iteration", "training_loss", "test_loss", "test_accuracy"]).issubset(rows[0].keys())
    assert rows[5]["test_loss"] != "" and rows[4]["test_loss"] == ""
    g = _run([sys.executable, os.path.join(ROOT, "example", "graph.py")], cwd=tmp_path)
    assert g.returncode == 0, g.stderr
    assert os.path.getsize(os.path.join(tmp_path, "train_time.png")) > 100


def test_server_and_two_workers_via_launcher(tmp_path):
    """`make server` + `make first` + `make second` (reference Makefile:13-20), started by run-pytorch.py --local."""
    from distributed_ml_pytorch_b200.parallel.launcher import free_port
    r = _run([sys.executable, os.path.join(ROOT, "run-pytorch.py"), "--local", "3", "--cpu", "--port", str(free_port()), "--", "--model", "lenet",
              "--epochs", "1", "--steps-per-epoch", "12", "--log-interval", "1000", "--test-batch-size", "50", "--synthetic", "--num-push", "3",
              "--num-pull", "3", "--ckpt", str(tmp_path / "ck.pt")], cwd=tmp_path)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert os.path.exists(tmp_path / "log" / "node1.csv") and os.path.exists(tmp_path / "log" / "node2.csv")
    assert "server done" in r.stdout
    import torch
    sd = torch.load(tmp_path / "ck.pt", weights_only=False)
    keys = sd["model"].keys() if "model" in sd else sd.keys()
    assert "conv1.weight" in keys


def test_p2p_example_runs():
    r = _run([sys.executable, os.path.join(ROOT, "pytorch_p2p_ex.py")], cwd=ROOT, timeout=120)
    assert r.returncode == 0, r.stderr[-1500:]
    assert r.stdout.count("tensor(1.)") == 2


def test_azure_submit_is_guarded():
    r = _run([sys.executable, os.path.join(ROOT, "run-pytorch.py")], cwd=ROOT, timeout=60)
    assert r.returncode != 0 and "azureml-core is not installed" in (r.stderr + r.stdout)


def test_checkpoint_then_resume_continues_from_the_next_epoch(tmp_path):
    """--ckpt after every epoch, --resume picks up at epoch+1 with the saved weights (SURVEY.md §5 checkpoint / resume)."""
    common = [sys.executable, os.path.join(ROOT, "example", "main.py"), "--no-distributed", "--model", "lenet", "--steps-per-epoch", "6",
              "--log-interval", "1000", "--test-batch-size", "50", "--synthetic", "--ckpt", str(tmp_path / "ck.pt")]
    r = _run(common + ["--epochs", "1"], cwd=tmp_path)
    assert r.returncode == 0, r.stderr[-2000:]
    import torch
    first = torch.load(tmp_path / "ck.pt", weights_only=False)
    assert first["meta"]["epoch"] == 0
    r = _run(common + ["--epochs", "3", "--resume", str(tmp_path / "ck.pt")], cwd=tmp_path)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Training for epoch 0" not in r.stdout and "Training for epoch 1" in r.stdout and "Training for epoch 2" in r.stdout
    last = torch.load(tmp_path / "ck.pt", weights_only=False)
    assert last["meta"]["epoch"] == 2
    assert not torch.equal(first["model"]["conv1.weight"], last["model"]["conv1.weight"])


def test_two_sharded_servers_and_two_workers_via_launcher(tmp_path):
    """NORTH-STAR sharded server on CPU: run-pytorch.py --local 4 --cpu --servers 2 (ranks 0-1 serve one slice each, 2-3 train)."""
    from distributed_ml_pytorch_b200.parallel.launcher import free_port
    r = _run([sys.executable, os.path.join(ROOT, "run-pytorch.py"), "--local", "4", "--cpu", "--servers", "2", "--port", str(free_port()), "--",
              "--model", "lenet", "--epochs", "1", "--steps-per-epoch", "10", "--log-interval", "1000", "--test-batch-size", "50", "--synthetic",
              "--num-push", "2", "--num-pull", "2", "--ckpt", str(tmp_path / "ck.pt")], cwd=tmp_path)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert os.path.exists(tmp_path / "log" / "node2.csv") and os.path.exists(tmp_path / "log" / "node3.csv")
    assert r.stdout.count("server done") == 2
    from distributed_ml_pytorch_b200.models import build_model
    from distributed_ml_pytorch_b200.parallel import ParameterServer
    sd = ParameterServer.merge_shards([str(tmp_path / "ck.pt") + f".shard{i}" for i in range(2)], build_model("lenet"))
    assert "conv1.weight" in sd and "fc3.bias" in sd


def test_sync_mode_two_cpu_ranks_via_launcher(tmp_path):
    """--mode sync on CPU/gloo (NORTH-STAR synchronous data parallelism): no server rank, both ranks train, gradients are all-reduced."""
    from distributed_ml_pytorch_b200.parallel.launcher import free_port
    r = _run([sys.executable, os.path.join(ROOT, "run-pytorch.py"), "--local", "2", "--cpu", "--port", str(free_port()), "--", "--mode", "sync",
              "--model", "lenet", "--epochs", "1", "--steps-per-epoch", "8", "--log-interval", "1000", "--test-batch-size", "50", "--synthetic"],
             cwd=tmp_path)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert os.path.exists(tmp_path / "log" / "node0.csv") and os.path.exists(tmp_path / "log" / "node1.csv")
    assert r.stdout.count("Finished Training") == 2


def test_bench_reports_unavailable_without_a_gpu():
    """bench.py keeps its one-JSON-line contract on a machine without CUDA (the reference arm exits 0, as the driver expects)."""
    import json
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for CPU-only machines")
    for impl, rc in (("ours", 1), ("reference", 0)):
        r = _run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", impl], cwd=ROOT, timeout=120)
        assert r.returncode == rc, r.stderr[-800:]
        line = json.loads(r.stdout.strip().splitlines()[-1])
        assert line["impl"] == impl and "unavailable" in line


def test_trace_flag_writes_chrome_traces_for_server_and_workers(tmp_path):
    """--trace: one Chrome / Perfetto trace per process with steps, push / pull and the server's message handling."""
    import json
    from distributed_ml_pytorch_b200.parallel.launcher import free_port
    r = _run([sys.executable, os.path.join(ROOT, "run-pytorch.py"), "--local", "2", "--cpu", "--port", str(free_port()), "--", "--model", "lenet",
              "--epochs", "1", "--steps-per-epoch", "6", "--log-interval", "1000", "--test-batch-size", "50", "--synthetic", "--num-push", "2",
              "--num-pull", "3", "--trace", str(tmp_path / "t.json")], cwd=tmp_path)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    worker = json.load(open(tmp_path / "t.rank1.json"))["traceEvents"]
    server = json.load(open(tmp_path / "t.rank0.json"))["traceEvents"]
    names = [e["name"] for e in worker]
    assert names.count("step") == 6 and names.count("push") == 3 and names.count("pull request") == 2
    assert {"forward+loss", "backward", "optimizer.step"} <= set(names)
    assert all(e["dur"] >= 0 for e in worker if e["ph"] == "X")
    assert sum(e["name"] == "GradientUpdate" for e in server) == 3 and any(e["name"] == "ParameterRequest" for e in server)


def test_async_gloo_resume_reseeds_the_server(tmp_path):
    """--resume in asynchronous distributed mode (1 gloo server + 1 worker): the worker loads the checkpoint before it builds the optimizer,
    so the server starts from the resumed parameters and the first pull does not throw them away.  Each process writes its own file."""
    import torch
    from distributed_ml_pytorch_b200.parallel.launcher import free_port
    base = ["--model", "lenet", "--steps-per-epoch", "6", "--log-interval", "1000", "--test-batch-size", "50", "--synthetic", "--num-push", "2",
            "--num-pull", "2", "--lr", "0.05", "--ckpt", str(tmp_path / "ck.pt")]
    r = _run([sys.executable, os.path.join(ROOT, "run-pytorch.py"), "--local", "2", "--cpu", "--port", str(free_port()), "--", "--epochs", "1"] + base,
             cwd=tmp_path)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    server1 = torch.load(tmp_path / "ck.pt", weights_only=False)               # the server rank's final parameters (plain state_dict)
    worker1 = torch.load(tmp_path / "ck.pt.rank1", weights_only=False)         # the worker's replica + optimizer state
    assert worker1["format"] == "dmlb-checkpoint-v2" and worker1["meta"]["epoch"] == 0 and worker1["optimizer"]["layout"] == "logical"
    assert worker1["optimizer"]["accumulated_gradients"].numel() == 62006
    # resume from the server's parameters, zero further steps of learning rate: whatever the server holds at the end must be them
    r = _run([sys.executable, os.path.join(ROOT, "run-pytorch.py"), "--local", "2", "--cpu", "--port", str(free_port()), "--", "--epochs", "1",
              "--resume", str(tmp_path / "ck.pt")] + [a if a != "0.05" else "0.0" for a in base], cwd=tmp_path)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    server2 = torch.load(tmp_path / "ck.pt", weights_only=False)
    for k in server1:
        assert torch.equal(server1[k], server2[k]), k


def test_server_momentum_with_gain_via_launcher(tmp_path):
    """1 server + 2 workers on gloo with the NORTH-STAR server update (momentum 0.9, inverse-staleness scaling) and the default
    --server-gain auto: the run finishes, the server reports its gain, and the workers' losses stay finite and near ln(10)."""
    import csv
    from distributed_ml_pytorch_b200.parallel.launcher import free_port
    r = _run([sys.executable, os.path.join(ROOT, "run-pytorch.py"), "--local", "3", "--cpu", "--port", str(free_port()), "--", "--model", "lenet",
              "--epochs", "1", "--steps-per-epoch", "30", "--log-interval", "1000", "--test-batch-size", "50", "--synthetic", "--num-push", "2",
              "--num-pull", "2", "--momentum", "0.9", "--staleness", "inverse", "--lr", "0.05"], cwd=tmp_path)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert "server done" in r.stdout
    for node in (1, 2):
        with open(tmp_path / "log" / f"node{node}.csv") as f:
            rows = list(csv.DictReader(f))
        losses = [float(x["training_loss"]) for x in rows if x.get("training_loss") not in (None, "", "None")]
        assert losses and all(l == l and l < 5.0 for l in losses), losses[-5:]
