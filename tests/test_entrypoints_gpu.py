"""GPU paths of the entry points and of the checkpoint format (channels-last arena weights must round-trip logically)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_main_py_cuda_single_gpu(tmp_path):
    """`make gpu` (reference Makefile:25-26): example/main.py --no-distributed --cuda, here through the CUDA-graph TrainStep."""
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "example", "main.py"), "--no-distributed", "--cuda", "--synthetic", "--epochs", "1",
                        "--steps-per-epoch", "40", "--log-interval", "20", "--test-batch-size", "256", "--ckpt", str(tmp_path / "ck.pt")],
                       cwd=tmp_path, env=env, capture_output=True, text=True, timeout=500)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert "Finished Training" in r.stdout
    from distributed_ml_pytorch_b200.utils.csvlog import read_csv
    rows = read_csv(os.path.join(tmp_path, "log", "gpu.csv"))
    assert len(rows) == 40 and rows[20]["test_accuracy"] != ""
    sd = torch.load(tmp_path / "ck.pt", weights_only=False)["model"]
    assert tuple(sd["features.3.weight"].shape) == (192, 64, 5, 5) and sd["features.3.weight"].is_contiguous()


def test_main_py_async_world1_peer_transport(tmp_path):
    """One-rank distributed run on the NVLink transport (world size 1: the GPU is its own parameter server)."""
    from distributed_ml_pytorch_b200.parallel.launcher import free_port
    env = dict(os.environ, PYTHONPATH=ROOT, RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "example", "main.py"), "--cuda", "--synthetic", "--epochs", "1", "--steps-per-epoch", "30",
                        "--log-interval", "1000", "--test-batch-size", "256", "--rank", "0", "--world-size", "1", "--model", "lenet",
                        "--port", env["MASTER_PORT"]], cwd=tmp_path, env=env, capture_output=True, text=True, timeout=500)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert os.path.exists(tmp_path / "log" / "node0.csv")


def test_checkpoint_from_gpu_arena_loads_into_reference_layout(tmp_path):
    from distributed_ml_pytorch_b200.models import AlexNet
    from distributed_ml_pytorch_b200.parallel import FusedSGD
    from distributed_ml_pytorch_b200.utils.checkpoint import load_checkpoint, save_checkpoint
    torch.manual_seed(0)
    net = AlexNet().cuda()
    opt = FusedSGD(net.parameters(), lr=0.01, model=net)            # binds the arena: conv weights become channels-last views
    assert net.features[3].weight.permute(0, 2, 3, 1).is_contiguous()
    x = torch.randn(8, 3, 32, 32)
    y_gpu = net(x.cuda()).float().cpu()
    save_checkpoint(str(tmp_path / "ck.pt"), net, opt, epoch=0)
    cpu = AlexNet()
    load_checkpoint(str(tmp_path / "ck.pt"), cpu)
    for (n1, p1), (n2, p2) in zip(net.named_parameters(), cpu.named_parameters()):
        assert torch.equal(p1.detach().cpu(), p2.detach()), n1      # logical [Cout,Cin,kh,kw] values identical
    y_cpu = cpu(x)
    assert ((y_gpu - y_cpu).norm() / y_cpu.norm()).item() < 3e-2     # bf16 tensor cores vs fp32
    # and back: a CPU (reference-layout) checkpoint loads into the GPU arena and refreshes the bf16 shadows
    with torch.no_grad():
        for p in cpu.parameters():
            p.mul_(0.5)
    torch.save(cpu.state_dict(), tmp_path / "bare.pt")
    load_checkpoint(str(tmp_path / "bare.pt"), net)
    y2 = net(x.cuda()).float().cpu()
    assert ((y2 - cpu(x)).norm() / cpu(x).norm()).item() < 3e-2


def test_main_py_trains_from_gpu_resident_cifar(tmp_path):
    """With CIFAR-10 on disk and --cuda the image set lives in HBM and each batch is one gather/normalise kernel (utils/data.py)."""
    import pickle
    import numpy as np
    base = tmp_path / "data" / "cifar-10-batches-py"
    os.makedirs(base)
    rng = np.random.default_rng(0)
    for name in [f"data_batch_{i}" for i in range(1, 6)] + ["test_batch"]:
        with open(base / name, "wb") as f:
            pickle.dump({"data": rng.integers(0, 256, (512, 3072), dtype=np.uint8), "labels": rng.integers(0, 10, 512).tolist()}, f)
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "example", "main.py"), "--no-distributed", "--cuda", "--epochs", "1", "--log-interval", "20",
                        "--test-batch-size", "256", "--data-root", str(tmp_path / "data")], cwd=tmp_path, env=env, capture_output=True, text=True,
                       timeout=500)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    from distributed_ml_pytorch_b200.utils.csvlog import read_csv
    rows = read_csv(os.path.join(tmp_path, "log", "gpu.csv"))
    assert len(rows) == 5 * 512 // 64                     # every sample once per epoch, fixed batch shape


def test_async_overlap_checkpoint_restores_the_unpushed_update(tmp_path):
    """Overlap mode keeps no accumulator (the unpushed update is replica - snapshot): a checkpoint still carries it, in logical order,
    and a fresh optimizer resumes with exactly that much left to push."""
    from distributed_ml_pytorch_b200.models import build_model
    from distributed_ml_pytorch_b200.ops import functional as OF
    from distributed_ml_pytorch_b200.parallel import Asynchronous
    from distributed_ml_pytorch_b200.utils.checkpoint import load_model_state, load_optimizer_state, read_checkpoint, save_checkpoint

    def train(net, opt, steps, seed):
        g = torch.Generator().manual_seed(seed)
        for _ in range(steps):
            x = torch.randn(32, 3, 32, 32, generator=g).cuda()
            y = torch.randint(0, 10, (32,), generator=g).cuda()
            opt.zero_grad()
            OF.cross_entropy(net(x), y).backward()
            opt.step()

    torch.manual_seed(0)
    net = build_model("lenet").cuda().eval()
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=1000, n_pull=1000, model=net)       # pushes at step 0 only
    assert opt.overlap
    train(net, opt, 4, seed=1)
    unpushed = opt.accumulated_gradients.clone()
    assert unpushed.abs().sum().item() > 0
    server = opt.server_parameters().clone()
    replica = opt.arena.flat_params().clone()
    torch.testing.assert_close(server + unpushed[: server.numel()], replica, rtol=1e-5, atol=1e-6)   # replica = server + not yet pushed
    save_checkpoint(str(tmp_path / "ck.pt"), net, opt, epoch=0)
    opt.finish()
    opt.close()

    sd, meta, ost = read_checkpoint(str(tmp_path / "ck.pt"))
    net2 = build_model("lenet").cuda().eval()
    load_model_state(net2, sd)                                   # the server's parameters (what a resumed run seeds the server with)
    opt2 = Asynchronous(net2.parameters(), lr=0.05, n_push=1000, n_pull=1000, model=net2)
    load_optimizer_state(opt2, ost)
    assert opt2.idx == 4
    torch.testing.assert_close(opt2.accumulated_gradients, unpushed, rtol=0, atol=1e-7)
    torch.testing.assert_close(opt2.server_parameters(), server, rtol=0, atol=0)
    opt2._push()                                                  # what was left goes to the server exactly once
    torch.cuda.synchronize()
    torch.testing.assert_close(opt2.server_parameters(), server + unpushed[: server.numel()], rtol=1e-5, atol=1e-6)
    assert opt2.accumulated_gradients.abs().sum().item() == 0.0
    opt2.finish()
    opt2.close()
