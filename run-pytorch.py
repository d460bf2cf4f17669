#!/usr/bin/env python
"""Job launcher — the reference's run-pytorch.py submits example/main.py to an AzureML compute target
(run-pytorch.py:7-19).  Kept, behind an import guard (azureml is not installable here), and extended with the launch
mode this framework is built for: one process per GPU on an NVSwitch box.

    python run-pytorch.py                      # AzureML submit, as in the reference (needs azureml-core + config.json)
    python run-pytorch.py --local 8 -- --cuda --model alexnet --epochs 1      # 8 local GPU processes running example/main.py
    python run-pytorch.py --local 3 --cpu -- --model lenet                    # rank 0 = server, ranks 1-2 = workers (gloo)
"""
import argparse
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def submit_azureml(extra):
    try:
        from azureml.core import Experiment, ScriptRunConfig, Workspace
    except ImportError:
        raise SystemExit("azureml-core is not installed; use `python run-pytorch.py --local N -- <main.py flags>` to launch locally")
    ws = Workspace.from_config()
    experiment = Experiment(workspace=ws, name="single-cpu")
    config = ScriptRunConfig(source_directory="./example", script="main.py", compute_target="distbelief-single", arguments=list(extra))
    run = experiment.submit(config)
    aml_url = run.get_portal_url()
    print(aml_url)


def launch_local(n, cpu, extra, port, servers=1):
    extra = list(extra)
    sync = any(a == "--mode" and b == "sync" for a, b in zip(extra, extra[1:])) or "--mode=sync" in extra
    if sync:
        servers = 0                            # synchronous data parallelism: every rank trains, nobody serves
    procs = []
    for rank in range(n):
        env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(n), LOCAL_RANK=str(rank))
        cmd = [sys.executable, os.path.join(HERE, "example", "main.py"), "--rank", str(rank), "--world-size", str(n), "--master", "127.0.0.1",
               "--port", str(port)] + list(extra)
        if cpu and rank < servers:
            cmd.append("--server")            # reference topology: rank 0 is the parameter server (main.py:151,153); --servers K: ranks 0..K-1
        if servers > 1:
            cmd += ["--num-servers", str(servers)]
        procs.append(subprocess.Popen(cmd, env=env))
    rc = 0
    for p in procs:
        rc = max(rc, p.wait())
    return rc


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--local", type=int, default=0, help="launch N local processes instead of submitting to AzureML")
    ap.add_argument("--cpu", action="store_true", help="CPU/gloo topology: rank 0 serves, ranks 1.. train")
    ap.add_argument("--port", type=int, default=29500)
    ap.add_argument("--servers", type=int, default=1, help="with --cpu: shard the parameter server over ranks 0..K-1")
    ap.add_argument("extra", nargs=argparse.REMAINDER)
    a = ap.parse_args()
    extra = [e for e in a.extra if e != "--"]
    if a.local:
        sys.exit(launch_local(a.local, a.cpu, extra, a.port, a.servers))
    submit_azureml(extra)
