# Same targets as the reference Makefile (run / packaging), plus build / test / bench for the native library.
PY ?= python

setup:
	$(PY) -m pip install --no-index --find-links /opt/wheelhouse -r requirements.txt || true
	$(PY) build_native.py

install:
	$(PY) -m pip install --no-build-isolation --no-deps .

build:
	$(PY) build_native.py --harness

graph:
	$(PY) example/graph.py
	mkdir -p docs && mv train_time.png test_time.png docs

first:
	$(PY) example/main.py --rank 1 --world-size 3

second:
	$(PY) example/main.py --rank 2 --world-size 3

server:
	$(PY) example/main.py --rank 0 --world-size 3 --server

single:
	$(PY) example/main.py --no-distributed

gpu:
	$(PY) example/main.py --no-distributed --cuda

gpu8:
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 example/main.py --cuda

gpu8-sync:
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 example/main.py --cuda --mode sync

# one machine, CPU / gloo: the reference's 1 server + 2 workers, a server sharded over 2 ranks + 2 workers, synchronous data parallelism
local3:
	$(PY) run-pytorch.py --local 3 --cpu

local-sharded:
	$(PY) run-pytorch.py --local 4 --cpu --servers 2

local-sync:
	$(PY) run-pytorch.py --local 2 --cpu -- --mode sync

p2p:
	$(PY) pytorch_p2p_ex.py

test:
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu:
	$(PY) -m pytest tests -x -q -m gpu

bench:
	$(PY) bench.py

dist:
	$(PY) setup.py sdist bdist_wheel

upload: dist
	twine upload dist/*

upload-test: dist
	twine upload --repository-url https://test.pypi.org/legacy/ dist/*

install-test:
	$(PY) -m pip install --index-url https://test.pypi.org/simple/ pytorch-distbelief

.PHONY: setup install build graph first second server single gpu gpu8 gpu8-sync local3 local-sharded local-sync p2p test test-gpu bench dist upload upload-test install-test

# measurement helpers (B200 box).  N = GPUs
N ?= 8
TORCHRUN = $(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(N) --master-addr 127.0.0.1 --master-port 29511

bench-scale:
	for n in 1 2 4 8; do $(PY) bench.py --gpus $$n; done

baseline-nccl:
	$(TORCHRUN) baseline/nccl_protocol.py --steps 100

sweep:
	$(TORCHRUN) tools/comm_sweep.py --out gpurun_out/comm_sweep_$(N)gpu.json

sweep-movers:
	$(TORCHRUN) tools/comm_sweep.py --movers-only --min-kb 1024 --out gpurun_out/comm_movers_$(N)gpu.json

p2p-overlap:
	$(PY) pytorch_p2p_ex.py --backend nccl --overlap

evidence-8gpu:
	bash tools/run_8gpu.sh
