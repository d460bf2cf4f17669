#!/bin/bash
# One 8-GPU box: multi-GPU semantics tests, scaling bench, NCCL-protocol baseline, communication sweeps.  Logs -> gpurun_out/g8/
out=gpurun_out/g8; mkdir -p $out
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv | head -3 > $out/gpus.txt
timeout 420 python -m pytest tests/test_peer_gpu.py -q --tb=short -rs > $out/peer_tests_8gpu.log 2>&1; tail -3 $out/peer_tests_8gpu.log
for n in 8 4 1; do timeout 150 python bench.py --gpus $n --steps 300 --warmup 30 --no-second-dtype 2>$out/bench_async_$n.err | tail -1 > $out/bench_async_$n.json; cut -c1-200 $out/bench_async_$n.json; done
timeout 150 python bench.py --gpus 8 --mode sync --steps 300 --warmup 30 2>$out/bench_sync_8.err | tail -1 > $out/bench_sync_8.json; cut -c1-200 $out/bench_sync_8.json
for n in 8 4; do timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2960$n baseline/nccl_protocol.py --steps 100 2>$out/nccl_protocol_$n.err | grep '^{' > $out/nccl_protocol_$n.json; cat $out/nccl_protocol_$n.json; done
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29618 tools/comm_sweep.py --movers-only --min-kb 1024 --max-mb 1024 --out $out/comm_movers_8gpu.json > $out/comm_movers_8gpu.log 2>&1; grep -c '^{' $out/comm_movers_8gpu.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29614 tools/comm_sweep.py --max-mb 1024 --out $out/comm_sweep_4gpu.json > $out/comm_sweep_4gpu.log 2>&1; grep -c '^{' $out/comm_sweep_4gpu.log
P2P_PORT=29630 P2P_OVERLAP_OUT=$out/p2p_overlap.json timeout 120 python pytorch_p2p_ex.py --backend nccl --overlap > $out/p2p_overlap.log 2>&1; tail -6 $out/p2p_overlap.log
