#!/bin/bash
run() { echo "== $* ${EXTRA}"; env "$@" python bench.py --steps 300 --warmup 30 --no-second-dtype ${EXTRA} 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['dtype'], d['ms_per_step'], 'ms/step dev |', d['e2e']['ms_per_step'], 'e2e |', d['kernels_per_step'], 'kernels', d['clocks'])"; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,temperature.gpu --format=csv
run DMLB_CLUSTER_SPLITK=0
run DMLB_CLUSTER_SPLITK=0
run X=1
EXTRA="--dtype bf16" run DMLB_CLUSTER_SPLITK=0
