#!/bin/bash
run() { echo "== $* ${EXTRA}"; env "$@" python bench.py --steps 200 --warmup 20 --no-second-dtype ${EXTRA} 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['dtype'], d['ms_per_step'], 'ms/step dev |', d['e2e']['ms_per_step'], 'e2e |', d['kernels_per_step'], 'kernels', d['config']['parallelism'][:30])"; }
run DMLB_CLUSTER_SPLITK=0
run DMLB_CLUSTER_MAXKB=4
run DMLB_CLUSTER_MAXKB=7
EXTRA="--dtype bf16" run DMLB_CLUSTER_SPLITK=0
EXTRA="--dtype bf16" run DMLB_CLUSTER_MAXKB=4
EXTRA="--dtype bf16" run DMLB_CLUSTER_MAXKB=7
EXTRA="--dtype bf16" run DMLB_CLUSTER_MAXKB=100
