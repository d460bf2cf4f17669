#!/bin/bash
run() { echo "== $*"; env "$@" python bench.py --steps 200 --warmup 20 --no-second-dtype ${EXTRA} 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['dtype'], d['ms_per_step'], 'ms/step dev |', d['e2e']['ms_per_step'], 'e2e |', d['kernels_per_step'], 'kernels', d['config']['parallelism'][:30])"; }
run X=1
EXTRA="--mode local" run X=1
run DMLB_SMALLC=0
EXTRA="--dtype bf16" run X=1
EXTRA="--dtype bf16 --mode local" run X=1
