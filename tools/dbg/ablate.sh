#!/bin/bash
# step time of the flagship config with each launch-count optimisation switched off in turn
run() { echo "== $*"; env "$@" python bench.py --steps 150 --warmup 15 --no-second-dtype ${EXTRA} 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['dtype'], d['ms_per_step'], 'ms/step', d['kernels_per_step'], 'kernels', d['config']['parallelism'][:30])"; }
run X=1
run DMLB_FUSE_ACT_BWD=0
run DMLB_SMALLC=0
run DMLB_FUSE_ACT_BWD=0 DMLB_SMALLC=0
EXTRA="--mode local" run X=1
EXTRA="--dtype bf16" run X=1
EXTRA="--dtype bf16" run DMLB_FUSE_ACT_BWD=0 DMLB_SMALLC=0
