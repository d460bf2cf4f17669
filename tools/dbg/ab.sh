#!/bin/bash
run() { python bench.py --steps 300 --warmup 30 --no-second-dtype $* 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['dtype'], d['ms_per_step'], 'ms/step dev |', d['kernels_per_step'], 'kernels', d['config']['parallelism'][:20])"; }
echo "== HEAD"; run; run --mode local
echo "== OLD"; (cd _old && run; run --mode local)
echo "== HEAD again"; run
