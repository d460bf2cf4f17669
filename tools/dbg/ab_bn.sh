#!/bin/bash
run() { timeout 150 python bench.py --model resnet50 --input-hw 224 --classes 1000 --mode local --steps 20 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['value'])"; }
echo "== NEW"; run
cp csrc/norm_ops.cu /tmp/norm_new.cu; cp tools/dbg/norm_ops_old.cu.txt csrc/norm_ops.cu; python build_native.py 2>&1 | tail -1
echo "== OLD"; run
cp /tmp/norm_new.cu csrc/norm_ops.cu; python build_native.py 2>&1 | tail -1
echo "== NEW again"; run
