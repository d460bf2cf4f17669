#!/bin/bash
python tools/profile_step.py --steps 3 > gpurun_out/plain_cl.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_cl.log; exit 1; }
ncu --metrics gpu__time_duration.sum,launch__cluster_dim_z --clock-control none --csv --log-file gpurun_out/launches_cl.csv python tools/profile_step.py --steps 3 > gpurun_out/ncu_cl.log 2>&1
echo "rc=$?"
