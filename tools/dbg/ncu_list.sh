#!/bin/bash
# per-kernel durations of one eager AlexNet step (cold caches, serialised: compare shares), small-channel conv1 on and off
for sc in 1 0; do
  DMLB_SMALLC=$sc python tools/profile_step.py --steps 3 > gpurun_out/plain_sc$sc.log 2>&1 || { echo "plain run failed sc=$sc"; tail -5 gpurun_out/plain_sc$sc.log; exit 1; }
  DMLB_SMALLC=$sc ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_sc$sc.csv python tools/profile_step.py --steps 3 > gpurun_out/ncu_sc$sc.log 2>&1
  echo "sc=$sc rc=$?"
done
