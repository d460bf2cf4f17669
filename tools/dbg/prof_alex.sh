#!/bin/bash
o=gpurun_out/prof_r2; mkdir -p $o
python tools/profile_step.py --model alexnet --steps 5 > $o/alex_plain.log 2>&1 || { echo plain failed; tail -5 $o/alex_plain.log; exit 1; }
timeout 300 ncu --set full --clock-control none -s 90 -c 42 -o /tmp/alexnet_step python tools/profile_step.py --model alexnet --steps 5 > $o/alex_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/alexnet_step.ncu-rep --page raw --csv > $o/alexnet_step_raw.csv 2>/dev/null; ls -la $o/alexnet_step_raw.csv
