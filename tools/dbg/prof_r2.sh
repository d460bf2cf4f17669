#!/bin/bash
# ResNet-50 step launch list + world-1 capture of the communication kernels
o=gpurun_out/prof_r2; mkdir -p $o
python tools/profile_step.py --model resnet50 --hw 224 --batch 64 --steps 3 > $o/r50_plain.log 2>&1 || { echo "r50 plain failed"; tail -5 $o/r50_plain.log; }
timeout 400 ncu --nvtx --nvtx-include "step2/" --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file $o/r50_step.csv python tools/profile_step.py --model resnet50 --hw 224 --batch 64 --steps 3 > $o/r50_ncu.log 2>&1; echo "r50 ncu rc=$?"; wc -l $o/r50_step.csv
python tools/comm_world1.py 64 > $o/comm_plain.log 2>&1 || { echo "comm plain failed"; tail -5 $o/comm_plain.log; }
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"push_add_kernel|pull_kernel|push_bulk_kernel|pull_bulk_kernel|push_mailbox_kernel|owner_apply_kernel|sync_rs_update_ag_kernel" -s 9 -c 9 -o $o/comm_world1 python tools/comm_world1.py 64 > $o/comm_ncu.log 2>&1; echo "comm ncu rc=$?"; ls -la $o
