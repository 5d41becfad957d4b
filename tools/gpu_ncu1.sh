#!/bin/bash
# ncu visit 1: launch list (gpu__time_duration) of the training step; the plain run must exit 0 first
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/r02_train_launches.csv $CMD > gpurun_out/r02_ncu_run.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/r02_ncu_run.log | cut -c1-300
python tools/ncu_launch_summary.py gpurun_out/r02_train_launches.csv 4 > gpurun_out/r02_train_launches_summary.txt 2>&1; head -40 gpurun_out/r02_train_launches_summary.txt
gzip -f gpurun_out/r02_train_launches.csv
