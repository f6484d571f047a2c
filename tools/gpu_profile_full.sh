#!/bin/bash
# ncu --set full over the 32 launches of one bench step (run under gpurun, 1 GPU).  Plain run first; ncu only if it exits 0.
#   gpurun_out/step_full_raw.csv : `--page raw --csv` export of every launch (tools/ncu_step_table.py turns it into a table)
set -u
mkdir -p gpurun_out
python tools/one_step.py > gpurun_out/plain_step.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_step.log; exit 1; }
ncu --set full --clock-control none -s 64 -c 32 -f -o /tmp/step_full python tools/one_step.py > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_full.log
ncu -i /tmp/step_full.ncu-rep --page raw --csv > gpurun_out/step_full_raw.csv 2> gpurun_out/ncu_export.err
ls -la /tmp/step_full.ncu-rep gpurun_out/step_full_raw.csv
