#!/bin/bash
# ncu evidence for one bench step (run under gpurun, 1 GPU).  Plain run first; ncu only if it exits 0.
#   gpurun_out/step_metrics.csv : DRAM bytes + duration of the 32 launches of the third step (cold-cache, serialised)
set -u
mkdir -p gpurun_out
python tools/one_step.py > gpurun_out/plain_step.log 2>&1 &&
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -s 64 -c 32 --csv \
    --log-file gpurun_out/step_metrics.csv python tools/one_step.py > gpurun_out/ncu_step.log 2>&1
echo "step metrics rc=$?"; tail -2 gpurun_out/ncu_step.log
