#!/bin/bash
# One gpurun call for the round's final evidence: tests + smoke + bench (tools/gpu_check.sh), the ncu launch list of the bench
# command, the DRAM-traffic capture and the --set full capture of one step.  Each ncu pass only after its command exited 0 plain.
set -u
bash tools/gpu_check.sh 5 3
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
bash tools/gpu_profile_step.sh
bash tools/gpu_profile_full.sh
