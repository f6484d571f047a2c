#!/bin/bash
# One gpurun call: tests + smoke + bench, then ncu launch list and --set full captures of the named kernels.
# Usage: bash tools/gpu_round.sh "<kernel regex list, space separated>"
set -u
bash tools/gpu_check.sh 5 3
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
for PAT in ${1:-edgeconv_tc_kernel}; do
  ncu --set full --clock-control none --import-source on -k regex:$PAT -s 8 -c 2 -f -o gpurun_out/prof_$PAT $CMD > gpurun_out/ncu_full_$PAT.log 2>&1
  echo "full capture $PAT rc=$?"
done
