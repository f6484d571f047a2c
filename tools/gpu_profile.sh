#!/bin/bash
# ncu evidence for the bench command (run under gpurun, 1 GPU).  Plain run first; ncu only if it exits 0.
#   gpurun_out/launches.csv : every launch with its device time (cold-cache, serialised: compare SHARES)
#   gpurun_out/prof_<name>.ncu-rep : --set full capture of the kernel matching $1 (default edgeconv_kernel)
set -u
PAT=${1:-edgeconv_kernel}; SKIP=${2:-8}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$PAT -s $SKIP -c 3 -f -o gpurun_out/prof_$PAT $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
tail -3 gpurun_out/ncu_full.log
