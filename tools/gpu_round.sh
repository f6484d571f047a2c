#!/bin/bash
# One gpurun call: tests + smoke + bench, then ncu launch list and --set full captures of the named kernels.
# Usage: bash tools/gpu_round.sh "<kernel regex list, space separated>"
set -u
bash tools/gpu_check.sh 5 3
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
# each argument word: <kernel regex>[:<launches to skip>[:<launches to capture>]]
for SPEC in ${1:-edgeconv_tc_kernel}; do
  PAT=${SPEC%%:*}; REST=${SPEC#*:}; [ "$REST" = "$SPEC" ] && REST="8:2"
  SKIP=${REST%%:*}; CNT=${REST#*:}; [ "$CNT" = "$REST" ] && CNT=2
  ncu --set full --clock-control none --import-source on -k regex:$PAT -s $SKIP -c $CNT -f -o gpurun_out/prof_$PAT $CMD > gpurun_out/ncu_full_$PAT.log 2>&1
  echo "full capture $PAT (skip $SKIP, count $CNT) rc=$?"
  # summarise on the box: gpurun_out/ travels back only below 64 MiB
  python tools/ncu_summary.py gpurun_out/prof_$PAT.ncu-rep > gpurun_out/ncu_$PAT.txt 2>&1
  python tools/ncu_hot.py gpurun_out/prof_$PAT.ncu-rep 24 > gpurun_out/hot_$PAT.txt 2>&1
  [ -n "${KEEP_REP:-}" ] || rm -f gpurun_out/prof_$PAT.ncu-rep
done
