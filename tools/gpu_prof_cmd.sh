#!/bin/bash
# ncu --set full of one kernel of an arbitrary command:  bash tools/gpu_prof_cmd.sh <kernel regex> <skip> <count> <out name> -- <cmd...>
set -u
PAT=$1; SKIP=$2; CNT=$3; NAME=$4; shift 5
mkdir -p gpurun_out
timeout 300 "$@" > gpurun_out/plain_$NAME.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$PAT -s $SKIP -c $CNT -f -o gpurun_out/prof_$NAME "$@" > gpurun_out/ncu_$NAME.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_$NAME.log
