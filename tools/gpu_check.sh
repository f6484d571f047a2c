#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the bench line, then the ncu launch list of the same bench command.
# Usage (from the repo root, on the GPU box):  bash tools/gpu_check.sh [steps] [warmup]
set -u
S=${1:-3}; W=${2:-3}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py --steps $S --warmup $W > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench.json
timeout 600 python tools/quick_time.py 4096 > gpurun_out/quick_time.log 2>&1; echo "quick rc=$?"
