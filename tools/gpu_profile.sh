#!/bin/bash
# ncu evidence for the headline workload: launch list + one full capture of the dominant kernel.
# Usage (under gpurun): bash tools/gpu_profile.sh <tag> <kernel-regex> [bench args]
TAG=${1:-r01}; KRE=${2:-fused_linear_step}; shift 2
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e $*"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KRE -s 4 -c 2 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"
tail -3 gpurun_out/plain_$TAG.log
