#!/bin/bash
# Round-2 ncu evidence: launch list of the default bench, full captures of the headline step kernel, of the c2 GEMM / narrow
# head kernels at batch 65536, and of the cluster kernel; each only after the plain command has exited 0.
mkdir -p gpurun_out
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-secondary"
$CMD > gpurun_out/plain_r02.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r02_c3.csv $CMD > gpurun_out/ncu_list_r02.log 2>&1; echo "c3 list rc=$?"
$CMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fused_linear_step -s 4 -c 1 -f -o gpurun_out/prof_r02_c3 $CMD > gpurun_out/ncu_full_r02_c3.log 2>&1; echo "c3 full rc=$?"
CMD2="python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-secondary"
$CMD2 > gpurun_out/plain_r02_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_r02_c2.csv $CMD2 > /dev/null 2>&1; echo "c2 list rc=$?"
$CMD2 > /dev/null 2>&1 && ncu --set full --clock-control none -k regex:gemm_split3_kernel -s 6 -c 2 -f -o gpurun_out/prof_r02_c2_gemm $CMD2 > gpurun_out/ncu_full_r02_c2.log 2>&1; echo "c2 gemm rc=$?"
$CMD2 > /dev/null 2>&1 && ncu --set full --clock-control none -k regex:narrow_head_kernel -s 3 -c 1 -f -o gpurun_out/prof_r02_c2_head $CMD2 > /dev/null 2>&1; echo "c2 head rc=$?"
CMD5="python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-secondary"
$CMD5 > gpurun_out/plain_r02_c5.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_r02_c5.csv $CMD5 > /dev/null 2>&1; echo "c5 list rc=$?"
timeout 120 python tools/cs_one.py c2 100 200 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r02_cluster.csv python tools/cs_one.py c2 100 550 > /dev/null 2>&1; echo "cluster list rc=$?"
timeout 120 python simple/simple.py && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r02_simple_py.csv python simple/simple.py > /dev/null 2>&1; echo "simple rc=$?"
