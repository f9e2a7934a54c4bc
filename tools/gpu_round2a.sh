#!/bin/bash
# Round-2 first contact: new persistent kernels first, then the whole GPU suite, bench, sanitizer, ncu.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== persistent tests"; timeout 900 python -m pytest tests/test_persistent_gpu.py -m gpu -q --timeout=300 > gpurun_out/pytest_persistent.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/pytest_persistent.log
echo "== full suite"; timeout 2400 python -m pytest tests -m gpu -q --timeout=600 --deselect tests/test_persistent_gpu.py > gpurun_out/pytest_all.log 2>&1; echo "rc=$?"; tail -40 gpurun_out/pytest_all.log
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/smoke.log
echo "== bench"; timeout 1200 python bench.py > gpurun_out/bench_r02a.json 2> gpurun_out/bench_r02a.err; echo "rc=$?"; cut -c1-3000 gpurun_out/bench_r02a.json; tail -5 gpurun_out/bench_r02a.err
echo "== bench per-step"; timeout 600 python bench.py --no-persistent --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/bench_r02a_perstep.json 2> gpurun_out/bench_r02a_perstep.err; echo "rc=$?"; cut -c1-1500 gpurun_out/bench_r02a_perstep.json
for c in per-step multi tiny; do
  for t in memcheck racecheck; do
    echo "== sanitizer $t $c"; timeout 600 compute-sanitizer --tool $t python tools/sanitize_case.py $c > gpurun_out/sanitizer_${t}_${c}.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/sanitizer_${t}_${c}.log
  done
done
