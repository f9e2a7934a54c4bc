#!/bin/bash
# First-contact GPU run: smoke, parity tests, a short bench; each step under its own timeout.
# Usage (under gpurun): bash tools/gpu_check.sh [pytest-args]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke" ; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1 ; echo "smoke rc=$?" ; tail -5 gpurun_out/smoke.log
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q "$@" > gpurun_out/pytest.log 2>&1 ; echo "pytest rc=$?" ; tail -25 gpurun_out/pytest.log
echo "== bench" ; timeout 900 python bench.py --steps 100 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err ; echo "bench rc=$?" ; cat gpurun_out/bench.json ; tail -5 gpurun_out/bench.err
