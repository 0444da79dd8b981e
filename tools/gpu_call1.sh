#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for i in 1 2 3 4 5 6; do SSPS_TRACE=1 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-psrf 2>gpurun_out/tr$i.err | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('%.3e'%d['value'], d['ms_per_step'], 'e2e %.3e'%d['e2e']['value'])"; grep trace gpurun_out/tr$i.err | head -5; done
