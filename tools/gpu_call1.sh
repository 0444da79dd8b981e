#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
SSPS_LIB=$PWD/variants/prof.so DIAG_PROF=1 python tools/diag.py 100 128 4000 2>&1 | tail -1
python tools/diag.py 100 4096 400 verify 2>&1 | grep "rep 3"
for i in 1 2; do python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-psrf 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('%.3e'%d['value'], d['ms_per_step'], 'e2e %.3e'%d['e2e']['value'])"; done
timeout 200 python tools/m2_probe.py 128 60 4000 2>&1 | tail -2
