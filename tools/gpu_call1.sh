#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python tools/diag.py 100 4096 400 verify 2>&1 | grep "rep 3"
timeout 600 ncu --set full --import-source on --clock-control none -k regex:mh_sweeps -s 9 -c 1 -f -o gpurun_out/r1k_mh \
   python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-psrf > gpurun_out/ncu_r1k.log 2>&1; echo "ncu exit $?"
