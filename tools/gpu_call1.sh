#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
tools/gpu_call2.sh
for v in dflt nols16 isq; do echo "== m2 $v"; SSPS_LIB=$PWD/variants/$v.so timeout 200 python tools/m2_probe.py 128 60 4000 2>&1 | tail -3; done
