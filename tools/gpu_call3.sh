#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:mh_sweeps -s 8 -c 1 -f -o gpurun_out/r1i_mh \
   python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-psrf > gpurun_out/ncu_r1i.log 2>&1; echo "ncu exit $?"
