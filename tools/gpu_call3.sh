#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for n in "$@"; do
SSPS_LIB=$PWD/variants/$n.so timeout 600 ncu --section WarpStateStats --section SchedulerStats --section InstructionStats --clock-control none -k regex:mh_sweeps -s 8 -c 1 -f -o gpurun_out/ws_$n \
   python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-psrf > gpurun_out/ncu_$n.log 2>&1; echo "ncu $n exit $?"
done
