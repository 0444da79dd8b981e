#!/bin/bash
# GPU-side variant sweep: bench.py kernel rate for every variants/*.so (and SSPS_MAX_OCC for the base build)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-psrf"
for so in variants/*.so; do
  n=$(basename $so .so)
  SSPS_LIB=$PWD/$so $B 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$n', '%.3e'%d['value'], '%.3e'%d['roofline']['kernel_proposals_per_s'], d['ms_per_step'])"
done
for occ in 1 2; do
  SSPS_MAX_OCC=$occ SSPS_LIB=$PWD/variants/base.so $B 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('base_occ$occ', '%.3e'%d['value'], '%.3e'%d['roofline']['kernel_proposals_per_s'], d['ms_per_step'])"
done
