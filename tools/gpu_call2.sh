#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B="python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-psrf"
for rep in 1 2; do
for so in variants/*.so; do
  n=$(basename $so .so)
  SSPS_LIB=$PWD/$so $B 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$n', '%.3e'%d['value'], '%.3e'%d['roofline']['kernel_proposals_per_s'], d['ms_per_step'])"
done; done | tee gpurun_out/sweep2.txt
