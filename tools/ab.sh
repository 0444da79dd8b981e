#!/bin/bash
# usage: tools/ab.sh variantA variantB ...   (names under variants/; "base" = the in-tree library) -- kernel rate of bench.py per variant, two rounds
cd "$(dirname "$0")/.."
for round in 1 2; do
for v in "$@"; do
  if [ "$v" = base ]; then unset SSPS_LIB; else export SSPS_LIB=$PWD/variants/$v.so; fi
  python bench.py --steps ${AB_STEPS:-10} --warmup 5 --no-psrf --no-cpu-baseline --e2e-steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v', 'value %.4e kernel %.4e e2e %.4e' % (d['value'], d['roofline']['kernel_proposals_per_s'], d['e2e']['value']))"
done; done
