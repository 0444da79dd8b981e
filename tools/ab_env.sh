#!/bin/bash
# usage: tools/ab_env.sh "VAR=a VAR2=b" "VAR=c" ...   -- bench.py (kernel rate, e2e) under each environment setting ("-" = none), two rounds
cd "$(dirname "$0")/.."
for round in 1 2; do
for v in "$@"; do
  if [ "$v" = "-" ]; then e=""; else e="$v"; fi
  env $e python bench.py --steps ${AB_STEPS:-10} --warmup 5 --no-psrf --no-cpu-baseline --e2e-steps ${AB_E2E:-10} 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('[$v]', 'value %.4e kernel %.4e e2e %.4e' % (d['value'], d['roofline']['kernel_proposals_per_s'], d['e2e']['value']))"
done; done
