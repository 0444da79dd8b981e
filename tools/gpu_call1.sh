#!/bin/bash
# round-end style check: GPU tests, smoke, default bench, launch list, one full ncu capture of the MH kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json
if [ "$1" = "full" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-psrf > gpurun_out/ncu_l.log 2>&1; echo "launch list exit $?"
timeout 600 ncu --set full --import-source on --clock-control none -k regex:mh_sweeps -s 9 -c 1 -f -o gpurun_out/r1k_mh \
   python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-psrf > gpurun_out/ncu_r1k.log 2>&1; echo "ncu exit $?"
fi
