#!/bin/bash
# The GPU test suite under a library built with -DSSPS_BOUNDS (index checks on every scratch / Stage / slot accessor of the kernels,
# counted on the device).  compute-sanitizer is closed on the build pool; this is the memcheck stand-in.  Build here
# (tools/build_variant.sh bounds -DSSPS_BOUNDS), run on the GPU box:  tools/bounds_run.sh > gpurun_out/bounds.txt
cd "$(dirname "$0")/.."
export SSPS_LIB=$PWD/variants/bounds.so
python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -15
