#!/bin/bash
# usage: tools/build_variant.sh NAME [nvcc flags...]  -> variants/NAME.so (+ variants/NAME.ptxas with registers / spills of the MH kernel)
set -e
name=$1; shift
cd "$(dirname "$0")/.."
mkdir -p variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 -Xcompiler -fPIC -shared -Xptxas -v "$@" \
  -o variants/$name.so ssps_b200/csrc/ssps_capi.cu 2> variants/$name.ptxas.raw
grep -A2 "mh_sweeps_kernelI12PhiloxSourceLb0" variants/$name.ptxas.raw | grep -E "Used|spill" > variants/$name.ptxas || true
cat variants/$name.ptxas
