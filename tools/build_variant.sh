#!/bin/bash
# usage: tools/build_variant.sh <name> [extra nvcc flags ...]  -> rl-cattle-herding_b200/variants/libchs_<name>.so
# Builds a side-by-side copy of the library from the working tree for A/B runs (CHS_LIB=... python tools/ab_bench.py);
# the in-tree libchs.so is left alone.  variants/ is git-ignored (*.so) and travels to the GPU box with gpurun.
set -e
here=$(cd "$(dirname "$0")/.." && pwd)
name=$1; shift
mkdir -p "$here/rl-cattle-herding_b200/variants"
cd "$here/rl-cattle-herding_b200/csrc"
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -ftz=true -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared "$@" \
  -o "../variants/libchs_$name.so" chs_api.cu > "/tmp/build_$name.log" 2>&1 || { grep -iE " error" "/tmp/build_$name.log" | head; exit 1; }
echo "built variants/libchs_$name.so"
