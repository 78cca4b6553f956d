#!/bin/sh
# Builds the microbenchmarks next to their sources (binaries are git-ignored).  usage: sh tools/microbench/build.sh
set -e
cd "$(dirname "$0")"
for f in *.cu; do
  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o "${f%.cu}" "$f"
done
