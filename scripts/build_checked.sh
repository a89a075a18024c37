#!/bin/bash
# Development build with bounds assertions (SR_ASSERT in csrc/*.cuh) -> seqrush_b200/libseqrush_b200_checked.so.
# Run any workload against it with SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_checked.so.
set -e
cd "$(dirname "$0")/.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -diag-suppress 550 \
     -DSRGPU_CHECKED -o seqrush_b200/libseqrush_b200_checked.so seqrush_b200/csrc/capi.cu
