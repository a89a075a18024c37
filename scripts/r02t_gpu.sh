#!/bin/bash
mkdir -p gpurun_out
SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_checked.so timeout 600 python -m pytest tests/test_gpu_bench_parity.py -x -q -k "gmem or c3s or c4s" > gpurun_out/r02t_checked_parity.log 2>&1; echo "checked parity rc=$?"; tail -3 gpurun_out/r02t_checked_parity.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02t_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r02t_gpu_tests.log
for T in 0 1000 1900 3900; do echo "SRGPU_SPLIT_T=$T"; SRGPU_SPLIT_T=$T timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | tail -1; done > gpurun_out/r02t_split_c2.log 2>&1; cat gpurun_out/r02t_split_c2.log
