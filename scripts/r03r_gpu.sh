#!/bin/bash
# bounds-asserting build (-DSRGPU_CHECKED) over the whole GPU suite at HEAD; then the plain suite
mkdir -p gpurun_out
SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_checked.so timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r03r_checked_gpu_tests.log 2>&1; echo "checked rc=$?"; tail -3 gpurun_out/r03r_checked_gpu_tests.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r03r_gpu_tests.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/r03r_gpu_tests.log
