#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02i_gpu_tests.log 2>&1
echo "gpu tests rc=$?" >> gpurun_out/r02i_gpu_tests.log
tail -15 gpurun_out/r02i_gpu_tests.log
export SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_checked.so
python scripts/sanitize_case.py big > gpurun_out/r02i_checked_case.log 2>&1; echo "checked case rc=$?"; tail -3 gpurun_out/r02i_checked_case.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02i_checked_gpu_tests.log 2>&1
echo "checked-build gpu tests rc=$?" >> gpurun_out/r02i_checked_gpu_tests.log
tail -5 gpurun_out/r02i_checked_gpu_tests.log
