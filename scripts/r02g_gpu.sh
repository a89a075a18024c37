#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02g_gpu_tests.log 2>&1
echo "gpu tests rc=$?" >> gpurun_out/r02g_gpu_tests.log
tail -25 gpurun_out/r02g_gpu_tests.log
timeout 900 python scripts/c5_slice.py 16 > gpurun_out/r02g_c5_slice.log 2>&1; tail -5 gpurun_out/r02g_c5_slice.log
