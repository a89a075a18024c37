#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r02l_no_alloc_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/r02l_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02l_gpu_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -4 gpurun_out/r02l_gpu_tests.log
