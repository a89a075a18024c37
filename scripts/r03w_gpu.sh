#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_bench_parity.py -x -q -m gpu -k "overflow or threshold" > gpurun_out/r03w_gpu_tests.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r03w_gpu_tests.log
