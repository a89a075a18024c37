#!/bin/bash
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_bench_parity.py -x -q -m gpu -k "overflow or gmem" > gpurun_out/r04k_gpu_tests.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/r04k_gpu_tests.log
