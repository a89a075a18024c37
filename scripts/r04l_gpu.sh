#!/bin/bash
mkdir -p gpurun_out
timeout 140 python -m pytest tests -x -q -m gpu > gpurun_out/r04l_gpu_tests.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r04l_gpu_tests.log
