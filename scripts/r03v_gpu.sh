#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r03v_gpu_tests.log 2>&1; echo "plain rc=$?"; tail -15 gpurun_out/r03v_gpu_tests.log
