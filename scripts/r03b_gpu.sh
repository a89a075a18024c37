#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r03b_gpu_tests.log; cat gpurun_out/r03b_gpu_tests.log
