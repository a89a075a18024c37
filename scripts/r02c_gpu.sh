#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02c_gpu_tests.log 2>&1
echo "gpu tests rc=$?" >> gpurun_out/r02c_gpu_tests.log
tail -15 gpurun_out/r02c_gpu_tests.log
timeout 400 python scripts/tune.py c2 1500 128x4,256x2,64x8 > gpurun_out/r02c_tune_c2.log 2>&1
cat gpurun_out/r02c_tune_c2.log
timeout 400 python scripts/tune.py c4s 400 256x2,128x4 > gpurun_out/r02c_tune_c4s.log 2>&1
cat gpurun_out/r02c_tune_c4s.log
