#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02q_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/r02q_gpu_tests.log
timeout 600 python scripts/c5_slice.py 16 > gpurun_out/r02q_c5_slice.log 2>&1; tail -2 gpurun_out/r02q_c5_slice.log
python scripts/tune.py c2 600 128x4 > gpurun_out/r02p_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:align_kernel -s 0 -c 1 -o gpurun_out/r02p_align -f python scripts/tune.py c2 600 128x4 > gpurun_out/r02p_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r02p_plain.log
