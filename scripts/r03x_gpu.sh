#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03x_fringe_deal.log; : > $L
SRGPU_TRACE=1 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split|seqrush_b200" | tail -2 >> $L
SRGPU_SPLIT_T=0 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|seqrush_b200" | tail -1 >> $L
cat $L
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r03x_gpu_tests.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r03x_gpu_tests.log
