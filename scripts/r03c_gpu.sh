#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03c_overlap.log; : > $L
for cfg in "1000" "500" "0"; do
  echo "SRGPU_SPLIT_T=$cfg" >> $L
  SRGPU_TRACE=1 SRGPU_SPLIT_T=$cfg timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split" | tail -2 >> $L
done
cat $L
timeout 900 python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r03c_gpu_tests.log; cat gpurun_out/r03c_gpu_tests.log
