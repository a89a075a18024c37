#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03d_split_T.log; : > $L
for cfg in 1000 700 500; do
  echo "SRGPU_SPLIT_T=$cfg" >> $L
  SRGPU_TRACE=1 SRGPU_SPLIT_T=$cfg timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split" | tail -2 >> $L
done
for cfg in 1000 500; do
  echo "c4x SRGPU_SPLIT_T=$cfg" >> $L
  SRGPU_TRACE=1 SRGPU_SPLIT_T=$cfg timeout 300 python scripts/tune.py c4x 600 128x4 2>&1 | grep -E "cfg|split" | tail -2 >> $L
done
cat $L
timeout 900 python -m pytest tests/test_gpu_bench_parity.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r03d_gpu_tests.log; cat gpurun_out/r03d_gpu_tests.log
