#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03q_leaf_kernel.log; : > $L
for lc in 32 24 16; do
  echo "SRGPU_LEAF_CTAS=$lc" >> $L
  SRGPU_LEAF_CTAS=$lc SRGPU_TRACE=1 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split|seqrush_b200" | tail -2 >> $L
done
cat $L
timeout 1200 python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r03q_gpu_tests.log; cat gpurun_out/r03q_gpu_tests.log
