#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03m_two_pass.log; : > $L
for tp in 1 0; do
  echo "SRGPU_TWO_PASS=$tp" >> $L
  SRGPU_TWO_PASS=$tp SRGPU_TRACE=1 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split|seqrush_b200" | tail -2 >> $L
done
cat $L
timeout 1200 python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r03m_gpu_tests.log; cat gpurun_out/r03m_gpu_tests.log
