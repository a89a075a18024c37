#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03t_mirror_orient.log; : > $L
for m in 1 0; do
  echo "SRGPU_MIRROR_ORIENT=$m" >> $L
  SRGPU_MIRROR_ORIENT=$m timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|seqrush_b200" | tail -1 >> $L
done
cat $L
timeout 1200 python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r03t_gpu_tests.log; cat gpurun_out/r03t_gpu_tests.log
