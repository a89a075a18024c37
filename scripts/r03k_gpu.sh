#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03k_bulk_rows.log; : > $L
SRGPU_TRACE=1 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split|seqrush_b200" >> $L
SRGPU_SPLIT_T=0 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split|seqrush_b200" >> $L
cat $L
timeout 900 python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r03k_gpu_tests.log; cat gpurun_out/r03k_gpu_tests.log
