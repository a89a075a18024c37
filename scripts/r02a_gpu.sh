#!/bin/bash
# round 2, first GPU call: parity at the benchmarked configurations + launch-shape experiments
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > gpurun_out/r02a_smi.txt
timeout 900 python -m pytest tests/test_gpu_bench_parity.py -x -q > gpurun_out/r02a_bench_parity.log 2>&1
echo "parity rc=$?" >> gpurun_out/r02a_bench_parity.log
timeout 400 python scripts/tune.py c2 1500 128x4,256x2 > gpurun_out/r02a_tune_staged.log 2>&1
SRGPU_SMEM_BUDGET=0 timeout 600 python scripts/tune.py c2 1500 128x4,64x8,32x16,128x5 > gpurun_out/r02a_tune_gmem.log 2>&1
tail -3 gpurun_out/r02a_bench_parity.log; cat gpurun_out/r02a_tune_staged.log gpurun_out/r02a_tune_gmem.log
