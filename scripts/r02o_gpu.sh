#!/bin/bash
mkdir -p gpurun_out
for n in 4 6 8; do echo "SRGPU_ORIENT_CTAS=$n"; SRGPU_ORIENT_CTAS=$n timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | tail -1; done > gpurun_out/r02o_orient_ctas_c2.log; cat gpurun_out/r02o_orient_ctas_c2.log
for n in 2 4 8; do echo "SRGPU_ORIENT_CTAS=$n"; SRGPU_ORIENT_CTAS=$n timeout 300 python scripts/tune.py c4x 1500 128x4 2>&1 | tail -1; done > gpurun_out/r02o_orient_ctas_c4x.log; cat gpurun_out/r02o_orient_ctas_c4x.log
timeout 900 python -m pytest tests/test_gpu_bench_parity.py -x -q > gpurun_out/r02o_parity.log 2>&1; tail -2 gpurun_out/r02o_parity.log
