#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bench_parity.py -x -q -k "gmem or c3s or c4s" > gpurun_out/r02e_parity.log 2>&1; tail -2 gpurun_out/r02e_parity.log
timeout 400 python scripts/tune.py c2 1500 128x4,128x5,256x2 > gpurun_out/r02e_tune_c2.log 2>&1; cat gpurun_out/r02e_tune_c2.log
SRGPU_STAGE_SEQS=0 timeout 400 python scripts/tune.py c2 1500 128x4,128x5,64x8 > gpurun_out/r02e_tune_c2_gmem.log 2>&1; cat gpurun_out/r02e_tune_c2_gmem.log
timeout 400 python scripts/tune.py c4s 400 256x2,128x4 > gpurun_out/r02e_tune_c4s.log 2>&1; cat gpurun_out/r02e_tune_c4s.log
SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_prof.so timeout 300 python scripts/tune.py c2 600 128x4 2>&1 | grep -E "prof cta 0|cfg" | tail -5
