#!/bin/bash
mkdir -p gpurun_out
SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_checked.so timeout 600 python -m pytest tests/test_gpu_bench_parity.py -x -q -k "gmem or c3s or c4s" > gpurun_out/r02m_checked_parity.log 2>&1; echo "checked parity rc=$?"; tail -2 gpurun_out/r02m_checked_parity.log
timeout 900 python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_parity.py -x -q > gpurun_out/r02m_parity.log 2>&1; echo "parity rc=$?"; tail -2 gpurun_out/r02m_parity.log
timeout 400 python scripts/tune.py c2 1500 128x4,256x2 > gpurun_out/r02m_tune_c2.log 2>&1; cat gpurun_out/r02m_tune_c2.log
timeout 400 python scripts/tune.py c4s 400 128x4 > gpurun_out/r02m_tune_c4s.log 2>&1; cat gpurun_out/r02m_tune_c4s.log
