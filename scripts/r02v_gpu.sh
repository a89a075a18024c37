#!/bin/bash
mkdir -p gpurun_out
for T in 0 500 1000 1900; do SRGPU_TRACE=1 SRGPU_SPLIT_T=$T timeout 300 python scripts/tune.py c4x 1500 128x4 2>&1 | grep -E "srgpu|cfg" | tail -2; done > gpurun_out/r02v_split_c4x.log 2>&1; cat gpurun_out/r02v_split_c4x.log
for T in 0 500 1000; do SRGPU_TRACE=1 SRGPU_SPLIT_T=$T timeout 300 python scripts/tune.py c3 3000 128x4 2>&1 | grep -E "srgpu|cfg" | tail -2; done > gpurun_out/r02v_split_c3.log 2>&1; cat gpurun_out/r02v_split_c3.log
