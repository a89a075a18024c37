#!/bin/bash
mkdir -p gpurun_out
for T in 500 1000 1900; do SRGPU_TRACE=1 SRGPU_SPLIT_T=$T timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "srgpu|cfg" | tail -2; done > gpurun_out/r02u_split_phases.log 2>&1; cat gpurun_out/r02u_split_phases.log
