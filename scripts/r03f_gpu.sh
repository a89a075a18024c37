#!/bin/bash
# cycle counters of align_kernel and subalign_kernel in split mode (measurement build)
mkdir -p gpurun_out
SRGPU_TRACE=1 SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_prof.so timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "prof|cfg|split" | tail -40 > gpurun_out/r03f_prof_split.log; cat gpurun_out/r03f_prof_split.log
