#!/bin/bash
mkdir -p gpurun_out
SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_prof.so timeout 300 python scripts/tune.py c2 600 128x4 2>&1 | grep -E "prof cta 0|cfg" | tail -8 > gpurun_out/r02s_prof_phases.log; cat gpurun_out/r02s_prof_phases.log
