#!/bin/bash
mkdir -p gpurun_out
SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_prof.so timeout 300 python scripts/tune.py c2 600 128x4 > gpurun_out/r02d_prof_128x4.log 2>&1
grep -E "prof cta 0|cfg" gpurun_out/r02d_prof_128x4.log | tail -8
SRGPU_LIB=$PWD/seqrush_b200/libseqrush_b200_prof.so timeout 300 python scripts/tune.py c2 600 256x2 > gpurun_out/r02d_prof_256x2.log 2>&1
grep -E "prof cta 0|cfg" gpurun_out/r02d_prof_256x2.log | tail -8
python scripts/tune.py c2 600 128x4 > gpurun_out/r02d_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:align_kernel -s 1 -c 1 \
    -o gpurun_out/r02d_align -f python scripts/tune.py c2 600 128x4 > gpurun_out/r02d_ncu.log 2>&1
echo "rc=$?"; cat gpurun_out/r02d_plain.log
