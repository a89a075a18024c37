#!/bin/bash
# ncu --set full + source of the two alignment kernels at HEAD (split mode, 600-pair c2 slice)
mkdir -p gpurun_out
python scripts/tune.py c2 600 128x4 > gpurun_out/r03h_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:align_kernel -s 0 -c 2 -o gpurun_out/r03h_align -f python scripts/tune.py c2 600 128x4 > gpurun_out/r03h_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r03h_plain.log; tail -3 gpurun_out/r03h_ncu.log
