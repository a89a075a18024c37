#!/bin/bash
mkdir -p gpurun_out
python scripts/tune.py c2 600 128x4 > gpurun_out/r02p_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:align_kernel -s 0 -c 1 -o gpurun_out/r02p_align -f python scripts/tune.py c2 600 128x4 > gpurun_out/r02p_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r02p_plain.log
