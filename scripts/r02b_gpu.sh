#!/bin/bash
# ncu --set full with source on the align kernel (600-pair c2 slice, 128x4)
mkdir -p gpurun_out
python scripts/tune.py c2 600 128x4 > gpurun_out/r02b_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:align_kernel -s 1 -c 1 \
    -o gpurun_out/r02b_align -f python scripts/tune.py c2 600 128x4 > gpurun_out/r02b_ncu.log 2>&1
echo "rc=$?"; cat gpurun_out/r02b_plain.log; tail -5 gpurun_out/r02b_ncu.log; ls -la gpurun_out/
