#!/bin/bash
# ncu --set full of the two launches of the one-warp kernel (13 CTAs/SM: at 16 ncu cannot save/restore the workspaces)
mkdir -p gpurun_out
SRGPU_SUB_CTAS=13 python scripts/tune.py c2 600 128x4 > gpurun_out/r03p_plain.log 2>&1 &&
SRGPU_SUB_CTAS=13 ncu --set full --clock-control none --import-source on -k regex:subalign_kernel -s 0 -c 2 -o gpurun_out/r03p_subalign -f python scripts/tune.py c2 600 128x4 > gpurun_out/r03p_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r03p_plain.log; tail -3 gpurun_out/r03p_ncu.log
