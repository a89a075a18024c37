#!/bin/bash
# one-warp kernel occupancy: 128-register build (13 CTAs/SM by shared memory before, 16 now) against the 96-register build (20)
mkdir -p gpurun_out
L=gpurun_out/r03a_sub_ctas.log; : > $L
for cfg in "13 1000" "16 1000" "20 1000" "20 1500" "20 1900" "16 1500"; do
  set -- $cfg
  echo "SRGPU_SUB_CTAS=$1 SRGPU_SPLIT_T=$2" >> $L
  SRGPU_TRACE=1 SRGPU_SUB_CTAS=$1 SRGPU_SPLIT_T=$2 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split" | tail -2 >> $L
done
cat $L
