#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03n_split_T.log; : > $L
for T in 700 1500 2000 3000; do
  echo "SRGPU_SPLIT_T=$T" >> $L
  SRGPU_SPLIT_T=$T SRGPU_TRACE=1 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split|seqrush_b200" | tail -2 >> $L
done
cat $L
