#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03y_sub20.log; : > $L
for sc in 20 16; do
  echo "SRGPU_SUB_CTAS=$sc (96-register build)" >> $L
  SRGPU_SUB_CTAS=$sc SRGPU_TRACE=1 timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | grep -E "cfg|split|seqrush_b200" | tail -2 >> $L
done
cat $L
