#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03i_shapes_split.log; : > $L
SRGPU_TRACE=1 timeout 600 python scripts/tune.py c2 1500 128x4,128x5,128x6,256x2 2>&1 | grep -E "cfg|split" >> $L
cat $L
