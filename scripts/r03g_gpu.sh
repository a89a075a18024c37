#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r03g_shapes.log; : > $L
SRGPU_TRACE=1 timeout 600 python scripts/tune.py c2 1500 128x4,160x3,96x5 2>&1 | grep -E "cfg|split" >> $L
cat $L
