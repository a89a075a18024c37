#!/bin/bash
mkdir -p gpurun_out
for n in abort abort_shallow; do echo "== $n"; SRGPU_LIB=$PWD/seqrush_b200/libexp_$n.so timeout 300 python scripts/tune.py c2 1500 128x4 2>&1 | tail -2; done > gpurun_out/r02r_exp_shallow.log 2>&1; cat gpurun_out/r02r_exp_shallow.log
