#!/bin/bash
mkdir -p gpurun_out
python bench.py > gpurun_out/r03u_bench_c2.json 2> gpurun_out/r03u_bench_c2.err; echo "ours rc=$?"; cut -c1-330 gpurun_out/r03u_bench_c2.json
