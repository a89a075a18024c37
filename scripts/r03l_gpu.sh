#!/bin/bash
mkdir -p gpurun_out
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r03l_bench_ref.json 2> gpurun_out/r03l_bench_ref.err; echo "ref rc=$?"; cat gpurun_out/r03l_bench_ref.json | cut -c1-900
python bench.py > gpurun_out/r03l_bench_c2.json 2> gpurun_out/r03l_bench_c2.err; echo "ours rc=$?"; cat gpurun_out/r03l_bench_c2.json
