#!/bin/bash
mkdir -p gpurun_out
python bench.py --workload c4x --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02n_bench_c4x_n1.json 2> gpurun_out/r02n_bench_c4x_n1.err; echo "c4x n1 rc=$?"; cut -c1-300 gpurun_out/r02n_bench_c4x_n1.json; tail -2 gpurun_out/r02n_bench_c4x_n1.err
python scripts/tune.py c4x 1500 128x4 > gpurun_out/r02n_plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:"union_kernel|orient_kernel" -c 2 -o gpurun_out/r02n_union_orient -f python scripts/tune.py c4x 1500 128x4 > gpurun_out/r02n_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r02n_plain.log; tail -3 gpurun_out/r02n_ncu.log
