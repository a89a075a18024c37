#!/bin/bash
# record at HEAD: reference arm, default bench (c2), c4x and c3 at N=1, then the ncu launch list of the default bench command
mkdir -p gpurun_out
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r03o_bench_ref.json 2> gpurun_out/r03o_bench_ref.err; echo "ref rc=$?"
python bench.py > gpurun_out/r03o_bench_c2.json 2> gpurun_out/r03o_bench_c2.err; echo "ours rc=$?"; cut -c1-330 gpurun_out/r03o_bench_c2.json
python bench.py --workload c4x --scaling strong --no-cpu-baseline > gpurun_out/r03o_bench_c4x.json 2> gpurun_out/r03o_bench_c4x.err; echo "c4x rc=$?"; cut -c1-330 gpurun_out/r03o_bench_c4x.json
python bench.py --workload c3 --no-cpu-baseline > gpurun_out/r03o_bench_c3.json 2> gpurun_out/r03o_bench_c3.err; echo "c3 rc=$?"; cut -c1-330 gpurun_out/r03o_bench_c3.json
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r03o_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r03o_launches_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r03o_ncu.log 2>&1
echo "ncu rc=$?"
