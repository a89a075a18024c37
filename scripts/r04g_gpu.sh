#!/bin/bash
# final record at HEAD, one GPU: whole GPU suite, reference arm, default bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r04g_gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r04g_gpu_tests.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r04g_bench_ref.json 2> gpurun_out/r04g_bench_ref.err; echo "ref rc=$?"
python bench.py > gpurun_out/r04g_bench_c2.json 2> gpurun_out/r04g_bench_c2.err; echo "ours rc=$?"; cut -c1-330 gpurun_out/r04g_bench_c2.json
