#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02f_gpu_tests.log 2>&1
echo "gpu tests rc=$?" >> gpurun_out/r02f_gpu_tests.log
tail -25 gpurun_out/r02f_gpu_tests.log
timeout 600 python scripts/graph_bench.py c2 > gpurun_out/r02f_graph_c2.log 2>&1; cat gpurun_out/r02f_graph_c2.log
