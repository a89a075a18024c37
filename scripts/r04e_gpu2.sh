#!/bin/bash
mkdir -p gpurun_out
SRGPU_MIRROR_ORIENT=0 BENCH_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/r04e_1.json 2> gpurun_out/r04e_1.err
echo "mirror0 first run rc=$?"; grep -h "SrgpuError\|illegal" gpurun_out/r04e_1.err | head -2; cut -c1-120 gpurun_out/r04e_1.json
