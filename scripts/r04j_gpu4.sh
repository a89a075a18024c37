#!/bin/bash
mkdir -p gpurun_out
BENCH_TRACE=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 > gpurun_out/r04j_c2_weak_n4.json 2> gpurun_out/r04j_c2_weak_n4.err
echo "first run rc=$?"; grep -h "SrgpuError\|illegal" gpurun_out/r04j_c2_weak_n4.err | head -2 | cut -c1-300; cut -c1-160 gpurun_out/r04j_c2_weak_n4.json
