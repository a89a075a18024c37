#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/r03z_c2_weak_n2.json 2> gpurun_out/r03z_c2_weak_n2.err
echo "rc=$?"; cut -c1-330 gpurun_out/r03z_c2_weak_n2.json; tail -3 gpurun_out/r03z_c2_weak_n2.err
