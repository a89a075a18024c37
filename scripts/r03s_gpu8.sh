#!/bin/bash
# 8-GPU record at HEAD: strong scaling on the north_star shape (c4x), weak on c2. One rank per GPU.
mkdir -p gpurun_out
run() { # N workload scaling tag
  local N=$1 wl=$2 sc=$3 tag=$4
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $wl --scaling $sc > gpurun_out/r03s_${tag}.json 2> gpurun_out/r03s_${tag}.err
  echo "$tag rc=$?"; cut -c1-330 gpurun_out/r03s_${tag}.json
}
run 8 c4x strong c4x_strong_n8
run 8 c2 weak c2_weak_n8
