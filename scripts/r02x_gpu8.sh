#!/bin/bash
# 8-GPU record: strong scaling on the north_star shape (c4x), weak on c2. One rank per GPU.
mkdir -p gpurun_out
run() { # N workload scaling tag
  local N=$1 wl=$2 sc=$3 tag=$4
  if [ "$N" = 1 ]; then
    timeout 900 python bench.py --gpus 1 --workload $wl --scaling $sc --no-cpu-baseline > gpurun_out/r02x_${tag}.json 2> gpurun_out/r02x_${tag}.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $wl --scaling $sc > gpurun_out/r02x_${tag}.json 2> gpurun_out/r02x_${tag}.err
  fi
  echo "$tag rc=$?"; cut -c1-700 gpurun_out/r02x_${tag}.json
}
run 8 c4x strong c4x_strong_n8
run 4 c4x strong c4x_strong_n4
run 1 c4x strong c4x_strong_n1
run 8 c2 weak c2_weak_n8
