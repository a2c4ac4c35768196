#!/bin/bash
# BASELINE configs[3] (res15, 8192 clips/GPU, weak) and configs[2] (res8, 16384 clips split over the ranks, strong) at N GPUs.
# usage: scripts/scale_runs.sh N   ->  gpurun_out/scale_res15_N.json, gpurun_out/scale_res8_N.json
N=${1:-1}
mkdir -p gpurun_out
run() {  # name, extra args
  if [ "$N" = "1" ]; then
    python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline $2 > gpurun_out/scale_$1_$N.json 2> gpurun_out/scale_$1_$N.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 $2 \
      > gpurun_out/scale_$1_$N.json 2> gpurun_out/scale_$1_$N.err
  fi
  tail -c 600 gpurun_out/scale_$1_$N.json; echo
}
run res15 "--model res15"
run res8 "--model res8 --global-clips 16384"
