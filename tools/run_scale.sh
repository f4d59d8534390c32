#!/bin/bash
# bench.py at 8 / 4 / 2 ranks of one box exactly as the driver launches it (weak scaling, 2^20 particles per GPU);
# the 4096-landmark secondary leg only at 8 ranks.  Outputs under gpurun_out/.
export SLAM_BENCH_RANK_REPORT=1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
out=gpurun_out
for W in 8 4 2; do
  extra=""; [ $W -ne 8 ] && extra="--no-config5"
  timeout 420 $TR --nproc-per-node $W --master-port $((29800+W)) bench.py --gpus $W --steps 20 --warmup 5 $extra \
      > $out/r02_bench_${W}gpu.json 2> $out/r02_bench_${W}gpu.err
  head -c 260 $out/r02_bench_${W}gpu.json; echo
done
