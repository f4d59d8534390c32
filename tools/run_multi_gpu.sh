#!/bin/bash
# The multi-GPU evidence of a round in one gpurun call (8-GPU box): sharded parity tests at 4 and 8 ranks, bench.py
# at 8 and 4 ranks (weak scaling at the cfg-4 size per GPU, cfg 5 as a secondary leg, parity phase inside), strong
# scaling at the cfg-4 total size.  Outputs under gpurun_out/.
export SLAM_BENCH_RANK_REPORT=1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
out=gpurun_out
for W in 8 4; do
  timeout 400 $TR --nproc-per-node $W --master-port $((29700+W)) tests/sharded_worker.py > $out/r02_sharded_world$W.log 2>&1
  tail -2 $out/r02_sharded_world$W.log
done
for W in 8 4; do
  timeout 600 $TR --nproc-per-node $W --master-port $((29800+W)) bench.py --gpus $W --steps 20 --warmup 3 \
      > $out/r02_bench_${W}gpu.json 2> $out/r02_bench_${W}gpu.err
  grep "rank" $out/r02_bench_${W}gpu.err | cut -c1-400
  head -c 300 $out/r02_bench_${W}gpu.json; echo
done
for W in 8 4 2; do
  timeout 300 $TR --nproc-per-node $W --master-port $((29900+W)) bench.py --gpus $W --particles $((1048576/W)) \
      --steps 20 --warmup 3 --no-config5 --no-parity > $out/r02_strong_${W}gpu.json 2> $out/r02_strong_${W}gpu.err
  head -c 300 $out/r02_strong_${W}gpu.json; echo
done
