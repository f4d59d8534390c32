TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
out=gpurun_out
for W in 2 4; do timeout 300 $TR --nproc-per-node $W --master-port $((29730+W)) tests/sharded_worker.py > $out/r02_sharded_world$W.log 2>&1; tail -1 $out/r02_sharded_world$W.log; done
timeout 400 $TR --nproc-per-node 4 --master-port 29834 bench.py --gpus 4 --steps 20 --warmup 5 --no-config5 > $out/r02_bench_4gpu_final.json 2> $out/r02_bench_4gpu_final.err; head -c 200 $out/r02_bench_4gpu_final.json; echo
