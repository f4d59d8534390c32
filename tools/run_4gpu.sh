export SLAM_BENCH_RANK_REPORT=1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
out=gpurun_out
timeout 300 $TR --nproc-per-node 4 --master-port 29704 tests/sharded_worker.py > $out/r02_sharded_world4.log 2>&1; tail -1 $out/r02_sharded_world4.log
timeout 400 $TR --nproc-per-node 4 --master-port 29804 bench.py --gpus 4 --steps 20 --warmup 3 > $out/r02_bench_4gpu.json 2> $out/r02_bench_4gpu.err; head -c 250 $out/r02_bench_4gpu.json; echo
SLAM_BENCH_TRACE=1 timeout 300 $TR --nproc-per-node 2 --master-port 29802 bench.py --gpus 2 --steps 20 --warmup 3 > $out/r02_bench_2gpu.json 2> $out/r02_bench_2gpu.err; head -c 250 $out/r02_bench_2gpu.json; echo
