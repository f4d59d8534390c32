#!/bin/bash
# The 1-GPU evidence of a round in one gpurun call: the default bench line, a 200-step line, the reference arm, the
# FastSLAM1 secondary line, the cfg-5 per-GPU shape on one GPU, then (only after the plain command exited 0) the ncu
# launch list and one --set full capture of the hot kernels.
out=gpurun_out
python bench.py > $out/r02_bench_1gpu.json 2> $out/r02_bench_1gpu.err; tail -2 $out/r02_bench_1gpu.err
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-parity > $out/r02_bench_1gpu_200steps.json 2>> $out/r02_bench_1gpu.err
python bench.py --impl reference > $out/r02_bench_reference_arm.json 2>> $out/r02_bench_1gpu.err
python bench.py --landmarks 4096 --steps 20 --warmup 3 --no-cpu-baseline --no-parity > $out/r02_bench_nf4096_1gpu.json 2>> $out/r02_bench_1gpu.err
python tools/bench_fs1.py > $out/r02_fs1_secondary.json 2>> $out/r02_bench_1gpu.err
CMD="python bench.py --steps 6 --warmup 3 --burn-in 30 --no-cpu-baseline --no-parity"
$CMD > $out/r02_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 110 --csv --log-file $out/r02_launches.csv $CMD > $out/r02_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_fs2_update_fast|k_weights_fused|k_search_keep|k_predict_batch_fast|k_gather_pose|k_estimate|k_chain|k_materialise_rows" -s 280 -c 24 -o $out/r02_prof $CMD > $out/r02_ncu2.log 2>&1
ls -la $out | tail -12
