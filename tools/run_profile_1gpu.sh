#!/bin/bash
# The short form of tools/run_final_1gpu.sh (when only the update kernel changed): the 200-step line, then — only after
# the plain command exited 0 — the ncu launch list and one --set full capture of the hot kernels.
out=gpurun_out
timeout 70 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-parity > $out/r02_bench_1gpu_200steps.json 2> $out/r02_bench_1gpu_200steps.err
tail -c 300 $out/r02_bench_1gpu_200steps.json
CMD="python bench.py --steps 6 --warmup 3 --burn-in 30 --no-cpu-baseline --no-parity"
timeout 40 $CMD > $out/r02_plain.log 2>&1 && \
timeout 80 ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 110 --csv --log-file $out/r02_launches.csv $CMD > $out/r02_ncu1.log 2>&1
timeout 125 ncu --set full --clock-control none --import-source on -k regex:"k_fs2_update_fast|k_weights_fused|k_search_keep|k_predict_batch_fast|k_gather_pose|k_estimate|k_chain|k_materialise_rows" -s 280 -c 24 -o $out/r02_prof $CMD > $out/r02_ncu2.log 2>&1
ls -la $out | tail -8
