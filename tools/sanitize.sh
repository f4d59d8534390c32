#!/bin/bash
# One compute-sanitizer tool over small cases of the GPU test-suite (run under gpurun, one tool per call):
#   bash tools/sanitize.sh memcheck|racecheck|synccheck
# The plain run comes first (the tool only runs if it exits 0).  Output: gpurun_out/r02_sanitizer_<tool>.txt
tool=${1:-memcheck}
out=gpurun_out/r02_sanitizer_$tool.txt
SEL='test_drive_blocked_order_ragged_p or test_resample_search_window_cases or (test_resample_blocked_bit_exact_vs_oracle and not 1048576 and not 100000) or test_resample_moves_whole_map_deferred or test_update_recolumns_observed_rows_in_kernel or (test_forced_anomaly and 1000) or (test_cooperative_weights and 1000) or test_estimate_comes_with_the_update or test_device_rng'
CMD="python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -q -x -k"
$CMD "$SEL" > gpurun_out/r02_sanitizer_plain.log 2>&1 && \
compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 0 $CMD "$SEL" > $out 2>&1
tail -15 $out
grep -c "ERROR SUMMARY" $out
