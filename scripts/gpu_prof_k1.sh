#!/bin/bash
# ncu --set full capture of the K1 streaming kernels (one launch each), after the plain run exited 0.
mkdir -p gpurun_out
timeout 200 python scripts/k1_bench.py > gpurun_out/k1_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"col_stats_kernel|normalize_pack|label_encode" -c 40 -f -o gpurun_out/prof_k1 env REPS=1 python scripts/k1_bench.py > gpurun_out/ncu_k1.log 2>&1
echo "rc=$?"; cat gpurun_out/k1_plain.log | tail -8; tail -2 gpurun_out/ncu_k1.log
