#!/bin/bash
mkdir -p gpurun_out
export MODE=${MODE:-0} STEPS=3
timeout 120 python scripts/tc_bench.py > gpurun_out/tc_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_step_kernel -s 2 -c 1 -f -o gpurun_out/prof_tc python scripts/tc_bench.py > gpurun_out/ncu_tc.log 2>&1
echo "rc=$?"; cat gpurun_out/tc_plain.log; tail -2 gpurun_out/ncu_tc.log
