#!/bin/bash
# ncu evidence for the tensor-core step kernel: one full-set capture (with source) of two launches of the bench workload.
mkdir -p gpurun_out
CMD="python bench.py --steps 4 --warmup 3 --repeats 1 --no-cpu-baseline --no-e2e ${PROF_ARGS}"
$CMD > gpurun_out/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tc_step_kernel -s 3 -c 2 -f -o gpurun_out/prof_tc $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/ | head -20
