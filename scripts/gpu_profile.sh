#!/bin/bash
# ncu evidence for the dominant kernel: launch list (time shares) + one full-set capture.
mkdir -p gpurun_out
CMD="python bench.py --steps ${PROF_STEPS:-4} --warmup 3 --no-cpu-baseline --no-e2e ${PROF_ARGS}"
$CMD > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launch-list rc=$?"
$CMD > gpurun_out/prof_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${PROF_KERNEL:-mlp_grid_kernel} -s 3 -c 2 -f -o gpurun_out/prof_step $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/
