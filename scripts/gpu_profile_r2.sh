#!/bin/bash
# Round-2 evidence for profiles/: launch list of the default bench command (shares), one full-set capture of the step
# kernel, one of the wide GEMM.  Run under gpurun; every ncu command follows a plain run of the same command line.
mkdir -p gpurun_out
CMD="python bench.py --steps 4 --warmup 3 --repeats 1 --no-cpu-baseline --no-e2e --no-extras"
$CMD > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_tc_launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/prof_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tc_step_kernel -s 3 -c 2 -f -o gpurun_out/prof_tc $CMD > gpurun_out/ncu_full.log 2>&1
echo "tc full rc=$?"
WCMD="python bench.py --workload wide --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$WCMD > gpurun_out/prof_wide_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_wide_launches.csv $WCMD > gpurun_out/ncu_wide_launch.log 2>&1
echo "wide launch list rc=$?"
ls -la gpurun_out | head -30
