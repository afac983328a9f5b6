#!/bin/bash
# wide path: launch list of the bench command, then one --set full capture of the GEMM kernels (after the plain run exits 0)
mkdir -p gpurun_out
MODE=${1:-both}
timeout 300 python bench.py --workload wide --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/wide_plain.log 2>&1 || exit 1
if [ "$MODE" != full ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/wide_launches.csv python bench.py --workload wide --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/wide_ncu1.log 2>&1
echo "launch list rc=$?"
fi
if [ "$MODE" != list ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wide_gemm -s 18 -c 9 -f -o gpurun_out/prof_wide python bench.py --workload wide --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/wide_ncu2.log 2>&1
echo "full rc=$?"
fi
tail -1 gpurun_out/wide_plain.log | cut -c1-300; tail -2 gpurun_out/wide_ncu2.log
