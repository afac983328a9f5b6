#!/bin/bash
# One 8-GPU box: batch sweep + 100M-row epoch (configs[4], configs[2]) and the wide workload at N = 4, 8 (configs[3]).
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "GPUs: $NG"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29911 scripts/batch_sweep.py > gpurun_out/sweep_n$NG.log 2> gpurun_out/sweep_n$NG.err
echo "sweep rc=$?"; grep -v "^{" gpurun_out/sweep_n$NG.log | tail -9; tail -2 gpurun_out/sweep_n$NG.err
for n in 4 8; do
  [ $n -gt $NG ] && continue
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29900+n)) bench.py --workload wide --gpus $n --steps 20 --warmup 3 > gpurun_out/wide_n$n.json 2> gpurun_out/wide_n$n.err
  echo "wide N=$n rc=$?"; python -c "
import json;d=json.loads(open('gpurun_out/wide_n$n.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['e2e']['value'])"; tail -2 gpurun_out/wide_n$n.err
done
