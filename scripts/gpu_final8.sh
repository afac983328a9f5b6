#!/bin/bash
# round-end evidence on one 8-GPU box: wide workload at N = 1, 2, 4, 8 and the default (weather) bench at N = 8
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for n in 1 2 4 8; do
  [ $n -gt $NG ] && continue
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --workload wide --gpus 1 --steps 20 --warmup 3 > gpurun_out/wide_n1.json 2> gpurun_out/wide_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29900+n)) bench.py --workload wide --gpus $n --steps 20 --warmup 3 > gpurun_out/wide_n$n.json 2> gpurun_out/wide_n$n.err
  fi
  echo "wide N=$n rc=$? $(python -c "
import json;d=json.loads(open('gpurun_out/wide_n$n.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['e2e']['value'])")"
done
if [ $NG -ge 8 ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29808 bench.py --gpus 8 --steps 30 --warmup 3 > gpurun_out/scale_n8.json 2> gpurun_out/scale_n8.err
  echo "weather N=8 rc=$? $(python -c "
import json;d=json.loads(open('gpurun_out/scale_n8.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e']['value'])")"
fi
