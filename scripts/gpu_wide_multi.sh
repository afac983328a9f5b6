#!/bin/bash
# wide workload at N = 1 and at every GPU count the box offers; bench lines land in gpurun_out/wide_n*.json
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for n in 1 2 4 8; do
  [ $n -gt $NG ] && continue
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --workload wide --gpus 1 --steps 20 --warmup 3 > gpurun_out/wide_n1.json 2> gpurun_out/wide_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29900+n)) bench.py --workload wide --gpus $n --steps 20 --warmup 3 > gpurun_out/wide_n$n.json 2> gpurun_out/wide_n$n.err
  fi
  echo "N=$n rc=$?"; tail -c 1500 gpurun_out/wide_n$n.json; tail -3 gpurun_out/wide_n$n.err
done
