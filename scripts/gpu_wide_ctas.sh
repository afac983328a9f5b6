#!/bin/bash
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for c in 2 4 8 16 default; do
  if [ "$c" = default ]; then unset NCCL_MAX_CTAS; else export NCCL_MAX_CTAS=$c; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29931 bench.py --workload wide --gpus $NG --steps 20 --warmup 3 --no-e2e > gpurun_out/wide_ctas.json 2> gpurun_out/wide_ctas.err
  echo "NCCL_MAX_CTAS=$c rc=$? $(python -c "
import json;d=json.loads(open('gpurun_out/wide_ctas.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'])")"
done
