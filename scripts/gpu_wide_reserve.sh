#!/bin/bash
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for cfg in "8 8" "16 16" "16 default" "24 24"; do
  set -- $cfg
  export WX_WIDE_RESERVE_SMS=$1
  if [ "$2" = default ]; then unset NCCL_MAX_CTAS; else export NCCL_MAX_CTAS=$2; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29941 bench.py --workload wide --gpus $NG --steps 20 --warmup 3 --no-e2e > gpurun_out/wide_res.json 2> gpurun_out/wide_res.err
  echo "reserve=$1 NCCL_MAX_CTAS=$2 rc=$? $(python -c "
import json;d=json.loads(open('gpurun_out/wide_res.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'])")"
done
