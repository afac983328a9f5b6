#!/bin/bash
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -3
for r in 16 24; do
  export WX_WIDE_RESERVE_SMS=$r
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29951 bench.py --workload wide --gpus $NG --steps 20 --warmup 3 > gpurun_out/wide_n${NG}_r$r.json 2> gpurun_out/wide_res.err
  echo "reserve=$r rc=$? $(python -c "
import json;d=json.loads(open('gpurun_out/wide_n${NG}_r$r.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e']['value'])")"
done
