#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29902 bench.py --workload wide --gpus 2 --steps 20 --warmup 3 > gpurun_out/wide_n2.json 2> gpurun_out/wide_n2.err
echo "rc=$?"; python -c "
import json;d=json.loads(open('gpurun_out/wide_n2.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['e2e']['value'])"; tail -3 gpurun_out/wide_n2.err
