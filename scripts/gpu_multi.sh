#!/bin/bash
# multi-GPU pass: N-GPU DDP parity test + bench at N GPUs via torchrun
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -p no:cacheprovider --timeout 800 2>&1 | tail -15
for n in $(seq 1 $N); do :; done
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "rc=$?"; tail -3 gpurun_out/bench_n$N.err
python - <<PY
import json
for n in (1, $N):
    try:
        d = json.loads(open(f"gpurun_out/bench_n{n}.json").read().strip().splitlines()[-1])
        print(n, "value", d["value"], "ms/step", d["ms_per_step"], "e2e", d.get("e2e", {}).get("value"))
    except Exception as e:
        print(n, "no result", e)
PY
