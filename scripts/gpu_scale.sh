#!/bin/bash
# weak-scaling check: bench at N = 1, 2, 4, 8 (as many as the box has), as the driver launches it
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for n in 1 2 4 8; do
  [ $n -gt $NG ] && continue
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800+n)) bench.py --gpus $n --steps 30 --warmup 3 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  echo "N=$n rc=$?"
done
python - <<PY
import json
base=None
for n in (1,2,4,8):
    try:
        d = json.loads(open(f"gpurun_out/scale_n{n}.json").read().strip().splitlines()[-1])
        base = base or d["value"]
        print(n, "value %.3e" % d["value"], "ms/step %.4f" % d["ms_per_step"], "eff %.3f" % (d["value"]/(n*base)), "e2e %.3e" % d.get("e2e",{}).get("value",0), d["roofline"]["kernel"][:60])
    except Exception as e:
        print(n, "no result", e)
PY
tail -3 gpurun_out/scale_n8.err 2>/dev/null
