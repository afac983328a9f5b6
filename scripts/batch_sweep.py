"""BASELINE.json configs[4] and configs[2]: batch-size sweep 256 .. 1 Mi rows/step/GPU and one epoch over a 100 M-row
table, on however many GPUs torchrun gives it.  One process per GPU; the step is StepEngine.step_raw (fused
standardise + forward + CE + backward + DDP exchange + Adam, one launch per step).  CUDA events, max over ranks.
Run: python scripts/batch_sweep.py            (1 GPU)
     python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29911 scripts/batch_sweep.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import wxddp
from bench import synth_raw_on_device
from wxddp.engine import StepEngine

rank = int(os.environ.get("RANK", "0"))
local_rank = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
F = int(os.environ.get("F", 20))
TOTAL_ROWS = int(os.environ.get("TOTAL_ROWS", 100_000_000))
torch.cuda.set_device(local_rank)
device = torch.device("cuda", local_rank)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=device)

shard = TOTAL_ROWS // world // 4 * 4            # rows of the table this rank holds (contiguous shard)
raw, y = synth_raw_on_device(torch, shard, F, 1234 + rank, device)
cols = [raw[j] for j in range(F)]
torch.manual_seed(42)
eng = StepEngine(F, device=device, world=world, rank=rank, process_group=dist)
eng.fit_stats(cols)
if world > 1:
    eng.connect_peers()


def sync():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, n):
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    sync()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


out = {"n_gpus": world, "features": F, "table_rows_total": shard * world, "sweep": [], "epoch": None}
for B in (256, 1024, 4096, 16384, 65536, 262144, 1048576):
    nwin = shard // B
    steps = 200 if B <= 65536 else 50
    fn = lambda i, B=B, nwin=nwin: eng.step_raw(cols, y, row0=((i * 7919) % nwin) * B, batch=B)
    timed(fn, 5)
    l0 = eng.ctx.launch_count
    ms = timed(fn, steps)
    out["sweep"].append({"rows_per_step_per_gpu": B, "steps": steps, "us_per_step": 1e3 * ms / steps,
                         "samples_per_s": steps * B * world / (ms / 1e3),
                         "launches_per_step": (eng.ctx.launch_count - l0) / steps,
                         "hbm_gbs_per_gpu": steps * B * (4 * F + 8) / (ms / 1e3) / 1e9})

# one epoch over the whole sharded table at 1 Mi rows/step/GPU (the last window is ragged)
B = 1 << 20
n_steps = (shard + B - 1) // B
fn = lambda i: eng.step_raw(cols, y, row0=i * B, batch=min(B, shard - i * B) // 4 * 4)
timed(fn, 3)
ms = timed(fn, n_steps)
out["epoch"] = {"rows_total": shard * world, "steps": n_steps, "ms": ms, "samples_per_s": shard * world / (ms / 1e3),
                "final_loss": float(eng.loss.item())}
status = eng.ctx.tc_status()
out["tc_status"] = status
if rank == 0:
    for r in out["sweep"]:
        print("B/GPU %8d  %9.2f us/step  %.4e samples/s  %.1f launches/step  %.1f GB/s/GPU" % (
            r["rows_per_step_per_gpu"], r["us_per_step"], r["samples_per_s"], r["launches_per_step"],
            r["hbm_gbs_per_gpu"]))
    e = out["epoch"]
    print("epoch over %d rows: %d steps, %.2f ms, %.4e samples/s, loss %.4f" % (e["rows_total"], e["steps"], e["ms"],
                                                                                e["samples_per_s"], e["final_loss"]))
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
