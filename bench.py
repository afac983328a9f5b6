#!/usr/bin/env python
"""bench.py -- train samples/sec of the weather-MLP training step on N B200s (+ roofline, CPU baseline).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--rows-per-step B] [--features F] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one optimizer step of the hot path (standardise -> forward -> cross-entropy -> backward ->
Adam) over `rows-per-step` rows PER GPU of a synthetic weather table (weak scaling).  Workload, roofline
arithmetic and the CPU baseline are defined in DESIGN.md "Measurement".  One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

# Rank 0 prints ONE JSON line on stdout.  Libraries write there too (NCCL's version banner, its INFO lines when the
# environment asks for them): main() keeps a private copy of the real stdout for that line and points file descriptor 1 at
# stderr for everything else.
_RESULT_OUT = None


def emit(line: dict) -> None:
    out = _RESULT_OUT if _RESULT_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "train samples/sec"
UNIT = "samples/s"


def read_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def profiled_traffic(matches_profiled_workload: bool):
    """roofline.traffic is not measurable inside an un-profiled run: it is dram__bytes_read.sum + dram__bytes_write.sum per
    launch of the step kernel from the committed ncu --set full capture of this exact workload (profiles/traffic.json,
    which names the capture); any other workload has no capture -> null."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not matches_profiled_workload or not os.path.exists(path):
        return None, None
    with open(path) as fh:
        d = json.load(fh)
    return float(d["dram_bytes_per_launch"]), d["source"]


def median(xs):
    xs = sorted(xs)
    n = len(xs)
    return xs[n // 2] if n % 2 else 0.5 * (xs[n // 2 - 1] + xs[n // 2])


def timed_repeats(torch, dist, world, device, one_step, steps: int, repeats: int, barrier, fill_ms: float = 1000.0):
    """Time `steps` steps `repeats` times (CUDA events on the launch stream, barrier + synchronize on both sides of
    every repeat, max over ranks per repeat).  repeats == 0: as many as fill ~`fill_ms` of device time, at least 5, so
    that the clock sampler sees the region.  Returns the per-repeat milliseconds."""
    def once():
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for _ in range(steps):
            one_step()
        ev1.record()
        barrier()
        t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    times = [once()]
    if repeats <= 0:
        repeats = max(5, min(400, int(fill_ms / max(times[0], 1e-3)) + 1))
    while len(times) < repeats:
        times.append(once())
    return times


def synth_raw_on_device(torch, n_rows: int, f: int, seed: int, device):
    """Raw (un-standardised) fp32 columns + int64 labels, generated on the GPU (same family of distributions
    as oracle/synth.py; generation is outside every timed region)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    mu = [20.0, 65.0, 10.0, 50.0, 1013.0]
    sg = [8.0, 20.0, 6.0, 30.0, 7.0]
    raw = torch.empty(f, n_rows, dtype=torch.float32, device=device)
    for j in range(f):
        raw[j].normal_(mu[j % 5] + 3.0 * (j // 5), sg[j % 5] * (1.0 + 0.25 * (j // 5)), generator=g)
    if f > 5:
        raw[f - 1].fill_(42.5)      # constant column: exercises the std==0 guard (preprocess.py:36)
    z0 = (raw[0] - mu[0]) / sg[0]
    z1 = (raw[1 % f] - mu[1 % 5]) / sg[1 % 5]
    score = -1.6 + 0.9 * z1 - 0.7 * z0 + 0.6 * torch.randn(n_rows, device=device, generator=g)
    y = (score > 0).to(torch.int64)
    return raw, y


def cpu_baseline(f: int, rows_per_rank: int, steps: int, warmup: int):
    """The oracle's restatement of the reference step under gloo DDP, world_size 2 (the reference's topology,
    docker-compose.yml:124), on this host's cores.  Bounded sample: `steps` steps of rows_per_rank rows/rank."""
    from oracle import train_ref

    cores = os.cpu_count() or 2
    threads = max(1, min(32, cores // 2))
    r = train_ref.time_cpu_ddp(f=f, batch=rows_per_rank, steps=steps, warmup=warmup, world=2, threads=threads,
                               n_rows=max(4 * rows_per_rank, 145_460), per_sample=False)
    return r, threads


def run_reference(args, rank: int):
    if rank != 0:
        return 0
    f = args.features
    rows = args.cpu_rows_per_step if args.cpu_rows_per_step else args.rows_per_step
    rows_per_rank = max(1, rows // 2)
    r, threads = cpu_baseline(f, rows_per_rank, args.steps, args.warmup)
    val = r["samples_per_sec"]
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * r["seconds"] / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, rows_per_gpu=rows),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": r["cores"], "kind": "port",
                         "sample": f"{args.steps} steps x {rows_per_rank} rows/rank x 2 ranks, gloo DDP on localhost, "
                                   f"{threads} threads/rank, tensor-sliced batches (oracle/train_ref.py restatement of "
                                   f"jobs/train_lightning_ddp.py; Lightning/MLflow not installable)"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


def workload_config(args, rows_per_gpu: int) -> dict:
    return {"workload": f"weather MLP F={args.features} H=64 C=2 fp32: fused standardise+forward+CE+backward+Adam "
                        f"step (BASELINE.json configs[1]); {rows_per_gpu} rows/step/GPU streamed from a "
                        f"{args.table_rows}-row raw fp32 columnar table resident in HBM",
            "features": args.features, "rows_per_step_per_gpu": rows_per_gpu, "table_rows_per_gpu": args.table_rows,
            "dropout": "philox p=0.2", "optimizer": "Adam lr=0.01",
            "l2_policy": "inputs larger than L2: every step reads a different window of a table >> 126 MB",
            "parallelism": f"dp{args.gpus}"}


def read_tensor_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        if "bf16_tflops_sustained" in d:
            return float(d["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained: step of back-to-back GEMMs)"
    return 1416.4, "fallback (B200_PROFILING.md sustained cuBLAS bf16)"


def wide_config(args, hidden: int, rows: int) -> dict:
    return {"workload": f"wide tabular MLP {args.features}->{hidden}x3->2 (4 layers, hidden {hidden}; BASELINE.json "
                        f"configs[3]), bf16 tcgen05 GEMMs with fp32 accumulate, fp32 head/loss/Adam: forward + CE + "
                        f"backward + DDP mean + Adam per step; {rows} rows/step/GPU",
            "features": args.features, "hidden": hidden, "layers": 4, "rows_per_step_per_gpu": rows,
            "dropout": "philox p=0.2", "optimizer": "Adam lr=0.01 (fp32 master weights, bf16 copies re-emitted)",
            "l2_policy": "working set per step (activations + weights, > 1 GB) is far larger than the 126 MB L2",
            "parallelism": f"dp{args.gpus}"}


def run_wide(args, rank: int, local_rank: int, world: int) -> int:
    """BASELINE.json configs[3]: same JSON contract, tensor-pipe roofline."""
    f, hidden, nl, B = args.features, args.wide_hidden, 3, args.wide_rows
    weights = f * hidden + (nl - 1) * hidden * hidden + 2 * hidden
    flops_per_row = 6.0 * weights - 2.0 * f * hidden            # no dX for the first layer
    if args.impl == "reference":
        if rank != 0:
            return 0
        from oracle import wide_ref
        rows = min(B, 1024)
        r = wide_ref.time_cpu_steps(f, hidden, nl, rows, max(1, min(args.steps, 6)), warmup=1)
        line = {"impl": "reference", "metric": METRIC, "value": r["samples_per_sec"], "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * r["seconds"] / max(1, min(args.steps, 6)),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": wide_config(args, hidden, B),
                "cpu_baseline": {"value": r["samples_per_sec"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                 "sample": f"{max(1, min(args.steps, 6))} steps x {rows} rows, torch fp32 nn.Sequential + "
                                           f"autograd + torch.optim.Adam on the host cores ({r['seconds']:.1f} s)"},
                "e2e": {"value": r["samples_per_sec"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return 0

    import torch

    if args.warmup < 3:
        args.warmup = 3
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: wxddp has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    res = wide_measure(args, torch, dist, world, rank, local_rank, device, steps=args.steps, min_seconds=0.0,
                       with_e2e=not args.no_e2e)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    line = {"metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": wide_config(args, hidden, B), "roofline": res["roofline"], "clocks": res["clocks"],
            "gpu_launches": res["gpu_launches"], "final_loss": res["final_loss"], "kernel_status": res["kernel_status"]}
    if "e2e" in res:
        line["e2e"] = res["e2e"]
    if not args.no_cpu_baseline and world == 1:
        from oracle import wide_ref
        r = wide_ref.time_cpu_steps(f, hidden, nl, 1024, 4, warmup=1)
        line["cpu_baseline"] = {"value": r["samples_per_sec"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                "sample": f"4 steps x 1024 rows, torch fp32 nn.Sequential + autograd + torch.optim.Adam "
                                          f"on the host cores ({r['seconds']:.1f} s)"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def wide_measure(args, torch, dist, world, rank, local_rank, device, steps: int, min_seconds: float, with_e2e: bool) -> dict:
    """BASELINE.json configs[3] on the GPUs of this launch: `steps` optimizer steps per repeat, repeated until at least
    `min_seconds` of device time have been timed (so that the SUSTAINED tensor peak is the honest denominator); both
    the sustained and the burst fraction are reported.  Every rank returns; rank 0's dict is the one printed."""
    from wxddp.wide import WideEngine

    f, hidden, nl, B = args.features, args.wide_hidden, 3, args.wide_rows
    weights = f * hidden + (nl - 1) * hidden * hidden + 2 * hidden
    flops_per_row = 6.0 * weights - 2.0 * f * hidden            # no dX for the first layer
    torch.manual_seed(42)
    eng = WideEngine(f, hidden, nl, max_batch=B, device=device, world=world, rank=rank, process_group=dist,
                     seed=42 + 1000 * rank)
    n_win = 8
    g = torch.Generator(device=device)
    g.manual_seed(1234 + rank)
    table = torch.randn(n_win * B, f, device=device, generator=g)           # standardised features (K1's output)
    labels = (torch.rand(n_win * B, device=device, generator=g) < 0.3).to(torch.int64)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step_no = [0]

    def one_step():
        w = step_no[0] % n_win
        eng.step(table[w * B:(w + 1) * B], labels[w * B:(w + 1) * B])
        step_no[0] += 1

    for _ in range(max(3, args.warmup)):
        one_step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = eng.launch_count
    times = timed_repeats(torch, dist, world, device, one_step, steps, 0, barrier, fill_ms=1e3 * min_seconds)
    launches = (eng.launch_count - l0) // len(times)
    clocks = sampler.stop() if rank == 0 else None
    ms = median(times)
    final_loss = float(eng.grads[eng.P].item())
    status = eng.status()
    out = {}
    if with_e2e:
        hx = [table[i * B:(i + 1) * B].cpu().pin_memory() for i in range(4)]
        hy = [labels[i * B:(i + 1) * B].cpu().pin_memory() for i in range(4)]
        for i in range(3):
            eng.step_host(hx[i % 4], hy[i % 4])
        barrier()
        t0 = time.perf_counter()
        for i in range(steps):
            last = eng.step_host(hx[i % 4], hy[i % 4])
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        out["e2e"] = {"value": steps * B * world / dt, "unit": UNIT, "h2d_bytes_per_step": B * (4 * f + 8),
                      "d2h_bytes_per_step": 4, "ms_per_step": 1e3 * dt / steps, "last_loss": last,
                      "api": "wxddp.wide.WideEngine.step_host(pinned fp32 features, int64 labels) -> loss"}
    peak, peak_src = read_tensor_peak()
    burst = read_burst_peak()
    step_s = (ms / 1e3) / steps
    achieved = B * flops_per_row / step_s / 1e12
    out.update({
        "value": steps * B * world / (ms / 1e3), "ms_per_step": ms / steps, "steps": steps, "repeats": len(times),
        "timed_seconds": sum(times) / 1e3, "gpu_launches": int(launches), "final_loss": final_loss, "kernel_status": status,
        "clocks": clocks, "rows_per_step_per_gpu": B,
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                     "frac_sustained": achieved / peak, "frac_burst": (achieved / burst) if burst else None,
                     "peak_burst": burst, "traffic": None,
                     "kernel": "wide_gemm2_kernel (tcgen05.mma.cta_group::2 kind::f16, 256x256x64 tiles per SM pair, TMA "
                               "128B-swizzle tensor maps, 6-stage ring, TMEM double buffer): 7 of the 8 GEMM launches "
                               "per step; bias gradients, all-reduce and Adam run on a side stream under them",
                     "peak_source": peak_src, "algorithmic_flops_per_row": flops_per_row,
                     "note": "achieved = algorithmic FLOPs of the whole step / step time, i.e. it charges the head, "
                             "bias-gradient, cast and Adam kernels and the DDP exchange to the GEMMs; frac = frac_sustained "
                             "when the timed region is seconds long, frac_burst applies to sub-second regions"}})
    del eng, table, labels
    torch.cuda.empty_cache()
    return out


def read_burst_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        if "bf16_tflops" in d:
            return float(d["bf16_tflops"])
    return 1590.0


def main() -> int:
    global _RESULT_OUT
    sys.stdout.flush()
    _RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="wxddp", choices=["wxddp", "reference"])
    ap.add_argument("--features", type=int, default=20)
    ap.add_argument("--rows-per-step", type=int, default=1 << 20)
    ap.add_argument("--table-rows", type=int, default=1 << 24)
    ap.add_argument("--cpu-rows-per-step", type=int, default=0,
                    help="CPU arms: rows per step (0 = --rows-per-step, i.e. the GPU arm's configuration)")
    ap.add_argument("--cpu-steps", type=int, default=20, help="steps of the cpu_baseline leg inside the default run")
    ap.add_argument("--repeats", type=int, default=0,
                    help="timed repeats of --steps steps each (0 = as many as fill ~1 s, at least 5); the median is reported")
    ap.add_argument("--kernel", default="auto", choices=["auto", "tc", "ffma"], help="step kernel at N=1")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the configs[2]/[3]/[4] sections (s3_epoch, wide, sweep) of the default line")
    ap.add_argument("--s3-rows", type=int, default=100_000_000, help="total rows of the configs[2] table (split over the GPUs)")
    ap.add_argument("--wide-seconds", type=float, default=4.0, help="timed seconds of the wide section of the default line")
    ap.add_argument("--workload", default="weather", choices=["weather", "wide"],
                    help="weather = BASELINE.json configs[1] (default, the headline); wide = configs[3], the 4-layer "
                         "hidden-4096 bf16 tcgen05 MLP")
    ap.add_argument("--wide-rows", type=int, default=16384, help="rows/step/GPU of the wide workload")
    ap.add_argument("--wide-hidden", type=int, default=4096)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.workload == "wide":
        return run_wide(args, rank, local_rank, world)
    if args.impl == "reference":
        return run_reference(args, rank)

    import torch

    import wxddp

    if args.warmup < 3:
        args.warmup = 3
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: wxddp has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)

    f, B, T = args.features, args.rows_per_step, args.table_rows
    if T < 2 * B:
        T = args.table_rows = 2 * B
    ctx = wxddp.Context(f, device=device)
    P = ctx.P
    raw, y = synth_raw_on_device(torch, T, f, 1234 + rank, device)
    cols = [raw[j] for j in range(f)]
    stats = ctx.col_stats(cols)
    if world > 1:
        parts = [torch.empty_like(stats) for _ in range(world)]
        dist.all_gather(parts, stats)
        stats = ctx.stats_merge(torch.stack(parts).contiguous())
    torch.manual_seed(42)
    from torch import nn
    net = nn.Sequential(nn.Linear(f, 64), nn.ReLU(), nn.Dropout(0.2), nn.Linear(64, 2))
    params = torch.cat([p.detach().reshape(-1) for p in net.parameters()]).to(device)
    m = torch.zeros(P, device=device)
    v = torch.zeros(P, device=device)
    state = torch.zeros(2, dtype=torch.int64, device=device)
    loss = torch.zeros(1, device=device)
    grads = torch.zeros(P + 1, device=device)
    feats = None
    if world > 1:
        feats, _ = ctx.normalize_pack(cols, stats, impute=wxddp.IMPUTE_MEAN)
    n_windows = (T - B) // B + 1
    step_no = [0]
    bound = ctx.bind_columns(cols)          # pointer array validated once: the per-step call is the C ABI call only
    use_tc = args.kernel != "ffma" and ctx.tc_ok_raw(cols, y, 0) and B % 4 == 0
    step_raw = ctx.tc_train_step_raw if use_tc else ctx.train_step_raw
    kernel_name = ("tc_step_kernel (tcgen05 tf32x3, wx_tc_train_step_raw)" if use_tc
                   else "mlp_grid_kernel<F,TRAIN,RAW_F32> (FFMA, wx_train_step_raw)")
    grad_fn = ctx.mlp_grad
    if world > 1 and args.kernel != "ffma" and ctx.tc_ok_rowmajor(feats, y, 0) and (B * f) % 4 == 0:
        grad_fn = ctx.tc_mlp_grad

    fused_ddp = False
    if world > 1 and use_tc:
        # DDP exchange inside the step kernel: every rank maps every rank's inbox (CUDA IPC) once
        handles = [None] * world
        dist.all_gather_object(handles, ctx.peer_export())
        ctx.peer_connect(rank, world, handles)
        dist.barrier()
        fused_ddp = True
        kernel_name = "tc_step_kernel (tcgen05 tf32x3) with in-kernel [P+1] all-reduce over NVLink peer memory"

    def one_step():
        s = step_no[0]
        row0 = (s % n_windows) * B
        if world == 1 or fused_ddp:
            step_raw(params, m, v, state, bound, stats, y, loss, row0=row0, batch=B,
                     impute=wxddp.IMPUTE_MEAN, dropout=wxddp.DROPOUT_PHILOX, seed=42 + 1000 * rank)
        else:
            # DDP: local sums pre-scaled by 1/(W*B), SUM across ranks (what torch's reducer does, :136),
            # loss scalar rides in the same bucket (element P), then Adam
            grad_fn(params, feats, y, row0=row0, batch=B, dropout=wxddp.DROPOUT_PHILOX, seed=42 + 1000 * rank,
                    step=s, scale=1.0 / (world * B), out=grads)
            dist.all_reduce(grads)
            ctx.adam_step(params, grads, m, v, s + 1, n=P)
        step_no[0] = s + 1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        one_step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launch_count
    times = timed_repeats(torch, dist, world, device, one_step, args.steps, args.repeats, barrier)
    ms = median(times)
    launches = (ctx.launch_count - launches0) // len(times)
    clocks = sampler.stop() if rank == 0 else None
    final_loss = float(loss.item()) if (world == 1 or fused_ddp) else float(grads[P].item())
    tc_status = ctx.tc_status()

    # ---- N > 1: the replicas must be bit-identical, and one fused step (in-kernel NVLink exchange) must agree with the
    #      same step taken through the NCCL bucket path (wx_tc_mlp_grad -> all_reduce -> wx_adam_step) from the same state
    ddp_check = None
    if world > 1 and fused_ddp:
        gathered = [torch.empty_like(params) for _ in range(world)]
        dist.all_gather(gathered, params)
        identical = all(bool(torch.equal(gathered[0], gq)) for gq in gathered[1:])
        p0, m0, v0, s0 = params.clone(), m.clone(), v.clone(), state.clone()
        step0 = int(s0[0].item())
        row0 = 3 * B if T >= 4 * B else 0
        step_raw(params, m, v, state, bound, stats, y, loss, row0=row0, batch=B, impute=wxddp.IMPUTE_MEAN,
                 dropout=wxddp.DROPOUT_PHILOX, seed=42 + 1000 * rank)
        torch.cuda.synchronize()
        m_fused, loss_fused = m.clone(), float(loss.item())
        params.copy_(p0); m.copy_(m0); v.copy_(v0); state.copy_(s0)
        grad_fn(params, feats, y, row0=row0, batch=B, dropout=wxddp.DROPOUT_PHILOX, seed=42 + 1000 * rank, step=step0,
                scale=1.0 / (world * B), out=grads)
        dist.all_reduce(grads)
        ctx.adam_step(params, grads, m, v, step0 + 1, n=P)
        torch.cuda.synchronize()
        upd = float((0.1 * grads[:P]).norm().item())               # m_new - 0.9 m_old = 0.1 g
        rel = float((m_fused - m).norm().item()) / max(upd, 1e-30)
        ddp_check = {"replicas_identical": bool(identical), "vs_nccl_rel": rel,
                     "vs_nccl_max_abs": float((m_fused - m).abs().max().item()), "grad_l2": upd / 0.1,
                     "grad_max_abs": float(grads[:P].abs().max().item()),
                     "loss_fused": loss_fused, "loss_nccl": float(grads[P].item()), "tc_status": tc_status,
                     "what": "params all-gathered after the timed loop and compared bit for bit; then one step from the same "
                             "state through both paths: |m_fused - m_nccl| / |0.1 g| (the fused path standardises raw columns in "
                             "the kernel, the NCCL path reads the packed table: <= 1 ulp apart per cell; after thousands of "
                             "steps the mean gradient is a small difference of large per-row terms, so the relative figure is "
                             "read next to vs_nccl_max_abs and the gradient's own size)"}
        params.copy_(p0); m.copy_(m0); v.copy_(v0); state.copy_(s0)

    # ---- e2e: public API with HOST buffers; H2D of the step's raw rows + labels and D2H of the loss inside ----
    e2e = None
    if not args.no_e2e:
        from wxddp.engine import StepEngine
        eng = StepEngine(f, device=device, world=world, rank=rank, process_group=dist if world > 1 else None)
        eng.set_stats(stats)
        if world > 1:
            eng.connect_peers()
        n_host = 4
        host_cols = [torch.empty(f, B, dtype=torch.float32).pin_memory() for _ in range(n_host)]
        # labels travel as one byte per row (the 0/1 codes of preprocess.py:23-25); the int64 cast of :46 happens on the device
        host_y = [torch.empty(B, dtype=torch.uint8).pin_memory() for _ in range(n_host)]
        for i in range(n_host):
            host_cols[i].copy_(raw[:, i * B:(i + 1) * B])
            host_y[i].copy_(y[i * B:(i + 1) * B])
        for i in range(3):
            eng.submit_host(host_cols[i % n_host], host_y[i % n_host])
        eng.drain()
        dts = []
        for _ in range(5):
            barrier()
            t0 = time.perf_counter()
            n_read = 0
            for i in range(args.steps):
                # H2D of batch i overlaps the kernel of batch i-1; every step's loss is read back (one call late)
                r = eng.submit_host(host_cols[i % n_host], host_y[i % n_host])
                n_read += r is not None
            last = eng.drain()
            n_read += 1
            assert n_read == args.steps
            barrier()
            tt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dts.append(float(tt.item()))
        dt = median(dts)
        # what the host can deliver at all: the same pinned buffers copied by every rank at once, nothing else running
        ceil_dst = torch.empty(f, B, dtype=torch.float32, device=device)
        barrier()
        t0 = time.perf_counter()
        for i in range(10):
            ceil_dst.copy_(host_cols[i % n_host], non_blocking=True)
        barrier()
        tt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        h2d_ceiling = 10 * B * 4 * f / float(tt.item()) / 1e9
        del ceil_dst
        e2e = {"value": args.steps * B * world / dt, "unit": UNIT, "h2d_bytes_per_step": B * (4 * f + 1),
               "h2d_ceiling_gbs_per_gpu": h2d_ceiling,
               "h2d_ceiling_note": "pinned -> device copies of the same buffers issued by all ranks at once (max over ranks): "
                                   "the rate this host sustains per GPU with nothing else running",
               "d2h_bytes_per_step": 4, "ms_per_step": 1e3 * dt / args.steps, "last_loss": last, "repeats": len(dts),
               "h2d_gbs_per_gpu": B * (4 * f + 1) * args.steps / dt / 1e9,
               "api": "wxddp.engine.StepEngine.submit_host(pinned raw fp32 columns, uint8 labels) -> loss of the previous step; drain() for the last (H2D of batch i overlaps the kernel of batch i-1)"}

    # ---- the other configurations BASELINE.json names, in the same line (rows N2, N3, N1 of the verdict's table) ----
    extras = {}
    if not args.no_extras:
        def timed_once(fn, n):
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            ev0.record()
            for i in range(n):
                fn(i)
            ev1.record()
            barrier()
            t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        def step_at(row0, rows):
            if world == 1 or fused_ddp:
                step_raw(params, m, v, state, bound, stats, y, loss, row0=row0, batch=rows, impute=wxddp.IMPUTE_MEAN,
                         dropout=wxddp.DROPOUT_PHILOX, seed=42 + 1000 * rank)
            else:
                grad_fn(params, feats, y, row0=row0, batch=rows, dropout=wxddp.DROPOUT_PHILOX, seed=42 + 1000 * rank,
                        step=step_no[0], scale=1.0 / (world * rows), out=grads)
                dist.all_reduce(grads)
                step_no[0] += 1
                ctx.adam_step(params, grads, m, v, step_no[0], n=P)

        # configs[4]: batch-size sweep, rows per step and GPU from 256 to 1 Mi (latency-bound -> throughput-bound)
        sweep = []
        for rows in (256, 1024, 4096, 16384, 65536, 262144, 1048576):
            if rows > T // 2:
                continue
            nwin = T // rows
            n = 200 if rows <= 65536 else 50
            fn = lambda i, rows=rows, nwin=nwin: step_at(((i * 7919) % nwin) * rows, rows)
            timed_once(fn, 5)
            ts = [timed_once(fn, n) for _ in range(3)]
            t_ms = median(ts)
            sweep.append({"rows_per_step_per_gpu": rows, "rows_per_step": rows * world, "us_per_step": 1e3 * t_ms / n,
                          "samples_per_s": n * rows * world / (t_ms / 1e3)})
        extras["sweep"] = sweep
        # configs[2]: one epoch over a 100 M-row table sharded by rows over the GPUs, 1 Mi rows per step and GPU
        del raw, cols, bound
        feats = None
        torch.cuda.empty_cache()
        shard = args.s3_rows // world // 4 * 4
        raw, y = synth_raw_on_device(torch, shard, f, 4321 + rank, device)
        cols = [raw[j] for j in range(f)]
        bound = ctx.bind_columns(cols)
        stats = ctx.col_stats(cols)
        if world > 1:
            parts = [torch.empty_like(stats) for _ in range(world)]
            dist.all_gather(parts, stats)
            stats = ctx.stats_merge(torch.stack(parts).contiguous())
        n_steps = (shard + B - 1) // B
        if world == 1 or fused_ddp:
            ep = lambda i: step_at(i * B, min(B, shard - i * B) // 4 * 4)
            timed_once(ep, 3)
            ep_ms = [timed_once(ep, n_steps) for _ in range(3)]
            extras["s3_epoch"] = {"rows_total": shard * world, "rows_per_gpu": shard, "steps": n_steps,
                                  "rows_per_step_per_gpu": B, "ms": median(ep_ms), "ms_runs": ep_ms,
                                  "samples_per_s": shard * world / (median(ep_ms) / 1e3), "final_loss": float(loss.item()),
                                  "kernel_status": ctx.tc_status(),
                                  "table": f"{shard} rows x {f} raw fp32 columns per GPU ({shard * f * 4 / 1e9:.1f} GB) resident in HBM"}
        del raw, cols, bound
        torch.cuda.empty_cache()
        if world > 1:
            extras[f"gpu_reference_config_w{world}"] = reference_config_ddp(torch, dist, device, rank, world)
        # configs[3]: the wide MLP, timed for seconds so that the sustained tensor peak is the denominator
        w = wide_measure(args, torch, dist, world, rank, local_rank, device, steps=args.steps,
                         min_seconds=args.wide_seconds, with_e2e=not args.no_e2e)
        extras["wide"] = w

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    value = args.steps * B * world / (ms / 1e3)
    peak, peak_src = read_peaks()
    bytes_per_row = 4 * f + 8                                   # fp32 features + int64 label
    kernel_s = (ms / 1e3) / args.steps                          # one step-kernel launch per step and GPU
    achieved = B * bytes_per_row / kernel_s / 1e9
    flops_per_row = 2 * (2 * f * 64 + 3 * 64 * 2)
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this exact workload
    # (profiles/README.md: 92.46 MB + 2.5 MB); any other workload has no capture -> null
    traffic, traffic_src = profiled_traffic(use_tc and f == 20 and B == (1 << 20))
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "kernel": kernel_name if (world == 1 or fused_ddp)
            else ("tc_step_kernel (wx_tc_mlp_grad)" if grad_fn == ctx.tc_mlp_grad else "mlp_grid_kernel<F,TRAIN,ROWMAJOR> (wx_mlp_grad)")
            + " + NCCL all-reduce + adam_kernel",
            "traffic_source": traffic_src,
            "peak_source": peak_src, "algorithmic_bytes_per_row": bytes_per_row,
            "fp32_pipe": {"achieved_tflops": B * flops_per_row / kernel_s / 1e12, "peak_tflops": 74.4,
                          "frac": B * flops_per_row / kernel_s / 1e12 / 74.4,
                          "note": "148 SM x 128 lanes x 2 x 1.965 GHz; the step is ~67 FLOP/B, above the fp32 ridge"}}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, B),
            "roofline": roof, "clocks": clocks, "gpu_launches": int(launches), "final_loss": final_loss,
            "kernel_status": tc_status, "repeats": len(times), "ms_per_step_min": min(times) / args.steps, "ms_per_step_max": max(times) / args.steps,
            "timing": f"median of {len(times)} repeats of {args.steps} steps each (CUDA events, barrier + synchronize around "
                      f"every repeat, max over ranks); gpu_launches counts one repeat"}
    if e2e is not None:
        line["e2e"] = e2e
    if ddp_check is not None:
        line["ddp_check"] = ddp_check
    line.update(extras)
    if not args.no_cpu_baseline and world == 1:
        rows_cpu = args.cpu_rows_per_step if args.cpu_rows_per_step else B
        r, threads = cpu_baseline(f, max(1, rows_cpu // 2), args.cpu_steps, 2)
        line["cpu_baseline"] = {"value": r["samples_per_sec"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                "sample": f"{args.cpu_steps} steps x {max(1, rows_cpu // 2)} rows/rank x 2 ranks "
                                          f"(gloo DDP, localhost, {threads} threads/rank, tensor-sliced batches), "
                                          f"{r['seconds']:.1f} s of CPU work; host has {os.cpu_count()} cores"}
        # the reference's literal configuration (batch 4 per rank, per-sample loader), for the record
        from oracle import train_ref
        r4 = train_ref.time_cpu_ddp(f=f, batch=4, steps=1500, warmup=50, world=2, threads=1, per_sample=True)
        line["cpu_reference_config"] = {"value": r4["samples_per_sec"], "unit": UNIT, "cores": 2,
                                        "sample": "1500 steps x 4 rows/rank x 2 ranks, 1 thread/rank, per-sample "
                                                  "__getitem__+collate (train_lightning_ddp.py:49,122)"}
    if world == 1 and not args.no_cpu_baseline:
        line["gpu_reference_config"] = reference_config_on_gpu(torch, wxddp, device)
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def reference_config_ddp(torch, dist, device, rank: int, world: int) -> dict:
    """The reference's own DDP regime on GPUs: WORLD_SIZE ranks x 4 rows per step (train_lightning_ddp.py:122,:136,
    docker-compose.yml:124 has WORLD_SIZE=2), 145,460 x 5 table, sampler-ordered rows, 2048 optimizer steps per launch of
    the resident kernel with the gradient exchange inside it.  Printed as gpu_reference_config_w<N>."""
    from wxddp.engine import StepEngine

    f, n, batch = 5, 116_368, 4
    g = torch.Generator(device=device)
    g.manual_seed(7)
    x = torch.randn(n, f, device=device, generator=g)
    y = (torch.rand(n, device=device, generator=g) < 0.13).to(torch.int64)
    per_rank = n // world // batch * batch
    order = torch.randperm(n, device=device, generator=g)[rank * per_rank:(rank + 1) * per_rank].contiguous()
    torch.manual_seed(42)
    eng = StepEngine(f, device=device, world=world, rank=rank, process_group=dist)
    eng.connect_peers()
    per_launch = 2048
    n_steps = per_rank // batch
    buf = torch.empty(per_launch, device=device)

    def epoch():
        s = 0
        while s < n_steps:
            k = min(per_launch, n_steps - s)
            eng.steps_small(x, y, order[s * batch:(s + k) * batch], k, batch, buf)
            s += k

    epoch()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    epoch()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    eng.check()
    return {"value": n_steps * batch * world / (ms / 1e3), "unit": UNIT, "us_per_step": 1e3 * ms / n_steps, "world": world,
            "sample": f"one epoch: {n_steps} optimizer steps of {batch} rows per rank x {world} ranks (F=5), {per_launch} steps "
                      f"per launch of wx_train_steps_small, [P+1] bucket exchanged over NVLink inside the kernel every step, "
                      f"{ms:.1f} ms", "last_loss": float(buf[0].item())}


def reference_config_on_gpu(torch, wxddp, device) -> dict:
    """The reference's literal shape (configs[0]: 145,460 rows x 5 features, global batch 8 = 2 ranks x 4,
    train_lightning_ddp.py:122 + docker-compose.yml:124) on ONE B200: sampler-ordered rows, 2048 optimizer steps per
    launch of the resident small-step kernel (wx_train_steps_small).  Reported next to cpu_reference_config."""
    f, n, batch = 5, 116_368, 8
    ctx = wxddp.Context(f, device=device)
    g = torch.Generator(device=device)
    g.manual_seed(7)
    x = torch.randn(n, f, device=device, generator=g)
    y = (torch.rand(n, device=device, generator=g) < 0.13).to(torch.int64)
    order = torch.randperm(n, device=device, generator=g)
    torch.manual_seed(42)
    from torch import nn
    net = nn.Sequential(nn.Linear(f, 64), nn.ReLU(), nn.Dropout(0.2), nn.Linear(64, 2))
    p = torch.cat([q.detach().reshape(-1) for q in net.parameters()]).to(device)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    state = torch.zeros(2, dtype=torch.int64, device=device)
    per_launch = 2048
    n_steps = n // batch
    buf = torch.empty(per_launch, device=device)

    def epoch():
        s = 0
        while s < n_steps:
            k = min(per_launch, n_steps - s)
            ctx.train_steps_small(p, m, v, state, x, y, buf, n_steps=k, batch=batch,
                                  row_idx=order[s * batch:(s + k) * batch], dropout=wxddp.DROPOUT_PHILOX, seed=42)
            s += k

    epoch()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    epoch()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    return {"value": n_steps * batch / (ms / 1e3), "unit": UNIT, "us_per_step": 1e3 * ms / n_steps,
            "sample": f"one epoch: {n_steps} optimizer steps of {batch} sampler-ordered rows (F=5), {per_launch} steps per "
                      f"launch of wx_train_steps_small, {ms:.1f} ms", "last_loss": float(buf[0].item())}


if __name__ == "__main__":
    sys.exit(main())
