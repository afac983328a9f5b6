import os, sys, subprocess, time, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from wxddp.wide import WideEngine
dev = torch.device("cuda", 0)
H = 4096; B = 16384
eng = WideEngine(20, H, 3, max_batch=B, device=dev)
a = torch.randn(B, H, device=dev).bfloat16(); w = torch.randn(H, H, device=dev).bfloat16()
def run(name, fn, secs=3.0):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,temperature.gpu", "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
    t0 = time.time(); n = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < secs:
        for _ in range(50): fn()
        n += 50
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    p.terminate(); out = p.communicate()[0].strip().splitlines()
    rows = [[float(x) for x in l.split(",")] for l in out if l.count(",") == 2]
    rows = rows[len(rows)//3:]
    clk = statistics.median(r[0] for r in rows); pw = statistics.median(r[1] for r in rows); tc = max(r[2] for r in rows)
    fl = 2.0 * B * H * H
    ideal = fl / (148 * 8192 * clk * 1e6) * 1e3
    print(f"{name}: {ms:.4f} ms {fl/ms/1e9:.0f} TF  clk {clk:.0f} MHz  power {pw:.0f} W  temp {tc:.0f}  tensor-bound at this clock {ideal:.4f} ms -> {100*ideal/ms:.1f}%", flush=True)
run("cublas", lambda: torch.matmul(a, w.t()))
run("mine fwd", lambda: eng.gemm(a, w, True, True))
