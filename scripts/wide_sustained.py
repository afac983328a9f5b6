"""Sustained wide-MLP step time (20 -> 4096 x3 -> 2, 16,384 rows): median over repeats of 20 steps filling SECS seconds."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from wxddp.wide import WideEngine
dev = torch.device("cuda", 0)
H = int(os.environ.get("H", 4096)); B = int(os.environ.get("B", 16384)); F = 20; SECS = float(os.environ.get("SECS", 4))
eng = WideEngine(F, H, 3, max_batch=B, device=dev, dropout_p=float(os.environ.get('DROP', 0.2)))
x = torch.randn(8 * B, F, device=dev); y = (torch.rand(8 * B, device=dev) < 0.3).long()
n = [0]
def step():
    w = n[0] % 8; eng.step(x[w * B:(w + 1) * B], y[w * B:(w + 1) * B]); n[0] += 1
for _ in range(5): step()
torch.cuda.synchronize()
times = []; total = 0.0
while total < SECS * 1e3:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): step()
    e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1); times.append(t / 20); total += t
weights = F * H + 2 * H * H + 2 * H
flops = (6.0 * weights - 2.0 * F * H) * B
ms = statistics.median(times)
print(f"{os.environ.get('TAG','')} sustained {ms:.4f} ms/step (first {times[0]:.4f}, min {min(times):.4f})  {flops/ms/1e9:.0f} TFLOP/s  status {eng.status()}", flush=True)
