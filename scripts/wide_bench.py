"""Wide MLP (20 -> 4096 x3 -> 2) step time and tensor-pipe fraction on one B200, plus the three GEMM shapes on their own."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from wxddp.wide import WideEngine
dev = torch.device("cuda", 0)
H = int(os.environ.get("H", 4096)); L = 3; F = 20
PEAK = float(os.environ.get("PEAK", 1416.4))
Bs = [int(b) for b in os.environ.get("BS", "4096,16384").split(",")]
eng = WideEngine(F, H, L, max_batch=max(Bs), device=dev)
def timeit(fn, reps=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for B in Bs:
    a = torch.randn(B, H, device=dev).bfloat16(); w = torch.randn(H, H, device=dev).bfloat16(); dz = torch.randn(B, H, device=dev).bfloat16()
    fl = 2.0 * B * H * H
    for name, fn in (("fwd  A[B,K] W[N,K]  (K,K)", lambda: eng.gemm(a, w, True, True)),
                     ("dX   dZ[B,K] W[K,N]  (K,MN)", lambda: eng.gemm(dz, w, True, False)),
                     ("dW   dZ[K,M] H[K,N]  (MN,MN)", lambda: eng.gemm(dz, a, False, False))):
        ms = timeit(fn)
        print(f"B={B:6d} {name}: {ms:7.3f} ms  {fl/ms/1e9:7.1f} TFLOP/s  {100*fl/ms/1e9/PEAK:5.1f}% of {PEAK}")
    ms = timeit(lambda: torch.matmul(a, w.t()))
    print(f"B={B:6d} cuBLAS bf16 same shape     : {ms:7.3f} ms  {fl/ms/1e9:7.1f} TFLOP/s")
    del a, w, dz
    if os.environ.get('ONLY_GEMM'): continue
    x = torch.randn(B, F, device=dev); y = (torch.rand(B, device=dev) < 0.3).long()
    l0 = eng.launch_count
    ms = timeit(lambda: eng.step(x, y), reps=10)
    nl = (eng.launch_count - l0) / 13
    weights = F * H + (L - 1) * H * H + 2 * H
    flops = (6.0 * weights - 2.0 * F * H) * B
    print(f"B={B:6d} full step: {ms:7.3f} ms  {B/ms*1e3:.4e} samples/s  {flops/ms/1e9:7.1f} TFLOP/s ({100*flops/ms/1e9/PEAK:5.1f}% of {PEAK})  {nl:.0f} launches/step  loss {float(eng.grads[-1]):.4f} status {eng.status()}")
