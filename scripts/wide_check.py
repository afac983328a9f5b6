"""GPU bring-up check of the wide path: the three GEMM layouts against torch, then one fwd/bwd step against the oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from wxddp.wide import WideEngine
from oracle import wide_ref
torch.manual_seed(0)
dev = torch.device("cuda", 0)
H = int(os.environ.get("H", 512)); B = int(os.environ.get("B", 384)); F = 20; L = 3
eng = WideEngine(F, H, L, max_batch=max(B, 512), device=dev)
print("P", eng.P, "status", eng.status())
def rel(a, b): return float((a - b).norm() / b.norm())
for (M, N, K) in ((256, 512, 128), (384, 256, 64), (128, 768, 512), (4096, 64, 384)):
    a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
    bias = torch.randn(N, device=dev)
    want = a.float() @ b.float().t()
    got = eng.gemm(a, b, True, True, bias=bias); torch.cuda.synchronize()
    print(f"KK  {M}x{N}x{K}: rel {rel(got, want + bias):.2e} status {eng.status()}")
    bt = b.t().contiguous()            # [K, N]
    got = eng.gemm(a, bt, True, False); torch.cuda.synchronize()
    print(f"K,MN {M}x{N}x{K}: rel {rel(got, want):.2e} status {eng.status()}")
    at = a.t().contiguous()            # [K, M]
    got = eng.gemm(at, bt, False, False); torch.cuda.synchronize()
    print(f"MN,MN {M}x{N}x{K}: rel {rel(got, want):.2e} status {eng.status()}")
# full step vs oracle
x = torch.randn(B, F); y = (torch.rand(B) < 0.3).long()
flat = eng.params.cpu()
for drop in (0, 2):
    g = eng.loss_and_grads(x.to(dev), y.to(dev), dropout=drop, step=3).cpu()
    want = wide_ref.loss_and_grads(flat, x, y, F, H, L, dropout=(drop == 2), seed=42, step=3)
    print(f"dropout={drop}: loss {float(g[-1]):.6f} vs {float(want[-1]):.6f}; grads rel-L2 {rel(g[:-1], want[:-1]):.2e}; status {eng.status()}")
    off = 0
    for l, (lo, hi) in enumerate(eng.slices + [eng.head_slice]):
        hi = min(hi, eng.P)
        print(f"   slice {l}: rel {rel(g[lo:hi], want[lo:hi]):.2e}")
