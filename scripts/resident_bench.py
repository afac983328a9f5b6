"""us/step of the resident multi-step kernel (wx_train_steps_small, batch > 64) vs one launch per step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, wxddp
F = int(os.environ.get("F", 20)); dev = torch.device("cuda", 0)
c = wxddp.Context(F, device=dev)
g = torch.Generator(device=dev).manual_seed(1)
n = 1 << 22
x = torch.randn(n, F, device=dev, generator=g); y = (torch.rand(n, device=dev, generator=g) < 0.2).long()
for B in (8, 64, 256, 1024, 4096, 8192, 16384):
    steps = 200
    p = torch.randn(c.P, device=dev, generator=g) * 0.2; m = torch.zeros_like(p); v = torch.zeros_like(p)
    st = torch.zeros(2, dtype=torch.int64, device=dev); losses = torch.zeros(steps, device=dev); l1 = torch.zeros(1, device=dev)
    def resident(): c.train_steps_small(p, m, v, st, x, y, losses, n_steps=steps, batch=B, seed=3)
    def single():
        for s in range(steps): c.train_step(p, m, v, st, x, y, l1, row0=s * B, batch=B, seed=3)
    res = {}
    for name, fn in (("resident", resident), ("per-launch", single)):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        res[name] = 1e3 * e0.elapsed_time(e1) / steps
    print(f"F={F} B={B:6d}: resident {res['resident']:7.2f} us/step ({B/res['resident']:.1f} M rows/s)   per-launch FFMA {res['per-launch']:7.2f} us/step   status {c.tc_status()}")
