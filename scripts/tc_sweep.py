"""Fixed cost vs per-tile cost of the tensor-core step kernel: time wx_tc_train_step at 1, 2, 4, ... tiles per CTA."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, wxddp
f = int(os.environ.get("F", 20)); mode = int(os.environ.get("MODE", 2)); raw = int(os.environ.get("RAW", 1))
c = wxddp.Context(f)
g = torch.Generator(device="cuda").manual_seed(1)
NMAX = 128 * 148 * 64
x = torch.randn(NMAX * 2, f, device="cuda", generator=g); y = (torch.rand(NMAX * 2, device="cuda", generator=g) < 0.2).long()
cols_t = (x.t().contiguous() * 5 + 30)
cols = [cols_t[j] for j in range(f)]
stats = c.col_stats(cols)
bound = c.bind_columns(cols)
p = torch.randn(c.P, device="cuda", generator=g) * 0.2; m = torch.zeros_like(p); v = torch.zeros_like(p)
st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
res = []
for tiles in (1, 2, 4, 8, 16, 32, 55, 64):
    B = 128 * 148 * tiles
    def step(i):
        if raw:
            c.tc_train_step_raw(p, m, v, st, bound, stats, y, loss, row0=(i % 2) * B, batch=B, impute=1, dropout=mode, seed=3)
        else:
            c.tc_train_step(p, m, v, st, x, y, loss, row0=(i % 2) * B, batch=B, dropout=mode, seed=3)
    for i in range(5): step(i)
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    n = 200
    e0.record()
    for i in range(n): step(i)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    res.append((tiles, us))
    print(f"tiles/CTA {tiles:3d} rows {B:8d}: {us:7.2f} us/step  status={c.tc_status()}")
(t1, u1), (t2, u2) = res[-3], res[-1]
per = (u2 - u1) / (t2 - t1)
print(f"per tile {per:.3f} us ({per * 1.965e3:.0f} cycles at 1965 MHz), fixed {res[0][1] - per:.2f} us")
