import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, wxddp
f = int(sys.argv[1]) if len(sys.argv) > 1 else 20
B = 1 << 20
c = wxddp.Context(f)
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(4 * B, f, device="cuda", generator=g); y = (torch.rand(4 * B, device="cuda", generator=g) < 0.2).long()
p = (torch.randn(c.P, device="cuda", generator=g) * 0.2); m = torch.zeros_like(p); v = torch.zeros_like(p)
st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
dbg = torch.zeros(1184 * 4, dtype=torch.int64, device="cuda")
for i in range(3):
    c.train_step(p, m, v, st, x, y, loss, row0=i * B, batch=B, seed=1)
c.lib.wxdev_cta_times(c._h, dbg.data_ptr())
c.train_step(p, m, v, st, x, y, loss, row0=3 * B, batch=B, seed=1)
torch.cuda.synchronize()
d = dbg.cpu().numpy().reshape(-1, 4)
d = d[d[:, 1] > 0]
t0 = d[:, 1].min()
print("ctas", len(d), "distinct sms", len(set(d[:, 0])))
start = (d[:, 1] - t0) / 1e3; loop = (d[:, 2] - d[:, 1]) / 1e3; ex = (d[:, 3] - t0) / 1e3
print("start us: min %.1f max %.1f" % (start.min(), start.max()))
print("main loop us: min %.1f med %.1f max %.1f" % (loop.min(), np.median(loop), loop.max()))
print("exit us: min %.1f med %.1f max %.1f" % (ex.min(), np.median(ex), ex.max()))
order = np.argsort(loop)
print("slowest ctas (cta, smid, loop us):", [(int(i), int(d[i, 0]), round(float(loop[i]), 1)) for i in order[-8:]])
print("fastest:", [(int(i), int(d[i, 0]), round(float(loop[i]), 1)) for i in order[:4]])
hist, edges = np.histogram(loop, bins=10); print("hist", hist.tolist(), [round(e) for e in edges.tolist()])
