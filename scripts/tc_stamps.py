"""Developer aid (needs a build with WX_NVCC_EXTRA=-DWX_TC_TIMERS=1): per-CTA globaltimer stamps of one step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, wxddp
f = 20; tiles = int(os.environ.get("TILES", 55)); B = 128 * 148 * tiles
c = wxddp.Context(f)
g = torch.Generator(device="cuda").manual_seed(1)
cols_t = torch.randn(f, 2 * B, device="cuda", generator=g) * 5 + 30
y = (torch.rand(2 * B, device="cuda", generator=g) < 0.2).long()
cols = [cols_t[j] for j in range(f)]; stats = c.col_stats(cols); bound = c.bind_columns(cols)
p = torch.randn(c.P, device="cuda", generator=g) * 0.2; m = torch.zeros_like(p); v = torch.zeros_like(p)
st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
dbg = torch.zeros(148 * 8 + 148 * 40, dtype=torch.int64, device="cuda")
c.lib.wxdev_cta_times(c._h, dbg.data_ptr())
for i in range(6):
    c.tc_train_step_raw(p, m, v, st, bound, stats, y, loss, row0=(i % 2) * B, batch=B, impute=1, dropout=2, seed=3)
torch.cuda.synchronize()
d = dbg.cpu().numpy()[:148 * 8].reshape(148, 8).astype(np.float64)
t0 = d[:, 0].min()
names = ["start", "prologue done", "main loop done", "slab ready", "after ticket 1", "after ticket 2", "adam done"]
for k, n in enumerate(names):
    col = d[:, k]; col = col[col > 0]
    if len(col): print(f"{n:16s} n={len(col):3d}  min {(col.min()-t0)/1e3:7.2f}  med {(np.median(col)-t0)/1e3:7.2f}  max {(col.max()-t0)/1e3:7.2f} us")
ph = dbg.cpu().numpy()
if ph.size >= 148 * 8 + 148 * 5 * 8:
    ph = ph[148 * 8:148 * 8 + 148 * 40].reshape(148, 5, 8).astype(np.float64)
    names = {0: ["wait_pw", "wait_a2free", "wait_d1", "phaseA", "softmax", "phaseB", "U+arrive", "loop"],
             1: ["wait_pw", "wait_a2free", "wait_d1", "phaseA", "softmax", "phaseB", "U+arrive", "loop"],
             2: ["philox+handoff", "wait_land", "wait_xa_free", "convert", "-", "-", "-", "loop"],
             3: ["wait_xk", "issue_g1", "wait_a2ready", "issue_g2", "-", "-", "-", "pre"],
             4: ["wait_empty", "issue", "-", "-", "-", "-", "-", "loop"]}
    for role, label in ((0, "workers g0"), (1, "workers g1"), (2, "helpers"), (3, "mma warp"), (4, "producer")):
        n_tiles = tiles / 2 if role < 2 else tiles
        print(label, {nm: int(ph[:, role, k].mean() / n_tiles) for k, nm in enumerate(names[role]) if nm != "-"}, "cycles per own tile")
