import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, wxddp
f = int(os.environ.get("F", 20)); B = int(os.environ.get("B", 1 << 20)); mode = int(os.environ.get("MODE", 2)); steps = int(os.environ.get("STEPS", 6))
c = wxddp.Context(f)
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(4 * B, f, device="cuda", generator=g); y = (torch.rand(4 * B, device="cuda", generator=g) < 0.2).long()
p = torch.randn(c.P, device="cuda", generator=g) * 0.2; m = torch.zeros_like(p); v = torch.zeros_like(p)
st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
fn = c.tc_train_step if os.environ.get("IMPL", "tc") == "tc" else c.train_step
for i in range(3): fn(p, m, v, st, x, y, loss, row0=(i % 4) * B, batch=B, dropout=mode, seed=3)
torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(steps): fn(p, m, v, st, x, y, loss, row0=(i % 4) * B, batch=B, dropout=mode, seed=3)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print(f"impl={os.environ.get('IMPL','tc')} mode={mode} F={f} B={B}: {ms*1e3:.1f} us/step {B/ms/1e6:.2f} G rows/s loss={loss.item():.4f} status={c.tc_status()}")
if os.environ.get("CTA_TIMES"):
    import numpy as np
    dbg = torch.zeros(1184 * 4, dtype=torch.int64, device="cuda")
    c.lib.wxdev_cta_times(c._h, dbg.data_ptr())
    fn(p, m, v, st, x, y, loss, row0=0, batch=B, dropout=mode, seed=3); torch.cuda.synchronize()
    d = dbg.cpu().numpy().reshape(-1, 4); d = d[d[:, 1] > 0]; t0 = d[:, 1].min()
    loop = (d[:, 2] - d[:, 1]) / 1e3; ex = (d[:, 3] - t0) / 1e3
    print("ctas", len(d), "main us min/med/max %.1f %.1f %.1f" % (loop.min(), np.median(loop), loop.max()), "exit max %.1f" % ex.max(),
          "slowest", [(int(i), int(d[i, 0]), round(float(loop[i]), 1)) for i in np.argsort(loop)[-4:]])
if os.environ.get("CTA_TIMES"):
    dbg2 = torch.zeros(1184 * 4 + 148 * 2 * 8 + 148 * 8, dtype=torch.int64, device="cuda")
    c.lib.wxdev_cta_times(c._h, dbg2.data_ptr())
    fn(p, m, v, st, x, y, loss, row0=0, batch=B, dropout=mode, seed=3); torch.cuda.synchronize()
    ph = dbg2[1184 * 4:1184 * 4 + 148 * 16].cpu().numpy().reshape(148, 2, 8)
    mm = dbg2[1184 * 4 + 148 * 16:].cpu().numpy().reshape(148, 8)
    names = ["wait_d1", "rng+y", "phaseA+softmax", "wait_a2free", "xm_row", "phaseB", "wait_xk+convert", "loop_gap"]
    tiles = (B // 128) / 148
    for g in range(2):
        print("group", g, {n: int(ph[:, g, k].mean() / (tiles / 2)) for k, n in enumerate(names)}, "cycles per tile (avg over CTAs)")
    print("mma thread cycles/tile:", {n: int(mm[:, k].mean() / tiles) for k, n in enumerate(["wait_xk", "issue_g1", "wait_a2ready", "issue_g2"])},
          "sampled completion latency: g1", int(mm[:, 4].mean() / (tiles / 8)), "g2", int(mm[:, 5].mean() / (tiles / 8)))
