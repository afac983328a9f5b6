import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, wxddp
from oracle import train_ref, philox_ref
torch.manual_seed(0)
def check(f, n, row0, mode, seed=11, step=5):
    c = wxddp.Context(f)
    g = torch.Generator().manual_seed(f * 1000 + n)
    N = row0 + n
    x = torch.randn(N, f, generator=g); y = (torch.rand(N, generator=g) < 0.3).long()
    p = torch.randn(c.P, generator=g) * 0.2
    xd, yd, pd = x.cuda(), y.cuda(), p.cuda()
    mask = None; mk = None
    if mode == wxddp.DROPOUT_MASK:
        keep = torch.rand(n, 64, generator=g) < 0.8
        mask = keep.float()
        mk = torch.from_numpy(np.packbits(keep.numpy(), axis=1, bitorder="little").view(np.int64).reshape(-1).copy()).cuda()
    elif mode == wxddp.DROPOUT_PHILOX:
        mask = torch.from_numpy(philox_ref.dropout_keep_mask(n, seed, step).astype(np.float32))
    out = c.tc_mlp_grad(pd, xd, yd, row0=row0, batch=n, dropout=mode, keep_mask=mk, seed=seed, step=step)
    torch.cuda.synchronize()
    st = c.tc_status()
    out = out.cpu().numpy()
    v1 = c.mlp_grad(pd, xd, yd, row0=row0, batch=n, dropout=mode, keep_mask=mk, seed=seed, step=step).cpu().numpy()
    loss, grad = train_ref.loss_and_grads(p, f, x[row0:row0 + n], y[row0:row0 + n], mask)
    gr = grad.numpy()
    den = np.abs(gr).max()
    e_or = np.abs(out[:-1] - gr).max() / den; e_v1 = np.abs(out[:-1] - v1[:-1]).max() / den; e_v1o = np.abs(v1[:-1] - gr).max() / den
    ok = np.allclose(out[:-1], gr, rtol=1e-5, atol=1e-7) and abs(out[-1] - loss.item()) <= 1e-5 * abs(loss.item())
    print(f"F={f} n={n} row0={row0} mode={mode} status={st} loss tc={out[-1]:.7f} ref={loss.item():.7f} v1={v1[-1]:.7f} "
          f"graderr tc-oracle={e_or:.2e} tc-v1={e_v1:.2e} v1-oracle={e_v1o:.2e} parity={'OK' if ok else 'FAIL'}", flush=True)
    if not ok:
        P = c.P; bad = np.argsort(-np.abs(out[:-1] - gr))[:5]
        print("   worst idx", bad, "tc", out[bad], "ref", gr[bad])
    return ok
allok = True
for f, n, row0, mode in [(20, 128, 0, 0), (20, 128, 0, 1), (20, 100, 0, 0), (5, 256, 0, 0), (20, 1000, 128, 1), (20, 128 * 148 * 3 + 77, 0, 0),
                         (20, 4096, 0, 2), (5, 70001, 4, 2), (29, 3000, 0, 1), (1, 500, 0, 0), (13, 999, 12, 2)]:
    try:
        allok &= check(f, n, row0, mode)
    except Exception as e:
        print("EXC", f, n, row0, mode, e); allok = False
print("ALL OK" if allok else "SOME FAILED")
# timing
f, B = 20, 1 << 20
c = wxddp.Context(f)
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(8 * B, f, device="cuda", generator=g); y = (torch.rand(8 * B, device="cuda", generator=g) < 0.2).long()
p = torch.randn(c.P, device="cuda", generator=g) * 0.2; m = torch.zeros_like(p); v = torch.zeros_like(p)
st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
for name, fn in [("tc", c.tc_train_step), ("v1", c.train_step)]:
    for mode in (0, 2):
        for i in range(3): fn(p, m, v, st, x, y, loss, row0=(i % 8) * B, batch=B, dropout=mode, seed=3)
        torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(20): fn(p, m, v, st, x, y, loss, row0=(i % 8) * B, batch=B, dropout=mode, seed=3)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"{name} mode={mode} F={f} B={B}: {ms*1e3:.1f} us/step  {B/ms/1e6:.2f} G rows/s  loss={loss.item():.4f} status={c.tc_status()}", flush=True)
# raw-column fused step: tc vs v1 (one Adam step from zero moments: m = 0.1 * grad)
from oracle import synth, preprocess_ref
for f, n, r0 in [(20, 30000, 0), (5, 9001, 4), (20, 128 * 148 + 64, 128)]:
    t = synth.synth_table(5, n + r0, f, null_rate=0.01, const_col=(f - 1 if f > 5 else None))
    c = wxddp.Context(f)
    cols = [torch.from_numpy(t["raw"][j].astype(np.float32)).cuda() for j in range(f)]
    stats = c.col_stats(cols)
    yy = torch.from_numpy(preprocess_ref.label_encode(t["labels"]).astype(np.int64)).cuda()
    p0 = (torch.randn(c.P, generator=torch.Generator().manual_seed(1)) * 0.2)
    res = []
    for fn in (c.train_step_raw, c.tc_train_step_raw):
        p = p0.cuda(); m = torch.zeros_like(p); v = torch.zeros_like(p); st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
        fn(p, m, v, st, cols, stats, yy, loss, row0=r0, batch=n, impute=1, dropout=2, seed=9)
        torch.cuda.synchronize()
        res.append((loss.item(), m.cpu().numpy(), p.cpu().numpy()))
    dm = np.abs(res[0][1] - res[1][1]).max() / np.abs(res[0][1]).max()
    print(f"raw F={f} n={n} row0={r0}: loss v1={res[0][0]:.7f} tc={res[1][0]:.7f} m relerr={dm:.2e} status={c.tc_status()}", flush=True)
f, B = 20, 1 << 20
c = wxddp.Context(f)
raw = torch.randn(f, 8 * B, device="cuda") * 5 + 20; cols = [raw[j] for j in range(f)]
stats = c.col_stats(cols); yy = (torch.rand(8 * B, device="cuda") < 0.2).long()
p = torch.randn(c.P, device="cuda") * 0.2; m = torch.zeros_like(p); v = torch.zeros_like(p)
st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
for name, fn in [("tc_raw", c.tc_train_step_raw), ("v1_raw", c.train_step_raw)]:
    for i in range(3): fn(p, m, v, st, cols, stats, yy, loss, row0=(i % 8) * B, batch=B, impute=1, dropout=2, seed=3)
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20): fn(p, m, v, st, cols, stats, yy, loss, row0=(i % 8) * B, batch=B, impute=1, dropout=2, seed=3)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"{name} philox F={f} B={B}: {ms*1e3:.1f} us/step  {B/ms/1e6:.2f} G rows/s  loss={loss.item():.4f} status={c.tc_status()}", flush=True)
