import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, wxddp
from oracle import train_ref, philox_ref
def check(f, n, row0, mode, seed=11, step=5, philox_as_mask=False):
    c = wxddp.Context(f)
    g = torch.Generator().manual_seed(f * 1000 + n)
    N = row0 + n
    x = torch.randn(N, f, generator=g); y = (torch.rand(N, generator=g) < 0.3).long()
    p = torch.randn(c.P, generator=g) * 0.2
    xd, yd, pd = x.cuda(), y.cuda(), p.cuda()
    mask = None; mk = None
    if mode == 2 or philox_as_mask:
        keepnp = philox_ref.dropout_keep_mask(n, seed, step)
        mask = torch.from_numpy(keepnp.astype(np.float32))
        if philox_as_mask:
            mk = torch.from_numpy(np.packbits(keepnp, axis=1, bitorder="little").view(np.int64).reshape(-1).copy()).cuda(); mode = 1
    out = c.tc_mlp_grad(pd, xd, yd, row0=row0, batch=n, dropout=mode, keep_mask=mk, seed=seed, step=step).cpu().numpy()
    v1 = c.mlp_grad(pd, xd, yd, row0=row0, batch=n, dropout=mode, keep_mask=mk, seed=seed, step=step).cpu().numpy()
    loss, grad = train_ref.loss_and_grads(p, f, x[row0:row0 + n], y[row0:row0 + n], mask)
    gr = grad.numpy(); den = np.abs(gr).max()
    print(f"F={f} n={n} row0={row0} mode={mode} pm={philox_as_mask} st={c.tc_status()} tc-oracle={np.abs(out[:-1]-gr).max()/den:.2e} v1-oracle={np.abs(v1[:-1]-gr).max()/den:.2e} loss {out[-1]:.7f}/{loss.item():.7f}", flush=True)
for n in (70001, 65536, 65540, 66000, 131072 + 5):
    check(5, n, 4, 2)
check(5, 70001, 4, 0); check(5, 70001, 4, 2, philox_as_mask=True); check(5, 70001, 0, 2); check(20, 70001, 0, 2); check(20, 300001, 0, 2)
check(5, 70001, 4, 2, seed=12); check(5, 70001, 4, 2, step=6)
