import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, wxddp
from oracle import train_ref
for f in (5, 20):
    c = wxddp.Context(f)
    g = torch.Generator().manual_seed(5)
    N = 200_000
    x = torch.randn(N, f, generator=g); y = (torch.rand(N, generator=g) < 0.2).long()
    p = torch.randn(c.P, generator=g) * 0.2
    xd, yd, pd = x.cuda(), y.cuda(), p.cuda()
    for n in (8, 64, 1000, 9472, 9480, 9472*2, 9472*2+8, 20000, 50000, 200000):
        out = c.mlp_grad(pd, xd, yd, row0=0, batch=n).cpu().numpy()
        loss, grad = train_ref.loss_and_grads(p, f, x[:n], y[:n], None)
        gerr = np.abs(out[:-1] - grad.numpy()).max() / np.abs(grad.numpy()).max()
        print(f"F={f} n={n} loss {out[-1]:.6f} ref {loss.item():.6f} rel {abs(out[-1]-loss.item())/loss.item():.2e} grad_err {gerr:.2e} launches {c.launch_count}", flush=True)
    # per-row losses through eval on single rows to find bad rows
    ev = c.mlp_eval(pd, xd, yd, row0=0, batch=50000).cpu().numpy()
    l2, c2 = train_ref.eval_batch(p, f, x[:50000], y[:50000])
    print(f"F={f} eval50000 loss {ev[0]/50000:.6f} ref {l2.item():.6f} correct {ev[1]} ref {c2}")
