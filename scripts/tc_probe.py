import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, wxddp
c = wxddp.Context(20)
def run(x, w, dh, terms=1):
    z = torch.full((128, 64), float("nan"), device="cuda"); gg = torch.full((64, 32), float("nan"), device="cuda")
    st = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    xd, wd, dd = x.cuda().contiguous(), w.cuda().contiguous(), dh.cuda().contiguous()
    rc = c.lib.wxdev_tc_selftest(c._h, xd.data_ptr(), wd.data_ptr(), dd.data_ptr(), terms, z.data_ptr(), gg.data_ptr(), st.data_ptr(), None)
    torch.cuda.synchronize()
    return z.cpu().numpy(), gg.cpu().numpy(), st.item()
np.set_printoptions(linewidth=200, precision=3, suppress=True)
ones_x = torch.ones(128, 32); ones_w = torch.ones(64, 32); ones_dh = torch.ones(128, 64)
z, g, st = run(ones_x, ones_w, ones_dh)
print("P1 all ones: status", st, "z uniq", np.unique(z)[:10], "g uniq", np.unique(g)[:10])
# P2: x[r][c] = r+1 ; w = e_0
x = (torch.arange(128).float() + 1)[:, None].repeat(1, 32); w = torch.zeros(64, 32); w[:, 0] = 1
z, g, st = run(x, w, ones_dh)
print("P2 row map: z[:,0] first 20:", z[:20, 0], " z[0,:8]", z[0, :8])
# P3: x ones, w[n][c] = c+1 -> sum of (c+1) read ; full = 528
w = (torch.arange(32).float() + 1)[None, :].repeat(64, 1)
z, g, st = run(ones_x, w, ones_dh)
print("P3 k coverage: z uniq", np.unique(z)[:10])
# P4: x ones, w[n][c] = (n+1) if c==0 -> z[r][n] = n+1
w = torch.zeros(64, 32); w[:, 0] = torch.arange(64).float() + 1
z, g, st = run(ones_x, w, ones_dh)
print("P4 n map: z[0,:16]", z[0, :16], "z[5,:8]", z[5, :8])
# P5: x[r][c] = 1 only at c==k0 for several k0, w[n][c] = c+1: which c is paired
for k0 in (0, 1, 4, 8, 9, 31):
    x = torch.zeros(128, 32); x[:, k0] = 1
    w = (torch.arange(32).float() + 1)[None, :].repeat(64, 1)
    z, g, st = run(x, w, ones_dh)
    print("P5 k0", k0, "z[0..9,0]", z[:10, 0])
# GEMM2 probes: g[m][n] = sum_k dh[k][m] x[k][n]
z, g, st = run(ones_x, ones_w, ones_dh)
print("G1 all ones g[:4,:8]", g[:4, :8], "rows 16-20 col0", g[16:20, 0], "uniq", np.unique(g)[:8])
dh = torch.zeros(128, 64); dh[:, 0] = 1.0   # only m=0
x = (torch.arange(32).float() + 1)[None, :].repeat(128, 1)   # x[k][n] = n+1
z, g, st = run(x, ones_w, dh)
print("G2 m=0 only, x[k][n]=n+1: g[0,:8]", g[0, :8], "g[1,:4]", g[1, :4], "nonzero rows", np.nonzero(np.abs(g).sum(1))[0][:10])
