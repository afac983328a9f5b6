import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, wxddp
c = wxddp.Context(20)
g = torch.Generator().manual_seed(3)
x = torch.randn(128, 32, generator=g); w = torch.randn(64, 32, generator=g) * 0.3; dh = torch.randn(128, 64, generator=g) * 0.01
x[:, 24:] = 0
zref = (x.double() @ w.double().T).numpy(); gref = (dh.double().T @ x.double()).numpy()
for terms in (1, 3, 11, 13):
    z = torch.full((128, 64), float("nan"), device="cuda"); gg = torch.full((64, 32), float("nan"), device="cuda")
    st = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    xd, wd, dd = x.cuda().contiguous(), w.cuda().contiguous(), dh.cuda().contiguous()
    rc = c.lib.wxdev_tc_selftest(c._h, xd.data_ptr(), wd.data_ptr(), dd.data_ptr(), terms, z.data_ptr(),
                              gg.data_ptr(), st.data_ptr(), None)
    torch.cuda.synchronize()
    z = z.cpu().numpy(); gg = gg.cpu().numpy()
    ez = np.abs(z - zref).max() / np.abs(zref).max(); eg = np.abs(gg - gref).max() / np.abs(gref).max()
    print(f"terms={terms} rc={rc} status={st.item()} z_err={ez:.3e} g_err={eg:.3e} nan_z={np.isnan(z).sum()} nan_g={np.isnan(gg).sum()}")
    if not ez < 1e-2:
        ok = np.abs(z - zref) < 1e-2 * np.abs(zref).max()
        print(" z ok fraction", ok.mean(), "rows ok", ok.all(1).sum(), "cols ok", ok.all(0).sum())
        print(" z[0,:4]", z[0, :4], "ref", zref[0, :4])
    if not eg < 1e-2:
        ok = np.abs(gg - gref) < 1e-2 * np.abs(gref).max()
        print(" g ok fraction", ok.mean(), "rows ok", ok.all(1).sum(), "cols ok", ok.all(0).sum())
        print(" g[0,:4]", gg[0, :4], "ref", gref[0, :4]); print(" g[1,:4]", gg[1, :4], "ref", gref[1, :4])
