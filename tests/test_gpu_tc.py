"""GPU parity tests of the tensor-core (tcgen05, tf32x3) step kernel, through the C ABI.  Same checkers and the same
1e-5 bounds as tests/test_gpu_parity.py; large batches additionally allow the contribution of a single row per entry
(a pre-activation within rounding of zero flips its ReLU gate -- a measure-zero discontinuity any summation order hits)."""
import os

import numpy as np
import pytest
import torch

from oracle import philox_ref, preprocess_ref, synth, train_ref

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RTOL, ATOL_G = 1e-5, 1e-7


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope="module")
def wx():
    import wxddp
    return wxddp


_CTX = {}


def ctx_for(wx, f):
    if f not in _CTX:
        _CTX[f] = wx.Context(f, device=0)
    return _CTX[f]


@pytest.mark.parametrize("tag", ["F5_B4", "F20_B4", "F20_B256", "F5_B1000"])
def test_tc_grad_matches_reference_module(wx, tag):
    g = np.load(os.path.join(GOLDEN, f"steps_{tag}.npz"))
    f = int(g["f"])
    c = ctx_for(wx, f)
    P = c.P
    for s in range(g["loss"].shape[0]):
        out = c.tc_mlp_grad(dev(g["p_before"][s]), dev(g["x"][s]), dev(g["y"][s]), dropout=wx.DROPOUT_MASK,
                            keep_mask=dev(g["mask"][s].view(np.int64))).cpu().numpy()
        np.testing.assert_allclose(out[P], g["loss"][s], rtol=RTOL)
        np.testing.assert_allclose(out[:P], g["grad"][s], rtol=RTOL, atol=ATOL_G)
    assert c.tc_status() == 0


@pytest.mark.parametrize("tag", ["F20_B256", "F5_B1000"])
def test_tc_fused_step_matches_reference_module(wx, tag):
    g = np.load(os.path.join(GOLDEN, f"steps_{tag}.npz"))
    f = int(g["f"])
    c = ctx_for(wx, f)
    P = c.P
    loss = torch.zeros(1, device="cuda")
    for s in range(g["loss"].shape[0]):
        p = dev(g["p_before"][s])
        m = dev(g["m_after"][s - 1]) if s else torch.zeros(P, device="cuda")
        v = dev(g["v_after"][s - 1]) if s else torch.zeros(P, device="cuda")
        state = torch.tensor([s, 0], dtype=torch.int64, device="cuda")
        c.tc_train_step(p, m, v, state, dev(g["x"][s]), dev(g["y"][s]), loss, dropout=wx.DROPOUT_MASK,
                        keep_mask=dev(g["mask"][s].view(np.int64)))
        assert state.cpu().tolist() == [s + 1, 0]
        np.testing.assert_allclose(loss.item(), g["loss"][s], rtol=RTOL)
        np.testing.assert_allclose(m.cpu().numpy(), g["m_after"][s], rtol=RTOL, atol=ATOL_G * 0.1)
        big = np.abs(g["grad"][s]) > 1e-6 * np.abs(g["grad"][s]).max()
        np.testing.assert_allclose(p.cpu().numpy()[big], g["p_after"][s][big], rtol=RTOL, atol=2e-7)
    assert c.tc_status() == 0


@pytest.mark.parametrize("f,n,row0,mode", [(20, 128, 0, 0), (20, 100, 0, 1), (5, 257, 0, 2), (20, 1000, 128, 1),
                                           (29, 3000, 0, 1), (31, 2000, 0, 2), (8, 1500, 0, 2), (16, 700, 0, 1), (24, 300, 0, 0), (1, 500, 0, 0), (13, 999, 12, 2), (7, 1, 0, 0),
                                           (20, 56909, 0, 0), (5, 70001, 4, 2)])
def test_tc_grad_against_oracle_and_ffma_kernel(wx, f, n, row0, mode):
    c = ctx_for(wx, f)
    g = torch.Generator().manual_seed(f * 1000 + n)
    N = row0 + n
    x = torch.randn(N, f, generator=g); y = (torch.rand(N, generator=g) < 0.3).long()
    p = torch.randn(c.P, generator=g) * 0.2
    seed, step = 11, 5
    mask = mk = None
    if mode == wx.DROPOUT_MASK:
        keep = torch.rand(n, 64, generator=g) < 0.8
        mask = keep.float()
        mk = dev(np.packbits(keep.numpy(), axis=1, bitorder="little").view(np.int64).reshape(-1).copy())
    elif mode == wx.DROPOUT_PHILOX:
        mask = torch.from_numpy(philox_ref.dropout_keep_mask(n, seed, step).astype(np.float32))
    xd, yd, pd = x.cuda(), y.cuda(), p.cuda()
    out = c.tc_mlp_grad(pd, xd, yd, row0=row0, batch=n, dropout=mode, keep_mask=mk, seed=seed, step=step).cpu().numpy()
    assert c.tc_status() == 0
    loss, grad = train_ref.loss_and_grads(p, f, x[row0:row0 + n], y[row0:row0 + n], mask)
    gr = grad.numpy()
    np.testing.assert_allclose(out[-1], loss.item(), rtol=RTOL)
    one_row = 3.0 / n if n >= 16384 else 0.0          # see module docstring
    np.testing.assert_allclose(out[:-1], gr, rtol=RTOL, atol=ATOL_G + one_row)
    assert np.linalg.norm(out[:-1] - gr) <= 1e-5 * np.linalg.norm(gr)
    ffma = c.mlp_grad(pd, xd, yd, row0=row0, batch=n, dropout=mode, keep_mask=mk, seed=seed, step=step).cpu().numpy()
    np.testing.assert_allclose(out[:-1], ffma[:-1], rtol=2 * RTOL, atol=2 * ATOL_G + one_row)
    # deterministic: bit-identical on repeat
    again = c.tc_mlp_grad(pd, xd, yd, row0=row0, batch=n, dropout=mode, keep_mask=mk, seed=seed, step=step).cpu().numpy()
    np.testing.assert_array_equal(out, again)


@pytest.mark.parametrize("f,n,r0", [(20, 30000, 0), (5, 9001, 4), (20, 128 * 148 + 64, 128)])
def test_tc_raw_fused_step_matches_ffma_raw_step(wx, f, n, r0):
    """wx_tc_train_step_raw (standardise in registers on the tensor-core path) vs wx_train_step_raw."""
    t = synth.synth_table(5, n + r0, f, null_rate=0.01, const_col=(f - 1 if f > 5 else None))
    c = ctx_for(wx, f)
    cols = [dev(t["raw"][j].astype(np.float32)) for j in range(f)]
    stats = c.col_stats(cols)
    yy = dev(preprocess_ref.label_encode(t["labels"]).astype(np.int64))
    p0 = torch.randn(c.P, generator=torch.Generator().manual_seed(1)) * 0.2
    res = []
    for fn in (c.train_step_raw, c.tc_train_step_raw):
        p = p0.cuda(); m = torch.zeros_like(p); v = torch.zeros_like(p)
        st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
        fn(p, m, v, st, cols, stats, yy, loss, row0=r0, batch=n, impute=1, dropout=wx.DROPOUT_PHILOX, seed=9)
        torch.cuda.synchronize()
        res.append((loss.item(), m.cpu().numpy()))
    assert c.tc_status() == 0
    np.testing.assert_allclose(res[1][0], res[0][0], rtol=RTOL)
    np.testing.assert_allclose(res[1][1], res[0][1], rtol=2 * RTOL, atol=2e-8 + 0.3 / n)     # m = 0.1 * grad


def test_tc_rejects_what_it_cannot_take(wx):
    c = ctx_for(wx, 5)
    p = torch.zeros(c.P, device="cuda"); x = torch.zeros(64, 5, device="cuda"); y = torch.zeros(64, dtype=torch.int64, device="cuda")
    with pytest.raises(wx.WxError):
        c.tc_mlp_grad(p, x, y, row0=1, batch=8)          # 20-byte row offset: not 16-byte aligned
    c32 = ctx_for(wx, 32)
    with pytest.raises(wx.WxError):
        c32.tc_mlp_grad(torch.zeros(c32.P, device="cuda"), torch.zeros(8, 32, device="cuda"), y[:8].contiguous())


def test_tc_selftest_gemms(wx):
    """The two tcgen05 GEMM shapes in isolation: 3-term tf32 split reaches fp32-level accuracy, 1 term tf32-level."""
    c = ctx_for(wx, 20)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(128, 32, generator=g); w = torch.randn(64, 32, generator=g) * 0.3; dh = torch.randn(128, 64, generator=g) * 0.01
    zref = (x.double() @ w.double().T).numpy(); gref = (dh.double().T @ x.double()).numpy()
    xd, wd, dd = x.cuda(), w.cuda(), dh.cuda()
    for terms, tol in ((1, 2e-3), (3, 2e-6)):
        z = torch.empty(128, 64, device="cuda"); gg = torch.empty(64, 32, device="cuda")
        st = torch.full((1,), -1, dtype=torch.int32, device="cuda")
        assert c.lib.wxdev_tc_selftest(c._h, xd.data_ptr(), wd.data_ptr(), dd.data_ptr(), terms, z.data_ptr(), gg.data_ptr(),
                                    st.data_ptr(), None) == 0
        torch.cuda.synchronize()
        assert st.item() == 0
        assert np.abs(z.cpu().numpy() - zref).max() <= tol * np.abs(zref).max()
        assert np.abs(gg.cpu().numpy() - gref).max() <= tol * np.abs(gref).max()


# ---------------------------------------------------------------------------------------------
# The configuration bench.py times (BASELINE.json configs[1]): F=20, 1 Mi rows per step, raw fp32 columns with nulls
# and a constant column, Philox dropout, mean imputation -- ~55 tile-slots per CTA, so the landing ring, every mbarrier
# phase bit and the TMEM-resident accumulators wrap many times.
# ---------------------------------------------------------------------------------------------
_BENCH = {}


def bench_table(n_total=(1 << 20) + 4096, f=20):
    """Raw table + oracle standardisation (oracle statistics), shared by the tests below."""
    if "t" not in _BENCH:
        rng = np.random.default_rng(77)
        raw = np.empty((f, n_total), dtype=np.float32)
        for j in range(f):
            raw[j] = np.round(rng.normal(synth._MU[j % 5] + 3.0 * (j // 5), synth._SIGMA[j % 5] * (1 + 0.25 * (j // 5)),
                                         size=n_total), 2).astype(np.float32)
        raw[f - 1] = 42.5                                              # constant column: std guard (preprocess.py:36)
        nulls = rng.random((f, n_total)) < 0.01
        nulls[f - 1] = False
        raw[nulls] = np.nan
        z0 = (np.nan_to_num(raw[0], nan=20.0) - 20.0) / 8.0
        z1 = (np.nan_to_num(raw[1], nan=65.0) - 65.0) / 20.0
        y = ((-1.6 + 0.9 * z1 - 0.7 * z0 + rng.normal(0, 0.6, size=n_total)) > 0).astype(np.int64)
        raw64 = raw.astype(np.float64)
        st = [preprocess_ref.column_stats(raw64[j]) for j in range(f)]
        stats = np.array([[c, m, (s or 0.0) ** 2 * (c - 1), preprocess_ref.std_guard(s)] for c, m, s in st])
        feats = np.empty((n_total, f), dtype=np.float32)
        for j in range(f):
            feats[:, j] = preprocess_ref.zscore(raw64[j], stats[j, 1], stats[j, 3]).astype(np.float32)
        feats[np.isnan(feats)] = 0.0                                   # WX_IMPUTE_MEAN
        _BENCH["t"] = dict(raw=raw, y=y, stats=stats, feats=feats)
    return _BENCH["t"]


def _oracle_grads(p, feats, y, n, seed, step):
    mask = torch.from_numpy(philox_ref.dropout_keep_mask(n, seed, step).astype(np.float32))
    loss, grad = train_ref.loss_and_grads(p, feats.shape[1], torch.from_numpy(feats), torch.from_numpy(y), mask)
    return float(loss), grad.numpy()


@pytest.mark.parametrize("n,row0", [(1 << 20, 0), ((1 << 20) - 77, 2052)])
def test_tc_grad_at_bench_configuration_matches_oracle(wx, n, row0):
    """wx_tc_mlp_grad on 1 Mi rows (8192 tiles over 148 CTAs), full and with a ragged last tile at row0 != 0."""
    t = bench_table()
    f = 20
    c = ctx_for(wx, f)
    p = torch.randn(c.P, generator=torch.Generator().manual_seed(4)) * 0.2
    seed, step = 42, 3
    loss_ref, g_ref = _oracle_grads(p, t["feats"][row0:row0 + n], t["y"][row0:row0 + n], n, seed, step)
    x = dev(t["feats"]); y = dev(t["y"])
    out = c.tc_mlp_grad(p.cuda(), x, y, row0=row0, batch=n, dropout=wx.DROPOUT_PHILOX, seed=seed, step=step).cpu().numpy()
    assert c.tc_status() == 0
    np.testing.assert_allclose(out[-1], loss_ref, rtol=RTOL)
    assert np.linalg.norm(out[:-1] - g_ref) <= 1e-5 * np.linalg.norm(g_ref)
    np.testing.assert_allclose(out[:-1], g_ref, rtol=RTOL, atol=ATOL_G + 3.0 / n)      # + one row's contribution


@pytest.mark.parametrize("n,row0", [(1 << 20, 0), ((1 << 20) - 77, 2052)])
def test_tc_raw_step_at_bench_configuration_matches_oracle_and_ffma(wx, n, row0):
    """wx_tc_train_step_raw exactly as bench.py launches it, against (i) the oracle: Adam's first step from zero state
    gives m = 0.1 * grad, v = 0.001 * grad^2 and the parameter update, and (ii) wx_train_step_raw (FFMA kernel)."""
    t = bench_table()
    f = 20
    c = ctx_for(wx, f)
    p0 = torch.randn(c.P, generator=torch.Generator().manual_seed(4)) * 0.2
    seed = 42
    loss_ref, g_ref = _oracle_grads(p0, t["feats"][row0:row0 + n], t["y"][row0:row0 + n], n, seed, 0)
    p_ref, m_ref, v_ref = p0.clone(), torch.zeros(c.P), torch.zeros(c.P)
    train_ref.adam_update(p_ref, torch.from_numpy(g_ref), m_ref, v_ref, 1)
    cols = [dev(t["raw"][j]) for j in range(f)]
    y = dev(t["y"]); stats = dev(t["stats"])
    res = {}
    for name, fn in (("tc", c.tc_train_step_raw), ("ffma", c.train_step_raw)):
        p = p0.cuda(); m = torch.zeros_like(p); v = torch.zeros_like(p)
        st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
        fn(p, m, v, st, cols, stats, y, loss, row0=row0, batch=n, impute=wx.IMPUTE_MEAN, dropout=wx.DROPOUT_PHILOX,
           seed=seed)
        torch.cuda.synchronize()
        assert st.cpu().tolist() == [1, 0]
        res[name] = (loss.item(), m.cpu().numpy(), p.cpu().numpy())
    assert c.tc_status() == 0
    for name in ("tc", "ffma"):
        loss, m, p = res[name]
        np.testing.assert_allclose(loss, loss_ref, rtol=RTOL, err_msg=name)
        assert np.linalg.norm(m - m_ref.numpy()) <= 1e-5 * np.linalg.norm(m_ref.numpy()), name
        np.testing.assert_allclose(m, m_ref.numpy(), rtol=RTOL, atol=0.1 * (ATOL_G + 3.0 / n), err_msg=name)
        big = np.abs(g_ref) > 1e-4 * np.abs(g_ref).max()              # Adam's first step is lr*sign(g): skip g ~ 0
        np.testing.assert_allclose(p[big], p_ref.numpy()[big], rtol=RTOL, atol=2e-7, err_msg=name)
    np.testing.assert_allclose(res["tc"][1], res["ffma"][1], rtol=2 * RTOL, atol=0.1 * (2 * ATOL_G + 3.0 / n))


def test_tc_raw_steps_track_the_oracle_over_several_steps(wx):
    """Five consecutive fused steps over different 256 Ki-row windows (device step counter drives Philox and Adam's
    bias correction): loss of every step and the final parameters against the oracle run on the same windows."""
    t = bench_table()
    f, n = 20, 1 << 18
    c = ctx_for(wx, f)
    p0 = torch.randn(c.P, generator=torch.Generator().manual_seed(6)) * 0.2
    cols = [dev(t["raw"][j]) for j in range(f)]
    y = dev(t["y"]); stats = dev(t["stats"])
    p = p0.cuda(); m = torch.zeros_like(p); v = torch.zeros_like(p)
    st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
    p_ref, m_ref, v_ref = p0.clone(), torch.zeros(c.P), torch.zeros(c.P)
    for s in range(5):
        row0 = (s % 4) * n
        c.tc_train_step_raw(p, m, v, st, cols, stats, y, loss, row0=row0, batch=n, impute=wx.IMPUTE_MEAN,
                            dropout=wx.DROPOUT_PHILOX, seed=42)
        loss_ref, g_ref = _oracle_grads(p_ref, t["feats"][row0:row0 + n], t["y"][row0:row0 + n], n, 42, s)
        train_ref.adam_update(p_ref, torch.from_numpy(g_ref), m_ref, v_ref, s + 1)
        np.testing.assert_allclose(loss.item(), loss_ref, rtol=1e-4)          # free-running: tolerance of the loss curve
    assert c.tc_status() == 0 and st.cpu().tolist() == [5, 0]
    # Adam's early steps are ~lr*sign(g): an entry whose gradient is within rounding of zero may move by a different
    # +-0.01 (hard part 4), so the parameters are compared as a distribution, not entry by entry
    d = np.abs(p.cpu().numpy() - p_ref.numpy())
    assert np.quantile(d, 0.99) <= 1e-3 and d.max() <= 0.11


def test_sgd_update_rule_in_the_fused_steps(wx):
    """WX_OPT_SGD (no counterpart in the reference, which only uses Adam, :88): p -= lr * g in the tails of both fused
    step kernels and in wx_adam_step, against the oracle's gradient and torch's SGD arithmetic."""
    f, n = 20, 8192
    g = torch.Generator().manual_seed(12)
    x = torch.randn(n, f, generator=g); y = (torch.rand(n, generator=g) < 0.3).long()
    c = wx.Context(f, device=0, optimizer="sgd", lr=0.05)
    p0 = torch.randn(c.P, generator=g) * 0.2
    _, grad = train_ref.loss_and_grads(p0, f, x, y, None)
    want = p0.clone()
    train_ref.sgd_update(want, grad, lr=0.05)
    xd, yd = x.cuda(), y.cuda()
    for name, fn in (("tc", c.tc_train_step), ("ffma", c.train_step)):
        p = p0.cuda(); m = torch.full_like(p, 7.0); v = torch.full_like(p, 9.0)
        st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
        fn(p, m, v, st, xd, yd, loss, dropout=wx.DROPOUT_OFF)
        torch.cuda.synchronize()
        np.testing.assert_allclose(p.cpu().numpy(), want.numpy(), rtol=1e-6, atol=1e-7, err_msg=name)
        assert st.cpu().tolist() == [1, 0] and bool((m == 7.0).all()) and bool((v == 9.0).all()), name
    p = p0.cuda()
    c.adam_step(p, grad.cuda(), torch.zeros_like(p), torch.zeros_like(p), 1, n=c.P)
    np.testing.assert_allclose(p.cpu().numpy(), want.numpy(), rtol=1e-6, atol=1e-8)
    assert c.tc_status() == 0
    c.close()
