"""GPU parity tests (run on the B200 box: pytest -m gpu).  Everything goes through the C ABI of
libwxddp.so via wxddp.Context; the oracle (oracle/) and the golden vectors produced by the reference's own
classes (tests/golden/) are the checkers.

Tolerances: bit-exact for labels, null masks, imputation indices, standardised values given
the same statistics, and the Philox mask stream; fp32 step arithmetic: rtol 1e-5 (+ atol 1e-7 on gradients,
summation order differs from torch's); statistics: rtol 1e-12 in float64.
"""
import os

import numpy as np
import pytest
import torch

from oracle import philox_ref, preprocess_ref, synth, train_ref

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RTOL, ATOL_G = 1e-5, 1e-7


@pytest.fixture(scope="module")
def wx():
    import wxddp
    return wxddp


_CTX = {}


def ctx_for(wx, f):
    if f not in _CTX:
        _CTX[f] = wx.Context(f, device=0)
    return _CTX[f]


def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def mask_dev(bits_u64):
    return dev(np.ascontiguousarray(bits_u64).view(np.int64))


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def unpack_mask(bits):
    return ((bits[:, None] >> np.arange(64, dtype=np.uint64)) & np.uint64(1)).astype(np.float32)


# ---------------------------------------------------------------------------------------------
# library identity
# ---------------------------------------------------------------------------------------------
def test_library_identity_and_arch(wx):
    lib = wx.load()
    assert lib.wx_build_arch() == b"sm_100a"
    assert torch.cuda.get_device_capability(0)[0] == 10
    c = ctx_for(wx, 5)
    assert c.P == 514 and c.num_sms > 0
    with pytest.raises(wx.WxError):
        wx.Context(33)
    with pytest.raises(wx.WxError):
        wx.Context(0)


# ---------------------------------------------------------------------------------------------
# K1
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [0, 1, 7, 1000, 100_003])
def test_label_encode_bit_exact(wx, n):
    rng = np.random.default_rng(n)
    pool = np.array(["rain", "no rain", None, "Rain", "RAIN", "rain ", "", " rain", "rainy", "rai", "niar"], dtype=object)
    strings = list(pool[rng.integers(0, len(pool), size=n)])
    off, data, valid = synth.to_arrow_strings(strings)
    c = ctx_for(wx, 5)
    want = preprocess_ref.label_encode(strings)
    got64, got32 = c.label_encode(dev(off), dev(data), dev(valid) if n else None, want_int32=True)
    assert got64.dtype == torch.int64
    np.testing.assert_array_equal(got64.cpu().numpy(), want.astype(np.int64))
    np.testing.assert_array_equal(got32.cpu().numpy(), want)
    # a null slot whose bytes happen to spell "rain" must still encode 0 (validity wins)
    if n >= 7:
        off2 = np.arange(0, 4 * n + 1, 4, dtype=np.int32)
        data2 = np.frombuffer(b"rain" * n, dtype=np.uint8).copy()
        v = rng.random(n) < 0.5
        got = c.label_encode(dev(off2), dev(data2), dev(np.packbits(v, bitorder="little")))
        np.testing.assert_array_equal(got.cpu().numpy(), v.astype(np.int64))


@pytest.mark.parametrize("f,n,dtype", [(5, 145_460, np.float64), (5, 145_460, np.float32), (20, 50_001, np.float64),
                                       (20, 50_001, np.float32), (1, 3, np.float64), (32, 2049, np.float32),
                                       (7, 1, np.float64)])
def test_col_stats_against_oracle(wx, f, n, dtype):
    t = synth.synth_table(11 + f, n, f, null_rate=0.01 if n > 10 else 0.0, const_col=(f - 1 if f >= 7 else None))
    raw = t["raw"].astype(dtype)
    c = ctx_for(wx, f)
    cols = [dev(raw[j]) for j in range(f)]
    got = c.col_stats(cols).cpu().numpy()
    for j in range(f):
        cnt, mean, std = preprocess_ref.column_stats(raw[j].astype(np.float64))
        assert got[j, 0] == cnt
        if cnt > 0:
            np.testing.assert_allclose(got[j, 1], mean, rtol=1e-12, atol=1e-12)
        sv = preprocess_ref.std_guard(std)
        if sv is None:
            assert np.isnan(got[j, 3])
        else:
            np.testing.assert_allclose(got[j, 3], sv, rtol=1e-11)
    if f >= 7 and n > 1:
        assert got[f - 1, 3] == 1.0        # constant column -> std 0 -> guard -> 1.0 (preprocess.py:36)


def test_col_stats_validity_bitmap_and_merge(wx):
    f, n = 5, 10_007
    t = synth.synth_table(3, n, f)
    raw = t["raw"]
    rng = np.random.default_rng(0)
    valid = rng.random((f, n)) > 0.1
    c = ctx_for(wx, f)
    cols = [dev(raw[j]) for j in range(f)]
    vbits = [dev(np.packbits(valid[j], bitorder="little")) for j in range(f)]
    vbits[2] = None
    got = c.col_stats(cols, validity=vbits).cpu().numpy()
    for j in range(f):
        col = raw[j].copy()
        if j != 2:
            col[~valid[j]] = np.nan
        cnt, mean, std = preprocess_ref.column_stats(col)
        assert got[j, 0] == cnt
        np.testing.assert_allclose([got[j, 1], got[j, 3]], [mean, std], rtol=1e-12)
    # Chan merge of shard statistics == statistics of the whole (multi-GPU rule)
    bounds = [0, 1234, 1235, 7000, n]
    parts = torch.stack([c.col_stats([col[a:b].contiguous() for col in cols]) for a, b in zip(bounds[:-1], bounds[1:])])
    merged = c.stats_merge(parts.contiguous()).cpu().numpy()
    whole = c.col_stats(cols).cpu().numpy()
    np.testing.assert_allclose(merged, whole, rtol=1e-12)


@pytest.mark.parametrize("f,n,dtype", [(5, 145_460, np.float64), (5, 1, np.float32), (20, 100_003, np.float32),
                                       (20, 255, np.float64), (1, 257, np.float64), (32, 5000, np.float32),
                                       (13, 4096, np.float64)])
@pytest.mark.parametrize("impute", [0, 1])
def test_normalize_pack_bit_exact_given_stats(wx, f, n, dtype, impute):
    t = synth.synth_table(100 + f, n, f, null_rate=0.02, const_col=(f - 1 if f >= 13 else None))
    raw = t["raw"].astype(dtype)
    raw64 = raw.astype(np.float64)
    # statistics from the oracle, injected on both sides -> element-wise path must agree bit for bit
    ostats = [preprocess_ref.column_stats(raw64[j]) for j in range(f)]
    inj = [(m, preprocess_ref.std_guard(s)) for (_, m, s) in ostats]
    want = preprocess_ref.preprocess_table(raw64, t["labels"], impute=impute, stats=inj)
    st = np.zeros((f, 4))
    for j, (cnt, m, s) in enumerate(ostats):
        sv = preprocess_ref.std_guard(s)
        st[j] = [cnt, m, 0.0, np.nan if sv is None else sv]
    c = ctx_for(wx, f)
    cols = [dev(raw[j]) for j in range(f)]
    feats, nm = c.normalize_pack(cols, dev(st), impute=impute)
    got = feats.cpu().numpy()
    assert got.shape == (n, f) and got.dtype == np.float32
    np.testing.assert_array_equal(np.isnan(got), np.isnan(want["features"]))
    ok = ~np.isnan(want["features"])
    if impute and n > 1:
        assert ok.all()
    np.testing.assert_array_equal(got[ok].view(np.uint32), want["features"][ok].view(np.uint32))
    np.testing.assert_array_equal(nm.cpu().numpy().view(np.uint32), want["nullmask"])
    # imputation indices (extension; pinned by the oracle only)
    idx, cnt = c.impute_index_list(nm)
    ref_idx = np.flatnonzero(np.isnan(raw64).T.reshape(-1))
    assert cnt == ref_idx.size
    np.testing.assert_array_equal(idx.cpu().numpy(), ref_idx)
    if cnt > 3:
        idx2, cnt2 = c.impute_index_list(nm, capacity=3)
        assert cnt2 == cnt and idx2.cpu().numpy().tolist() == ref_idx[:3].tolist()


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_normalize_pack_exact_rounding_at_scale(wx, dtype):
    """The kernel multiplies by 1/std and only divides when the fp32 rounding could differ: over 25M cells (a few of
    which sit next to a rounding midpoint) the result must still be float32((x - mean) / std) evaluated in IEEE float64."""
    f, n = 8, 3_145_728 + 17
    gen = torch.Generator(device="cuda").manual_seed(5)
    raw = (torch.randn(f, n, device="cuda", generator=gen, dtype=torch.float64) * 7.3 + 19.1).to(dtype)
    raw[:, ::1013] = float("nan")
    raw[3, 5:2000:7] = raw[3, 4]                      # repeated values
    c = ctx_for(wx, f)
    cols = [raw[j] for j in range(f)]
    stats = c.col_stats(cols)
    stats[3, 1] = raw[3, 4].double()                  # mean equal to a value that occurs: exact zeros in the output
    feats, nm = c.normalize_pack(cols, stats, impute=0)
    want = ((raw.double() - stats[:, 1:2]) / stats[:, 3:4]).float().t().contiguous()
    assert torch.equal(torch.isnan(feats), torch.isnan(want))
    assert torch.equal(torch.nan_to_num(feats).view(torch.int32), torch.nan_to_num(want).view(torch.int32))
    assert (feats[5:2000:7, 3] == 0).all()
    # same table through columns that are only element-aligned (direct-load kernels instead of the TMA ring / 16-byte loads)
    pad = torch.empty(f, n + 1, device="cuda", dtype=dtype)
    pad[:, 1:] = raw
    cols_u = [pad[j, 1:] for j in range(f)]
    assert cols_u[0].data_ptr() % 16 != 0
    feats_u, nm_u = c.normalize_pack(cols_u, stats, impute=0)
    assert torch.equal(torch.nan_to_num(feats_u).view(torch.int32), torch.nan_to_num(feats).view(torch.int32))
    assert torch.equal(nm_u, nm)
    st_u = c.col_stats(cols_u).cpu().numpy()
    st_a = c.col_stats(cols).cpu().numpy()
    np.testing.assert_array_equal(st_u[:, 0], st_a[:, 0])
    np.testing.assert_allclose(st_u[:, 1], st_a[:, 1], rtol=1e-13)
    np.testing.assert_allclose(st_u[:, 3], st_a[:, 3], rtol=1e-12)


def test_normalize_pack_extreme_statistics(wx):
    """std / mean far outside the comfortable range: subnormal fp32 results, overflow to inf, a std whose reciprocal is
    not representable -- all must equal the float64 divide."""
    f, n = 4, 1024
    rng = np.random.default_rng(11)
    raw = np.stack([rng.normal(0, 1, n), rng.normal(0, 1e-30, n), rng.normal(0, 1e30, n), rng.normal(5, 1e-3, n)])
    st = np.zeros((f, 4))
    st[:, 0] = n
    st[:, 1] = [0.0, 0.0, 0.0, 5.0]
    st[:, 3] = [1e39, 1e-300, 1e-12, 3e300]          # tiny quotients, huge ones (inf in fp32), overflowing, underflowing
    c = ctx_for(wx, f)
    for dtype in (np.float64, np.float32):
        r = raw.astype(dtype)
        feats, _ = c.normalize_pack([dev(r[j]) for j in range(f)], dev(st), impute=0)
        with np.errstate(over="ignore", under="ignore"):
            want = ((r.astype(np.float64) - st[:, 1:2]) / st[:, 3:4]).astype(np.float32).T
        np.testing.assert_array_equal(feats.cpu().numpy().view(np.uint32), want.view(np.uint32))


def test_col_stats_trend_and_outliers(wx):
    """The pivoted-sum statistics keep full accuracy on a column with a trend, a far-off level and a single outlier."""
    n = 1_000_003
    rng = np.random.default_rng(3)
    cols_np = np.stack([np.arange(n, dtype=np.float64) * 1e-3 + rng.normal(0, 1e-2, n),          # trend
                        1e9 + rng.normal(0, 1.0, n),                                             # far-off level
                        np.where(np.arange(n) == 17, 1e12, rng.normal(0, 1.0, n)),               # one outlier first tile
                        np.concatenate([np.zeros(n // 2), 1e6 + rng.normal(0, 1, n - n // 2)])]) # level jump
    c = ctx_for(wx, 4)
    got = c.col_stats([dev(cols_np[j]) for j in range(4)]).cpu().numpy()
    for j in range(4):
        cnt, m, s = preprocess_ref.column_stats(cols_np[j])
        assert got[j, 0] == cnt
        np.testing.assert_allclose(got[j, 1], m, rtol=1e-13)
        np.testing.assert_allclose(got[j, 3], s, rtol=1e-11)


@pytest.mark.parametrize("f,n", [(20, 1000), (20, 256), (5, 777), (8, 513), (32, 300)])
def test_k1_kernels_stay_inside_their_outputs(wx, f, n):
    """Guard words around the outputs of wx_normalize_pack (TMA bulk store of the ragged last tile, vector stores) and
    wx_label_encode (4 rows per thread): nothing outside [0, n) may be touched."""
    c = ctx_for(wx, f)
    t = synth.synth_table(5, n, f, null_rate=0.01)
    cols = [dev(t["raw"][j].astype(np.float32)) for j in range(f)]
    stats = c.col_stats(cols)
    pad = 64
    buf = torch.full((n * f + 2 * pad,), -12345.0, device="cuda")
    nmb = torch.full((n + 2 * pad,), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    out = buf[pad:pad + n * f].view(n, f)
    from wxddp import _lib
    check = _lib.check
    check(c.lib.wx_normalize_pack(c._h, _lib.ptr_array([x.data_ptr() for x in cols]), 0, None, n, stats.data_ptr(), 1,
                                  out.data_ptr(), nmb[pad:].data_ptr(), None))
    torch.cuda.synchronize()
    assert (buf[:pad] == -12345.0).all() and (buf[pad + n * f:] == -12345.0).all()
    assert (nmb[:pad] == 0x5A5A5A5A).all() and (nmb[pad + n:] == 0x5A5A5A5A).all()
    ref, _ = c.normalize_pack(cols, stats, impute=1)
    assert torch.equal(out, ref)
    off, data, valid = synth.to_arrow_strings(t["labels"])
    offd, datad, validd = dev(off), dev(data), dev(valid)
    lab = torch.full((n + 2 * pad,), -7, dtype=torch.int64, device="cuda")
    check(c.lib.wx_label_encode(c._h, offd.data_ptr(), datad.data_ptr(), validd.data_ptr(), n, lab[pad:].data_ptr(), None,
                                None))
    torch.cuda.synchronize()
    assert (lab[:pad] == -7).all() and (lab[pad + n:] == -7).all()
    np.testing.assert_array_equal(lab[pad:pad + n].cpu().numpy(), preprocess_ref.label_encode(t["labels"]))


def test_normalize_with_gpu_stats_within_one_ulp(wx):
    f, n = 20, 145_460
    t = synth.synth_table(1234, n, f, null_rate=0.0, const_col=f - 1)
    c = ctx_for(wx, f)
    cols = [dev(t["raw"][j]) for j in range(f)]
    feats, _ = c.normalize_pack(cols, c.col_stats(cols))
    want = preprocess_ref.preprocess_table(t["raw"], t["labels"])["features"]
    got = feats.cpu().numpy()
    ulp = np.abs(got.view(np.int32).astype(np.int64) - want.view(np.int32).astype(np.int64))
    assert ulp.max() <= 1, ulp.max()
    assert (ulp == 0).mean() > 0.999


def test_dataset_golden_from_reference_weatherdataset(wx):
    """raw table -> GPU stats + pack == what the reference's WeatherDataset read back from Spark-shaped parquet."""
    g = load("dataset_F5.npz")
    c = ctx_for(wx, 5)
    cols = [dev(g["raw"][j]) for j in range(5)]
    feats, nm = c.normalize_pack(cols, c.col_stats(cols))
    got = feats.cpu().numpy()
    want = g["features"]
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = ~np.isnan(want)
    ulp = np.abs(got[ok].view(np.int32).astype(np.int64) - want[ok].view(np.int32).astype(np.int64))
    assert ulp.max() <= 1
    strings = [None if s == "\0" else str(s) for s in g["label_strings"]]
    off, data, valid = synth.to_arrow_strings(strings)
    lab = c.label_encode(dev(off), dev(data), dev(valid))
    np.testing.assert_array_equal(lab.cpu().numpy(), g["labels"])


# ---------------------------------------------------------------------------------------------
# K2 / K3: teacher-forced against vectors produced by the reference's WeatherClassifier
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("tag", ["F5_B4", "F20_B4", "F20_B256", "F5_B1000"])
def test_grad_and_adam_match_reference_module(wx, tag):
    g = load(f"steps_{tag}.npz")
    f = int(g["f"])
    c = ctx_for(wx, f)
    P = c.P
    steps = g["loss"].shape[0]
    m = torch.zeros(P, device="cuda")
    v = torch.zeros(P, device="cuda")
    for s in range(steps):
        p = dev(g["p_before"][s])
        x, y, mk = dev(g["x"][s]), dev(g["y"][s]), mask_dev(g["mask"][s])
        out = c.mlp_grad(p, x, y, dropout=wx.DROPOUT_MASK, keep_mask=mk).cpu().numpy()
        np.testing.assert_allclose(out[P], g["loss"][s], rtol=RTOL)
        np.testing.assert_allclose(out[:P], g["grad"][s], rtol=RTOL, atol=ATOL_G)
        # same rows through the sampler-index path
        n = x.shape[0]
        perm = torch.randperm(n, generator=torch.Generator().manual_seed(s))
        inv = torch.argsort(perm)
        out2 = c.mlp_grad(p, x[perm].contiguous(), y[perm].contiguous(), row_idx=inv.cuda(), dropout=wx.DROPOUT_MASK,
                          keep_mask=mk).cpu().numpy()
        np.testing.assert_array_equal(out2, out)
        # Adam on the reference's own gradient (isolates K3)
        gr = dev(g["grad"][s])
        c.adam_step(p, gr, m, v, s + 1)
        np.testing.assert_allclose(p.cpu().numpy(), g["p_after"][s], rtol=RTOL, atol=1e-9)
        np.testing.assert_allclose(m.cpu().numpy(), g["m_after"][s], rtol=RTOL, atol=1e-12)
        np.testing.assert_allclose(v.cpu().numpy(), g["v_after"][s], rtol=RTOL, atol=1e-15)
        m.copy_(dev(g["m_after"][s])); v.copy_(dev(g["v_after"][s]))


@pytest.mark.parametrize("tag", ["F5_B4", "F20_B256", "F5_B1000"])
def test_fused_step_matches_reference_module(wx, tag):
    """wx_train_step (one launch: fwd+loss+bwd+reduce+Adam), teacher-forced on params/m/v per step."""
    g = load(f"steps_{tag}.npz")
    f = int(g["f"])
    c = ctx_for(wx, f)
    P = c.P
    steps = g["loss"].shape[0]
    loss = torch.zeros(1, device="cuda")
    for s in range(steps):
        p = dev(g["p_before"][s])
        m = dev(g["m_after"][s - 1]) if s else torch.zeros(P, device="cuda")
        v = dev(g["v_after"][s - 1]) if s else torch.zeros(P, device="cuda")
        state = torch.tensor([s, 0], dtype=torch.int64, device="cuda")
        c.train_step(p, m, v, state, dev(g["x"][s]), dev(g["y"][s]), loss, dropout=wx.DROPOUT_MASK,
                     keep_mask=mask_dev(g["mask"][s]))
        assert state.cpu().tolist() == [s + 1, 0]
        np.testing.assert_allclose(loss.item(), g["loss"][s], rtol=RTOL)
        np.testing.assert_allclose(m.cpu().numpy(), g["m_after"][s], rtol=RTOL, atol=ATOL_G * 0.1)
        # Adam's first steps move every weight by ~lr*sign(g): entries whose gradient is at rounding-noise level
        # (|g| < 1e-6 of the largest) are compared on the gradient/moment only, not on the +-lr jump
        big = np.abs(g["grad"][s]) > 1e-6 * np.abs(g["grad"][s]).max()
        np.testing.assert_allclose(p.cpu().numpy()[big], g["p_after"][s][big], rtol=RTOL, atol=2e-7)
        assert big.mean() > 0.5


def test_small_steps_kernel_free_running_loss_curve(wx):
    """wx_train_steps_small: 24 reference-sized steps (batch 4) in ONE launch vs the reference's own run."""
    g = load("steps_F5_B4.npz")
    c = ctx_for(wx, 5)
    steps, b = g["loss"].shape[0], int(g["batch"])
    x = dev(g["x"].reshape(steps * b, 5)); y = dev(g["y"].reshape(-1)); mk = mask_dev(g["mask"].reshape(-1))
    p = dev(g["p_before"][0]); m = torch.zeros(c.P, device="cuda"); v = torch.zeros(c.P, device="cuda")
    state = torch.zeros(2, dtype=torch.int64, device="cuda")
    losses = torch.zeros(steps, device="cuda")
    c.train_steps_small(p, m, v, state, x, y, losses, n_steps=steps, batch=b, dropout=wx.DROPOUT_MASK, keep_mask=mk)
    assert state[0].item() == steps
    # free running: round-off compounds through Adam, so the curve is compared at 1e-4
    np.testing.assert_allclose(losses.cpu().numpy(), g["loss"], rtol=1e-4)
    np.testing.assert_allclose(p.cpu().numpy(), g["p_after"][-1], rtol=1e-3, atol=1e-5)
    # and the same 24 steps as 24 launches of the grid kernel give the same curve
    p2 = dev(g["p_before"][0]); m2 = torch.zeros_like(m); v2 = torch.zeros_like(v)
    st2 = torch.zeros(2, dtype=torch.int64, device="cuda"); l2 = torch.zeros(1, device="cuda")
    out = []
    for s in range(steps):
        c.train_step(p2, m2, v2, st2, x, y, l2, row0=s * b, batch=b, dropout=wx.DROPOUT_MASK,
                     keep_mask=mk[s * b:(s + 1) * b].contiguous())
        out.append(l2.item())
    np.testing.assert_allclose(out, losses.cpu().numpy(), rtol=1e-5)


@pytest.mark.parametrize("f,batch,mode", [(5, 256, "mask"), (20, 1000, "philox"), (20, 4096, "philox"), (7, 72, "off"),
                                          (20, 16384, "philox")])
def test_resident_multi_cta_steps_equal_step_by_step_launches(wx, f, batch, mode):
    """wx_train_steps_small with 64 < batch: several co-resident CTAs, one grid barrier per step.  Must reproduce the
    same steps launched one by one through wx_train_step (same rows, same dropout bits), ragged last chunk included."""
    torch.manual_seed(f + batch)
    c = ctx_for(wx, f)
    steps = 6
    n = steps * batch + 5
    x = torch.randn(n, f, device="cuda")
    y = (torch.rand(n, device="cuda") < 0.3).long()
    p0 = (torch.randn(c.P, device="cuda") * 0.2)
    dm = {"mask": wx.DROPOUT_MASK, "philox": wx.DROPOUT_PHILOX, "off": wx.DROPOUT_OFF}[mode]
    mk = None
    if mode == "mask":
        bits = (torch.rand(steps * batch, 64) < 0.8)
        mk = mask_dev(np.packbits(bits.numpy(), axis=1, bitorder="little").view(np.uint64).reshape(-1))
    idx = torch.randperm(n, device="cuda")[:steps * batch].contiguous() if f == 5 else None
    p, m, v = p0.clone(), torch.zeros(c.P, device="cuda"), torch.zeros(c.P, device="cuda")
    st = torch.zeros(2, dtype=torch.int64, device="cuda")
    losses = torch.zeros(steps, device="cuda")
    c.train_steps_small(p, m, v, st, x, y, losses, n_steps=steps, batch=batch, row_idx=idx, dropout=dm, keep_mask=mk, seed=9)
    assert st[0].item() == steps and c.tc_status() == 0
    p2, m2, v2 = p0.clone(), torch.zeros_like(m), torch.zeros_like(v)
    st2 = torch.zeros(2, dtype=torch.int64, device="cuda"); l2 = torch.zeros(1, device="cuda")
    want = []
    for s in range(steps):
        c.train_step(p2, m2, v2, st2, x, y, l2, row0=s * batch, batch=batch, dropout=dm, seed=9,
                     row_idx=None if idx is None else idx[s * batch:(s + 1) * batch].contiguous(),
                     keep_mask=None if mk is None else mk[s * batch:(s + 1) * batch].contiguous())
        want.append(l2.item())
    np.testing.assert_allclose(losses.cpu().numpy(), want, rtol=2e-5)
    np.testing.assert_allclose(m.cpu().numpy(), m2.cpu().numpy(), rtol=2e-4, atol=1e-7)
    # Adam moves a weight whose gradient is rounding noise by +-lr either way; compare where the two runs agree on m
    np.testing.assert_allclose(p.cpu().numpy(), p2.cpu().numpy(), rtol=1e-3, atol=2e-4)


@pytest.mark.parametrize("tag,f", [("F5", 5), ("F20", 20)])
def test_eval_and_predict_match_reference_validation_step(wx, tag, f):
    g = load(f"eval_{tag}.npz")
    c = ctx_for(wx, f)
    p, x, y = dev(g["params"]), dev(g["x"]), dev(g["y"])
    for lo, hi, loss, acc in zip(g["lo"], g["hi"], g["loss"], g["acc"]):
        out = c.mlp_eval(p, x, y, row0=int(lo), batch=int(hi - lo)).cpu().numpy()
        np.testing.assert_allclose(out[0] / (hi - lo), loss, rtol=RTOL)
        assert np.float32(out[1] / (hi - lo)) == acc
    idx = torch.arange(1023, -1, -1, device="cuda")
    a = c.mlp_eval(p, x, y).cpu().numpy(); b = c.mlp_eval(p, x, y, row_idx=idx).cpu().numpy()
    np.testing.assert_allclose(a, b, rtol=1e-12)
    probs = c.mlp_predict(p, x).cpu().numpy()
    want = torch.softmax(train_ref.forward_with_mask(torch.from_numpy(g["params"]), f, torch.from_numpy(g["x"]), None), 1)
    np.testing.assert_allclose(probs, want.numpy(), rtol=RTOL, atol=1e-7)
    # forward() itself: raw logits (train_lightning_ddp.py:64), also where a probability would underflow
    lg_ref = train_ref.forward_with_mask(torch.from_numpy(g["params"]), f, torch.from_numpy(g["x"]), None).numpy()
    np.testing.assert_allclose(c.mlp_logits(p, x).cpu().numpy(), lg_ref, rtol=RTOL, atol=1e-6)
    from wxddp.module import WeatherClassifier
    model = WeatherClassifier(f)
    model.load_flat(torch.from_numpy(g["params"]))
    model = model.cuda().eval()
    with torch.no_grad():
        out = model(x)
        big = model(x * 1e4)                      # logit gaps far beyond 103: log(softmax) would give -inf here
    np.testing.assert_allclose(out.cpu().numpy(), lg_ref, rtol=RTOL, atol=1e-6)
    assert torch.isfinite(big).all()


# ---------------------------------------------------------------------------------------------
# dropout stream, edge cases, size-independent properties
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("f,b", [(5, 37), (20, 1000)])
def test_philox_dropout_stream_is_the_documented_one(wx, f, b):
    c = ctx_for(wx, f)
    rng = np.random.default_rng(f)
    x = rng.normal(size=(b, f)).astype(np.float32); y = (rng.random(b) < 0.3).astype(np.int64)
    p = (rng.normal(size=c.P) * 0.3).astype(np.float32)
    seed, step = 0x1234_5678_9ABC_DEF0, 41
    keep = philox_ref.dropout_keep_mask(b, seed, step)
    assert 0.75 < keep.mean() < 0.85
    bits = np.packbits(keep, axis=1, bitorder="little").view(np.uint64).reshape(-1)
    a = c.mlp_grad(dev(p), dev(x), dev(y), dropout=wx.DROPOUT_PHILOX, seed=seed, step=step).cpu().numpy()
    m = c.mlp_grad(dev(p), dev(x), dev(y), dropout=wx.DROPOUT_MASK, keep_mask=mask_dev(bits)).cpu().numpy()
    np.testing.assert_array_equal(a, m)            # same mask -> same arithmetic, bit for bit
    other = c.mlp_grad(dev(p), dev(x), dev(y), dropout=wx.DROPOUT_PHILOX, seed=seed, step=step + 1).cpu().numpy()
    assert not np.array_equal(a, other)


@pytest.mark.parametrize("f", [1, 2, 3, 4, 7, 8, 16, 31, 32])
@pytest.mark.parametrize("b", [1, 9])
def test_every_feature_count_against_oracle(wx, f, b):
    c = ctx_for(wx, f)
    rng = np.random.default_rng(100 * f + b)
    x = rng.normal(size=(b, f)).astype(np.float32); y = (rng.random(b) < 0.4).astype(np.int64)
    p = (rng.normal(size=c.P) * 0.2).astype(np.float32)
    keep = rng.random((b, 64)) < 0.8
    bits = np.packbits(keep, axis=1, bitorder="little").view(np.uint64).reshape(-1)
    loss, grad = train_ref.loss_and_grads(torch.from_numpy(p), f, torch.from_numpy(x), torch.from_numpy(y),
                                          torch.from_numpy(keep.astype(np.float32)))
    out = c.mlp_grad(dev(p), dev(x), dev(y), dropout=wx.DROPOUT_MASK, keep_mask=mask_dev(bits)).cpu().numpy()
    np.testing.assert_allclose(out[-1], loss.item(), rtol=RTOL)
    np.testing.assert_allclose(out[:-1], grad.numpy(), rtol=RTOL, atol=ATOL_G)


def test_nan_feature_makes_loss_nan_like_the_reference(wx):
    """The reference never imputes: a null cell reaches the trainer as NaN and the batch loss is NaN."""
    c = ctx_for(wx, 5)
    x = np.zeros((4, 5), np.float32); x[2, 3] = np.nan
    y = np.array([0, 1, 0, 1]); p = np.full(c.P, 0.1, np.float32)
    out = c.mlp_grad(dev(p), dev(x), dev(y)).cpu().numpy()
    assert np.isnan(out[-1])
    ref_loss, _ = train_ref.loss_and_grads(torch.from_numpy(p), 5, torch.from_numpy(x), torch.from_numpy(y), None)
    assert np.isnan(ref_loss.item())


def test_large_batch_properties_and_oracle(wx):
    """1 Mi-row step: determinism, additivity over a row split, and agreement with the CPU oracle."""
    f, n = 20, 1 << 20
    c = ctx_for(wx, f)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(n, f, generator=g); y = (torch.rand(n, generator=g) < 0.2).long()
    p = torch.randn(c.P, generator=g) * 0.2
    xd, yd, pd = x.cuda(), y.cuda(), p.cuda()
    a = c.mlp_grad(pd, xd, yd).cpu().numpy()
    b = c.mlp_grad(pd, xd, yd).cpu().numpy()
    np.testing.assert_array_equal(a, b)                                        # deterministic reduction order
    k = 300_001
    s = 1.0 / n
    lo = c.mlp_grad(pd, xd, yd, row0=0, batch=k, scale=s).cpu().numpy().astype(np.float64)
    hi = c.mlp_grad(pd, xd, yd, row0=k, batch=n - k, scale=s).cpu().numpy().astype(np.float64)
    np.testing.assert_allclose(lo + hi, a, rtol=2e-5, atol=1e-8)               # additivity over rows
    loss, grad = train_ref.loss_and_grads(p, f, x, y, None)
    np.testing.assert_allclose(a[-1], loss.item(), rtol=RTOL)
    np.testing.assert_allclose(a[:-1], grad.numpy(), rtol=1e-4, atol=1e-7)      # torch's own 1M-term fp32 sums are the looser side


def test_raw_fused_step_equals_pack_then_step(wx):
    """wx_train_step_raw (standardise in registers) == wx_normalize_pack followed by wx_train_step."""
    for f, dtype in [(5, np.float64), (20, np.float32)]:
        n = 12_003
        t = synth.synth_table(77, n, f, null_rate=0.01, const_col=(f - 1 if f > 5 else None))
        raw = t["raw"].astype(dtype)
        c = ctx_for(wx, f)
        cols = [dev(raw[j]) for j in range(f)]
        stats = c.col_stats(cols)
        y = dev(preprocess_ref.label_encode(t["labels"]).astype(np.int64))
        feats, _ = c.normalize_pack(cols, stats, impute=1)
        rng = np.random.default_rng(1)
        p0 = (rng.normal(size=c.P) * 0.2).astype(np.float32)
        res = []
        for which in range(2):
            p = dev(p0); m = torch.zeros(c.P, device="cuda"); v = torch.zeros(c.P, device="cuda")
            st = torch.zeros(2, dtype=torch.int64, device="cuda"); loss = torch.zeros(1, device="cuda")
            for s, (r0, b) in enumerate([(0, 4096), (4096, 4099), (8195, n - 8195)]):
                if which == 0:
                    c.train_step(p, m, v, st, feats, y, loss, row0=r0, batch=b, dropout=wx.DROPOUT_PHILOX, seed=9)
                else:
                    c.train_step_raw(p, m, v, st, cols, stats, y, loss, row0=r0, batch=b, impute=1,
                                     dropout=wx.DROPOUT_PHILOX, seed=9)
            res.append((p.cpu().numpy(), loss.item()))
        assert np.isfinite(res[0][1])
        np.testing.assert_array_equal(res[0][0], res[1][0])
        assert res[0][1] == res[1][1]


def test_bad_arguments_are_rejected_with_messages(wx):
    c = ctx_for(wx, 5)
    p = torch.zeros(c.P, device="cuda"); x = torch.zeros(8, 5, device="cuda"); y = torch.zeros(8, dtype=torch.int64, device="cuda")
    with pytest.raises(wx.WxError):
        c.mlp_grad(p, x, y, row0=4, batch=8)
    with pytest.raises(wx.WxError):
        c.mlp_grad(p, x.double(), y)
    with pytest.raises(wx.WxError):
        c.mlp_grad(p, x, y, dropout=wx.DROPOUT_MASK)
    with pytest.raises(wx.WxError):
        c.train_steps_small(p, p.clone(), p.clone(), torch.zeros(2, dtype=torch.int64, device="cuda"), x, y,
                            torch.zeros(1, device="cuda"), n_steps=1, batch=65)
    lib = wx.load()
    assert lib.wx_mlp_grad(None, None, None, None, None, 0, 1, 0, None, 0, 0, 1.0, None, None) < 0
    assert b"ctx" in lib.wx_last_error()
