"""CPU tests: the oracle against the golden vectors produced by the reference's own classes
(tests/golden/make_golden.py) and against independent restatements."""
import os
import shutil
import tempfile

import numpy as np
import pytest
import torch

from oracle import preprocess_ref, synth, train_ref

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.mark.parametrize("f", [5, 20])
def test_seeded_setup_matches_reference_order(f):
    g = load("setup.npz")
    tr, va, model = train_ref.seeded_setup(1000, f)
    assert np.array_equal(np.array(tr), g[f"train_idx_F{f}"])
    assert np.array_equal(np.array(va), g[f"val_idx_F{f}"])
    assert np.array_equal(train_ref.flatten(model).numpy(), g[f"params_F{f}"])


@pytest.mark.parametrize("tag", ["F5_B4", "F20_B4", "F20_B256", "F5_B1000"])
def test_step_matches_reference_module(tag):
    g = load(f"steps_{tag}.npz")
    f = int(g["f"])
    steps = g["loss"].shape[0]
    p_dim = train_ref.n_params(f)
    m = torch.zeros(p_dim)
    v = torch.zeros(p_dim)
    for s in range(steps):
        p = torch.from_numpy(g["p_before"][s].copy())
        x = torch.from_numpy(g["x"][s]); y = torch.from_numpy(g["y"][s])
        bits = g["mask"][s]
        mask = torch.from_numpy(((bits[:, None] >> np.arange(64, dtype=np.uint64)) & np.uint64(1)).astype(np.float32))
        loss, grad = train_ref.loss_and_grads(p, f, x, y, mask)
        assert np.float32(loss.item()) == g["loss"][s]                       # bit-exact
        np.testing.assert_array_equal(grad.numpy(), g["grad"][s])            # bit-exact
        train_ref.adam_update(p, grad, m, v, s + 1)
        np.testing.assert_array_equal(m.numpy(), g["m_after"][s])
        np.testing.assert_array_equal(v.numpy(), g["v_after"][s])
        np.testing.assert_array_equal(p.numpy(), g["p_after"][s])


def test_backward_against_float64_formulas():
    """Independent restatement of SURVEY 8(a) a9-a11 in numpy float64 (no autograd)."""
    g = load("steps_F20_B256.npz")
    f = 20
    p = g["p_before"][3].astype(np.float64)
    x = g["x"][3].astype(np.float64); y = g["y"][3]
    bits = g["mask"][3]
    mask = ((bits[:, None] >> np.arange(64, dtype=np.uint64)) & np.uint64(1)).astype(np.float64)
    w1 = p[:64 * f].reshape(64, f); b1 = p[64 * f:64 * f + 64]
    w2 = p[64 * f + 64:64 * f + 192].reshape(2, 64); b2 = p[-2:]
    z = x @ w1.T + b1
    a = np.maximum(z, 0) * mask * 1.25
    lg = a @ w2.T + b2
    mx = lg.max(1, keepdims=True)
    lse = mx[:, 0] + np.log(np.exp(lg - mx).sum(1))
    loss = np.mean(lse - lg[np.arange(len(y)), y])
    sm = np.exp(lg - lse[:, None])
    dl = sm.copy(); dl[np.arange(len(y)), y] -= 1; dl /= len(y)
    dw2 = dl.T @ a; db2 = dl.sum(0)
    dh = (dl @ w2) * mask * 1.25 * (z > 0)
    dw1 = dh.T @ x; db1 = dh.sum(0)
    grad = np.concatenate([dw1.ravel(), db1, dw2.ravel(), db2])
    assert abs(loss - g["loss"][3]) < 2e-6
    np.testing.assert_allclose(g["grad"][3], grad, rtol=2e-4, atol=2e-7)


@pytest.mark.parametrize("tag,f", [("F5", 5), ("F20", 20)])
def test_eval_matches_reference_validation_step(tag, f):
    g = load(f"eval_{tag}.npz")
    p = torch.from_numpy(g["params"])
    x = torch.from_numpy(g["x"]); y = torch.from_numpy(g["y"])
    for lo, hi, loss, acc in zip(g["lo"], g["hi"], g["loss"], g["acc"]):
        l, c = train_ref.eval_batch(p, f, x[lo:hi], y[lo:hi])
        assert np.float32(l.item()) == loss
        assert np.float32(c / (hi - lo)) == acc


def test_dataset_matches_reference_weatherdataset():
    g = load("dataset_F5.npz")
    strings = [None if s == "\0" else str(s) for s in g["label_strings"]]
    pp = preprocess_ref.preprocess_table(g["raw"], strings)
    # the element-wise oracle reproduces what the reference class read back from parquet
    np.testing.assert_array_equal(pp["features"], g["features"])
    np.testing.assert_array_equal(pp["labels"], g["labels"])
    assert np.isnan(g["features"]).sum() == sum(bin(int(m)).count("1") for m in pp["nullmask"])
    # and load_processed restates WeatherDataset itself
    tmp = tempfile.mkdtemp()
    try:
        names = synth.feature_names(5)
        preprocess_ref.write_parquet_dir(tmp, names, g["norm"], g["labels32"], n_parts=3)
        feats, labels, cols = train_ref.load_processed(tmp)
        assert cols == [f"{n}_norm" for n in names]
        np.testing.assert_array_equal(feats.numpy(), g["features"])
        np.testing.assert_array_equal(labels.numpy(), g["labels"])
        assert feats.dtype == torch.float32 and labels.dtype == torch.int64
        with pytest.raises(FileNotFoundError):
            train_ref.load_processed(os.path.join(tmp, "missing"))
        # no *_norm columns -> ValueError (train_lightning_ddp.py:39-40)
        import pyarrow as pa, pyarrow.parquet as pq
        bad = os.path.join(tmp, "bad", "data.parquet"); os.makedirs(bad)
        pq.write_table(pa.table({"a": [1.0], "label_encoded": [0]}), os.path.join(bad, "part-0.parquet"))
        with pytest.raises(ValueError):
            train_ref.load_processed(os.path.join(tmp, "bad"))
    finally:
        shutil.rmtree(tmp)


def test_fit_world1_matches_dataloader_loop():
    g = load("loop_W1.npz")
    h = train_ref.fit(torch.from_numpy(g["features"]), torch.from_numpy(g["labels"]), max_epochs=2,
                      sanity_check=False)
    np.testing.assert_array_equal(np.array(h["train_loss"], dtype=np.float32), g["train_loss"])
    np.testing.assert_allclose(h["val_loss"], g["val_loss"], rtol=1e-6)
    np.testing.assert_allclose(h["val_acc"], g["val_acc"], rtol=1e-6)
    np.testing.assert_array_equal(h["params"].numpy(), g["params"])


def test_label_encode_and_stats_semantics():
    s = ["rain", "no rain", None, "Rain", "RAIN", "rain ", "", "rain"]
    assert preprocess_ref.label_encode(s).tolist() == [1, 0, 0, 0, 0, 0, 0, 1]
    col = np.array([1.0, np.nan, 3.0, 5.0])
    n, mean, std = preprocess_ref.column_stats(col)
    assert (n, mean) == (3, 3.0) and abs(std - 2.0) < 1e-15          # sample stddev, nulls skipped
    assert preprocess_ref.column_stats(np.array([np.nan, 2.0]))[2] is None
    assert preprocess_ref.std_guard(0.0) == 1.0 and preprocess_ref.std_guard(2.5) == 2.5
    z = preprocess_ref.zscore(col, mean, std)
    assert np.isnan(z[1]) and z[0] == -1.0 and z[3] == 1.0
    const = preprocess_ref.preprocess_table(np.full((1, 8), 42.5), ["rain"] * 8)
    assert np.all(const["features"] == 0.0) and const["stats"][0][2] == 1.0


def test_impute_mean_mask_and_indices():
    t = synth.synth_table(5, 257, 7, null_rate=0.05)
    a = preprocess_ref.preprocess_table(t["raw"], t["labels"], impute=preprocess_ref.IMPUTE_NONE)
    b = preprocess_ref.preprocess_table(t["raw"], t["labels"], impute=preprocess_ref.IMPUTE_MEAN)
    nan_cells = np.isnan(a["features"])
    assert nan_cells.sum() > 0 and not np.isnan(b["features"]).any()
    assert np.array_equal(np.flatnonzero(nan_cells.reshape(-1)), b["imputed_idx"])
    assert np.all(b["features"][nan_cells] == 0.0)
    bits = (a["nullmask"][:, None] >> np.arange(7, dtype=np.uint32)) & 1
    assert np.array_equal(bits.astype(bool), nan_cells)


def test_arrow_string_layout_roundtrip():
    s = ["rain", None, "", "no rain"]
    off, data, valid = synth.to_arrow_strings(s)
    assert off.tolist() == [0, 4, 4, 4, 11] and bytes(data) == b"rainno rain" and valid[0] == 0b1101


def test_sampler_restatement_matches_torch_distributed_sampler():
    from torch.utils.data import DistributedSampler
    ds = list(range(1003))
    for world in (2, 8):
        for rank in range(world):
            for epoch in (0, 3):
                smp = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True, seed=42)
                smp.set_epoch(epoch)
                assert list(iter(smp)) == train_ref.epoch_indices(1003, epoch, rank, world)
            smp = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=False)
            assert list(iter(smp)) == train_ref.val_indices_for_rank(1003, rank, world)


# ---------------------------------------------------------------------------------------------
# wide-MLP oracle (oracle/wide_ref.py): pinned through the narrow model and through autograd
# ---------------------------------------------------------------------------------------------
def test_wide_oracle_reduces_to_the_reference_model():
    """hidden=64, one hidden layer, no bf16 rounding: wide_ref IS the reference's net (train_lightning_ddp.py:57-69),
    so it must reproduce train_ref (itself pinned by the golden vectors of the reference module)."""
    import torch
    from oracle import train_ref, wide_ref

    f, b = 5, 96
    torch.manual_seed(3)
    flat = wide_ref.init_params(f, 64, 1)
    assert flat.numel() == train_ref.n_params(f) == wide_ref.n_params(f, 64, 1)
    x, y = torch.randn(b, f), (torch.rand(b) < 0.4).long()
    mask = (torch.rand(b, 64) < 0.8)
    loss, g = train_ref.loss_and_grads(flat, f, x, y, mask.float())
    got = wide_ref.loss_and_grads(flat, x, y, f, 64, 1, dropout=True, emulate_bf16=False, masks=[mask])
    np.testing.assert_allclose(float(got[-1]), float(loss), rtol=1e-6)
    np.testing.assert_allclose(got[:-1].numpy(), g.numpy(), rtol=2e-5, atol=1e-8)


def test_wide_oracle_manual_backward_matches_autograd():
    import torch
    from oracle import wide_ref

    f, h, nl, b = 20, 128, 3, 64
    torch.manual_seed(5)
    flat = wide_ref.init_params(f, h, nl)
    x, y = torch.randn(b, f), (torch.rand(b) < 0.3).long()
    masks = [torch.from_numpy(wide_ref.keep_mask(42, 7, l + 1, b, h, 0.2)) for l in range(nl)]
    got = wide_ref.loss_and_grads(flat, x, y, f, h, nl, dropout=True, seed=42, step=7, emulate_bf16=False)
    want = wide_ref.torch_autograd_grads(flat, x, y, f, h, nl, masks)
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=2e-4, atol=1e-8)
    # the bf16-rounded variant stays close to fp32 (this is the tolerance the CUDA path is held to against fp32)
    e = wide_ref.loss_and_grads(flat, x, y, f, h, nl, dropout=True, seed=42, step=7, emulate_bf16=True)
    assert abs(float(e[-1]) - float(want[-1])) < 2e-2 * abs(float(want[-1]))
    cos = float(torch.dot(e[:-1], want[:-1]) / (e[:-1].norm() * want[:-1].norm()))
    assert cos > 0.999


def test_wide_keep_mask_contract():
    from oracle import philox_ref, wide_ref

    m = wide_ref.keep_mask(42, 3, 2, 64, 256, 0.2)
    assert m.shape == (64, 256) and abs(m.mean() - 0.8) < 0.01
    # element (row 5, col 19): counter (5, 19 // 8, layer, step); e = 3 -> word 1, high half
    w = philox_ref.philox4x32_10(np.uint32(5), np.uint32(2), np.uint32(2), np.uint32(3), 42, 0)
    bits = (int(w[1]) >> 16) & 0xFFFF
    assert bool(m[5, 19]) == (bits < 52429)
    assert not np.array_equal(m, wide_ref.keep_mask(42, 4, 2, 64, 256, 0.2))
    assert wide_ref.keep_mask(42, 3, 2, 8, 64, 0.0).all()
