"""GPU tests of the host mirror: the fit loop, the CLI drop-in, the checkpoint/metric surface, the Lightning-shaped
module and the ETL front end -- all running the CUDA kernels through the C ABI."""
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np
import pytest
import torch

from oracle import preprocess_ref, synth, train_ref

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def test_fit_parity_mode_tracks_reference_loop_w1():
    """trainer.fit(parity=True) vs the 2-epoch loop driven through the reference's own classes (loop_W1.npz)."""
    from wxddp import trainer

    g = np.load(os.path.join(GOLDEN, "loop_W1.npz"))
    h = trainer.fit(torch.from_numpy(g["features"]), torch.from_numpy(g["labels"]), device=0, batch_size=4,
                    max_epochs=2, parity=True, sanity_check=False, verbose=False)
    assert h["steps"] == len(g["train_loss"]) == 82
    np.testing.assert_allclose(h["train_loss"], g["train_loss"], rtol=2e-4, atol=1e-6)      # free-running, 82 Adam steps
    np.testing.assert_allclose(h["val_loss"], g["val_loss"], rtol=2e-4)
    np.testing.assert_allclose(h["val_acc"], g["val_acc"], rtol=1e-6)
    np.testing.assert_allclose(h["params"].numpy(), g["params"], rtol=2e-3, atol=2e-5)


def test_fit_parity_mode_matches_oracle_fit_with_sanity_draw():
    from wxddp import trainer

    t = synth.synth_table(8, 403, 20, const_col=19)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    feats, labels = torch.from_numpy(pp["features"]), torch.from_numpy(pp["labels"])
    ref = train_ref.fit(feats, labels, max_epochs=1, batch_size=16)
    h = trainer.fit(feats, labels, device=0, batch_size=16, max_epochs=1, parity=True, verbose=False)
    np.testing.assert_allclose(h["train_loss"], ref["train_loss"], rtol=2e-4, atol=1e-6)
    np.testing.assert_allclose(h["val_loss"], ref["val_loss"], rtol=2e-4)


def _make_processed_dir(tmp, n=3000, f=5, seed=3):
    t = synth.synth_table(seed, n, f)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    preprocess_ref.write_parquet_dir(os.path.join(tmp, "processed"), t["names"], pp["norm"], pp["labels32"], n_parts=2)
    return t, pp


def test_cli_drop_in_writes_checkpoints_and_metrics():
    """`python jobs/train_lightning_ddp.py` with no argv: exit 0, top-1 + last.ckpt, metric names, loadable model."""
    tmp = tempfile.mkdtemp()
    try:
        t, pp = _make_processed_dir(tmp)
        env = dict(os.environ, WX_DATA_DIR=os.path.join(tmp, "processed"), WX_MODEL_DIR=os.path.join(tmp, "models"),
                   WX_MAX_EPOCHS="3", WX_METRICS_DIR=os.path.join(tmp, "metrics"), WORLD_SIZE="1")
        r = subprocess.run([sys.executable, os.path.join(ROOT, "jobs", "train_lightning_ddp.py")], env=env,
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        assert "Training with 5 features" in r.stdout and "Training finished." in r.stdout
        files = sorted(os.listdir(os.path.join(tmp, "models")))
        assert "last.ckpt" in files and len(files) == 2 and files[1].startswith("weather-best-epoch=0")
        mfile = [f for f in os.listdir(os.path.join(tmp, "metrics")) if f.endswith(".jsonl")]
        recs = [json.loads(l) for l in open(os.path.join(tmp, "metrics", mfile[0]))]
        steps = [r_["step"] for r_ in recs if "train_loss" in r_]
        assert steps[:3] == [4, 9, 14]     # log_every_n_steps=5 (:139) on Lightning's step axis (_batches_that_stepped)
        vals = [r_ for r_ in recs if "val_loss" in r_]
        assert len(vals) == 3 and all("val_acc" in v for v in vals)
        assert vals[-1]["val_loss"] < 0.45 and vals[-1]["val_acc"] > 0.8   # it learns
        assert any(r_.get("artifact_path") == "best_checkpoints" for r_ in recs)
        # the deploy consumer's call (score.py): load_from_checkpoint(path, input_dim=5), softmax(model(x))
        from wxddp.module import WeatherClassifier
        best = os.path.join(tmp, "models", files[1])
        model = WeatherClassifier.load_from_checkpoint(best, input_dim=5).cuda().eval()
        x = torch.from_numpy(pp["features"][:256]).cuda()
        with torch.no_grad():
            probs = torch.softmax(model(x), dim=1).cpu()
        ref = train_ref.WeatherNet(5)
        ref.load_state_dict(torch.load(best, weights_only=False)["state_dict"])
        ref.eval()
        with torch.no_grad():
            want = torch.softmax(ref(torch.from_numpy(pp["features"][:256])), dim=1)
        np.testing.assert_allclose(probs.numpy(), want.numpy(), rtol=1e-5, atol=1e-7)
    finally:
        shutil.rmtree(tmp)


def test_cli_fails_loudly_when_data_is_missing():
    tmp = tempfile.mkdtemp()
    try:
        env = dict(os.environ, WX_DATA_DIR=os.path.join(tmp, "nothing"), WX_MODEL_DIR=os.path.join(tmp, "models"),
                   WX_DISABLE_MLFLOW="1")
        r = subprocess.run([sys.executable, os.path.join(ROOT, "jobs", "train_lightning_ddp.py")], env=env,
                           capture_output=True, text=True, timeout=300)
        assert r.returncode != 0 and "CRITICAL ERROR: Data not found" in r.stderr
    finally:
        shutil.rmtree(tmp)


def test_lightning_shaped_module_runs_the_kernels():
    """WeatherClassifier.training_step/validation_step on CUDA batches, driven the way Lightning drives them."""
    from wxddp.module import WeatherClassifier

    torch.manual_seed(0)
    m = WeatherClassifier(5).cuda()
    ref = train_ref.WeatherNet(5)
    ref.load_state_dict(m.state_dict())
    g = torch.Generator().manual_seed(1)
    x = torch.randn(64, 5, generator=g); y = (torch.rand(64, generator=g) < 0.3).long()
    # eval mode -> dropout off -> comparable with autograd on the CPU copy
    m.eval(); ref.eval()
    loss = m.training_step((x.cuda(), y.cuda()), 0)
    loss.backward()
    rl = torch.nn.functional.cross_entropy(ref(x), y)
    rl.backward()
    np.testing.assert_allclose(loss.item(), rl.item(), rtol=1e-5)
    for (k, p), (_, q) in zip(m.named_parameters(), ref.named_parameters()):
        np.testing.assert_allclose(p.grad.cpu().numpy(), q.grad.numpy(), rtol=1e-5, atol=1e-7, err_msg=k)
    # optimizer from configure_optimizers steps the same parameters Lightning would step
    opt = m.configure_optimizers(); opt.step()
    ropt = torch.optim.Adam(ref.parameters(), lr=0.01); ropt.step()
    big = (ref.net[0].weight.grad.abs() > 1e-6).numpy()
    np.testing.assert_allclose(m.net[0].weight.detach().cpu().numpy()[big], ref.net[0].weight.detach().numpy()[big],
                               rtol=1e-5, atol=2e-7)
    vl = m.validation_step((x.cuda(), y.cuda()), 0)
    with torch.no_grad():
        want = torch.nn.functional.cross_entropy(ref(x), y)
        acc = (ref(x).argmax(1) == y).float().mean()
    np.testing.assert_allclose(vl.item(), want.item(), rtol=1e-5)
    assert abs(m.logged["val_acc"].item() - acc.item()) < 1e-6
    # train mode uses the in-kernel dropout stream and still yields finite gradients
    m.train(); m.zero_grad()
    l2 = m.training_step((x.cuda(), y.cuda()), 1)
    l2.backward()
    assert torch.isfinite(l2) and all(torch.isfinite(p.grad).all() for p in m.parameters())


def test_etl_front_end_matches_oracle_and_feeds_the_trainer():
    import pyarrow as pa

    from wxddp import etl
    from wxddp.module import WeatherDataset

    t = synth.synth_table(21, 5000, 5, null_rate=0.01)
    cols = {n: pa.array(t["raw"][j], mask=np.isnan(t["raw"][j])) for j, n in enumerate(t["names"])}
    cols["Rain"] = pa.array(t["labels"], type=pa.string())
    table = pa.table(cols)
    res = etl.preprocess_table(table, device=0)
    want = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    np.testing.assert_array_equal(res["labels32"], want["labels32"])
    np.testing.assert_allclose(res["norm"], want["norm"], rtol=1e-12, atol=1e-12, equal_nan=True)
    np.testing.assert_array_equal(res["nullmask"].cpu().numpy().view(np.uint32), want["nullmask"])
    # the float64 `*_norm` columns come from wx_zscore_f64: given the SAME statistics they are the oracle's float64
    # expression (col - mean) / std_val bit for bit (preprocess.py:38-41), nulls included
    for j in range(5):
        z = preprocess_ref.zscore(t["raw"][j], res["stats"][j, 1], res["stats"][j, 3])
        np.testing.assert_array_equal(res["norm"][j].view(np.int64)[~np.isnan(z)], z.view(np.int64)[~np.isnan(z)])
        np.testing.assert_array_equal(np.isnan(res["norm"][j]), np.isnan(z))
    tmp = tempfile.mkdtemp()
    try:
        # the Arrow validity bitmaps that came in with the table are the null bitmaps of the written columns
        etl.write_parquet_dir(os.path.join(tmp, "data.parquet"), t["names"], res["norm"], res["labels32"],
                              validity=res["validity"])
        ds = WeatherDataset(tmp)
        assert ds.feature_cols == [f"{n}_norm" for n in t["names"]]
        ulp = np.abs(ds.features.numpy().view(np.int32).astype(np.int64) - want["features"].view(np.int32).astype(np.int64))
        ok = ~np.isnan(want["features"])
        assert ulp[ok].max() <= 1 and np.array_equal(np.isnan(ds.features.numpy()), ~ok)
        np.testing.assert_array_equal(ds.labels.numpy(), want["labels"])
        # the same directory through the GPU ingest path (Arrow buffers -> pinned chunks -> HBM -> cast + pack kernel):
        # bit-identical to what the reference's pandas/torch casts produce (train_lightning_ddp.py:31,45-46)
        from wxddp import ingest
        got = ingest.load_processed(tmp, device=0)
        assert got["feature_cols"] == ds.feature_cols
        np.testing.assert_array_equal(got["features"].cpu().numpy().view(np.int32), ds.features.numpy().view(np.int32))
        np.testing.assert_array_equal(got["labels"].cpu().numpy(), ds.labels.numpy())
        assert got["info"]["rows"] == 5000 and got["info"]["bytes_h2d"] >= 5000 * (5 * 8 + 4)
    finally:
        shutil.rmtree(tmp)


def test_ingest_of_a_multi_part_directory_with_unaligned_part_boundaries():
    """Spark writes several part files; their row counts are not multiples of 8, so the validity bitmaps of consecutive
    parts do not join on byte boundaries."""
    from wxddp import ingest
    from wxddp.module import WeatherDataset

    t = synth.synth_table(5, 10007, 20, null_rate=0.02, const_col=19)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    tmp = tempfile.mkdtemp()
    try:
        preprocess_ref.write_parquet_dir(os.path.join(tmp), t["names"], pp["norm"], pp["labels32"], n_parts=3)
        ds = WeatherDataset(tmp)
        got = ingest.load_processed(tmp, device=0)
        np.testing.assert_array_equal(got["features"].cpu().numpy().view(np.int32), ds.features.numpy().view(np.int32))
        np.testing.assert_array_equal(got["labels"].cpu().numpy(), ds.labels.numpy())
        assert got["info"]["files"] == 3
    finally:
        shutil.rmtree(tmp)


def test_host_buffer_steps_sync_and_pipelined_agree():
    """StepEngine.step_host (synchronous) and submit_host/drain (H2D of batch i overlapped with the kernel of batch
    i-1, loss one call late) run the same arithmetic in the same order as step_raw on device-resident columns."""
    from wxddp.engine import StepEngine

    f, b, n_steps = 20, 4096, 7
    t = synth.synth_table(21, b * n_steps, f, null_rate=0.01)
    raw = torch.from_numpy(t["raw"].astype(np.float32))
    y = torch.from_numpy(preprocess_ref.label_encode(t["labels"]).astype(np.int64))
    torch.manual_seed(42)
    p0 = None
    engines = []
    for _ in range(3):
        e = StepEngine(f, device=0, params=p0)
        p0 = e.params.cpu() if p0 is None else p0
        engines.append(e)
    dev_cols = [raw[j].cuda() for j in range(f)]
    stats = engines[0].fit_stats(dev_cols)
    for e in engines[1:]:
        e.set_stats(stats)
    host = [(raw[:, i * b:(i + 1) * b].contiguous().pin_memory(), y[i * b:(i + 1) * b].contiguous().pin_memory())
            for i in range(n_steps)]
    yd = y.cuda()
    ref = [float(engines[0].step_raw(dev_cols, yd, row0=i * b, batch=b).item()) for i in range(n_steps)]
    sync = [engines[1].step_host(c, l) for c, l in host]
    piped = []
    for c, l in host:
        r = engines[2].submit_host(c, l)
        if r is not None:
            piped.append(r)
    piped.append(engines[2].drain())
    assert engines[2].drain() is None
    # int32 labels (the dtype Spark wrote) through the same pipeline: widened on the device, same losses
    e32 = StepEngine(f, device=0, params=p0)
    e32.set_stats(stats)
    got32 = [e32.submit_host(c, l.to(torch.int32).pin_memory()) for c, l in host]
    got32 = [v for v in got32 if v is not None] + [e32.drain()]
    assert got32 == piped
    e8 = StepEngine(f, device=0, params=p0)                     # one byte per label
    e8.set_stats(stats)
    got8 = [e8.submit_host(c, l.to(torch.uint8).pin_memory()) for c, l in host]
    got8 = [v for v in got8 if v is not None] + [e8.drain()]
    assert got8 == piped
    assert sync == piped and len(piped) == n_steps
    assert torch.equal(engines[1].params, engines[2].params)
    # device-resident windows of one big table vs per-batch staging buffers: same rows, same kernel
    np.testing.assert_allclose(sync, ref, rtol=1e-6)
    np.testing.assert_allclose(engines[1].params.cpu().numpy(), engines[0].params.cpu().numpy(), rtol=1e-5, atol=1e-7)
