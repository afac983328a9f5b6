"""CPU tests of the host side: index streams, checkpoint schema, metric names, dataset type, and the N>1
protocol (bucket of P+1 pre-scaled sums, SUM all-reduce, Adam) over gloo with world_size 2."""
import json
import os
import shutil
import tempfile

import numpy as np
import pytest
import torch

from oracle import preprocess_ref, synth, train_ref

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_index_streams_match_torch_samplers():
    from torch.utils.data import DistributedSampler, TensorDataset, random_split
    from wxddp import trainer

    torch.manual_seed(42)
    tr, va = trainer.split_indices(1000)
    torch.manual_seed(42)
    full = TensorDataset(torch.zeros(1000, 1))
    a, b = random_split(full, [800, 200])
    assert tr == a.indices and va == b.indices
    for world in (2, 4, 8):
        for rank in range(world):
            s = DistributedSampler(list(range(801)), num_replicas=world, rank=rank, shuffle=True, seed=42)
            s.set_epoch(3)
            assert trainer.distributed_order(801, 3, rank, world, shuffle=True) == list(iter(s))
            s = DistributedSampler(list(range(201)), num_replicas=world, rank=rank, shuffle=False)
            assert trainer.distributed_order(201, 0, rank, world, shuffle=False) == list(iter(s))
    # RandomSampler under DataLoader(shuffle=True): same draws from the global generator
    from torch.utils.data import DataLoader
    torch.manual_seed(7)
    dl = DataLoader(list(range(50)), batch_size=50, shuffle=True)
    want = next(iter(dl)).tolist()
    torch.manual_seed(7)
    trainer._draw_loader_base_seed()
    assert trainer.random_sampler_order(50) == want


def test_init_params_like_reference_matches_golden_setup():
    from wxddp import trainer
    from wxddp.engine import init_params_like_reference

    g = np.load(os.path.join(GOLDEN, "setup.npz"))
    for f in (5, 20):
        torch.manual_seed(42)
        tr, va = trainer.split_indices(1000)
        p = init_params_like_reference(f)
        assert tr == g[f"train_idx_F{f}"].tolist()
        np.testing.assert_array_equal(p.numpy(), g[f"params_F{f}"])       # bit-identical to the reference module's init


def test_checkpoint_schema_and_top1_policy():
    from wxddp.module import WeatherClassifier
    from wxddp.surface import TopOneCheckpoint, build_checkpoint, flat_from_state_dict

    tmp = tempfile.mkdtemp()
    try:
        f = 5
        torch.manual_seed(1)
        flat = torch.randn(train_ref.n_params(f))
        m, v = torch.rand_like(flat), torch.rand_like(flat)
        ck = TopOneCheckpoint(os.path.join(tmp, "models"))
        for epoch, vl in enumerate([0.52, 0.449, 0.47, 0.301]):
            ck.on_validation_end(lambda cb: build_checkpoint(flat + epoch, m, v, f, epoch, 100 * (epoch + 1), callbacks=cb),
                                 epoch, vl)
        files = sorted(os.listdir(os.path.join(tmp, "models")))
        assert files == ["last.ckpt", "weather-best-epoch=03-val_loss=0.30.ckpt"]          # top-1 + last (:103-110)
        assert ck.best_model_path.endswith("weather-best-epoch=03-val_loss=0.30.ckpt")
        d = torch.load(ck.best_model_path, weights_only=False)
        for key in ["epoch", "global_step", "pytorch-lightning_version", "state_dict", "loops", "callbacks",
                    "optimizer_states", "lr_schedulers", "hparams_name", "hyper_parameters"]:
            assert key in d
        assert d["pytorch-lightning_version"] == "2.1.0" and d["hyper_parameters"] == {"input_dim": 5}
        assert list(d["state_dict"]) == ["net.0.weight", "net.0.bias", "net.3.weight", "net.3.bias"]
        assert d["state_dict"]["net.0.weight"].shape == (64, 5) and d["state_dict"]["net.3.weight"].shape == (2, 64)
        assert d["epoch"] == 3 and d["global_step"] == 400
        # what score.py does (dags/azure_auto_deploy.py:93): load_from_checkpoint(path, input_dim=5)
        model = WeatherClassifier.load_from_checkpoint(ck.best_model_path, input_dim=5)
        np.testing.assert_array_equal(flat_from_state_dict(model.state_dict()).numpy(), (flat + 3).numpy())
        # and a plain torch module with the reference's layer names accepts the state_dict unchanged
        ref = train_ref.WeatherNet(5)
        ref.load_state_dict(d["state_dict"])
        # the Adam state is a loadable torch optimizer state
        opt = torch.optim.Adam(ref.parameters(), lr=0.01)
        opt.load_state_dict(d["optimizer_states"][0])
        # NaN never becomes best
        ck.on_validation_end(lambda cb: build_checkpoint(flat, m, v, f, 4, 500, callbacks=cb), 4, float("nan"))
        assert ck.best_model_path.endswith("epoch=03-val_loss=0.30.ckpt")
    finally:
        shutil.rmtree(tmp)


def test_metric_logger_fallback_uses_reference_names():
    from wxddp.surface import MetricLogger

    tmp = tempfile.mkdtemp()
    try:
        lg = MetricLogger("file:/nonexistent", tmp, params={"input_dim": 5})
        if lg.mlflow is not None:
            pytest.skip("mlflow installed")
        lg.log_metrics({"train_loss": 0.5, "epoch": 0}, step=5)
        lg.log_metrics({"val_loss": 0.4, "val_acc": 0.9, "epoch": 0}, step=29)
        lg.log_artifact("/x/last.ckpt", "best_checkpoints")
        lg.close()
        recs = [json.loads(l) for l in open(lg.path)]
        assert os.path.basename(lg.path).startswith("weather_forecasting-")
        assert recs[0]["params"] == {"input_dim": 5}
        assert recs[1]["train_loss"] == 0.5 and recs[1]["step"] == 5
        assert set(recs[2]) >= {"val_loss", "val_acc", "epoch"}
        assert recs[3]["artifact_path"] == "best_checkpoints"
    finally:
        shutil.rmtree(tmp)


def test_weatherdataset_type_matches_reference_golden():
    from wxddp.module import WeatherDataset

    g = np.load(os.path.join(GOLDEN, "dataset_F5.npz"))
    tmp = tempfile.mkdtemp()
    try:
        preprocess_ref.write_parquet_dir(tmp, synth.feature_names(5), g["norm"], g["labels32"], n_parts=3)
        ds = WeatherDataset(tmp)
        assert ds.features.dtype == torch.float32 and ds.labels.dtype == torch.int64 and len(ds) == 600
        np.testing.assert_array_equal(ds.features.numpy(), g["features"])      # the reference class's own output
        np.testing.assert_array_equal(ds.labels.numpy(), g["labels"])
        x, y = ds[3]
        assert x.shape == (5,) and y.shape == ()
        with pytest.raises(FileNotFoundError):
            WeatherDataset(os.path.join(tmp, "nope"))
    finally:
        shutil.rmtree(tmp)


def test_module_has_no_cpu_path():
    from wxddp import WxError
    from wxddp.module import WeatherClassifier

    m = WeatherClassifier(5)
    assert [k for k, _ in m.named_parameters()] == ["net.0.weight", "net.0.bias", "net.3.weight", "net.3.bias"]
    assert isinstance(m.configure_optimizers(), torch.optim.Adam)
    with pytest.raises(WxError):
        m.training_step((torch.zeros(4, 5), torch.zeros(4, dtype=torch.long)), 0)


# ---- world_size 2 over gloo: the DDP bucket protocol of StepEngine/trainer.fit ------------------------------
class _CpuDoubleEngine:
    """Test double with StepEngine's interface: the local gradient comes from the oracle instead of the CUDA
    kernel, everything else (bucket layout, pre-scaling by 1/(W*B), SUM all-reduce, Adam, eval reduce) is the
    protocol under test."""

    def __init__(self, n_features, device, world, rank, process_group, params, seed):
        self.F, self.P = n_features, train_ref.n_params(n_features)
        self.device = torch.device("cpu")
        self.world, self.rank, self.dist = world, rank, process_group
        self.params = params.clone()
        self.m, self.v = torch.zeros(self.P), torch.zeros(self.P)
        self.bucket = torch.zeros(self.P + 1)
        self.steps_done = 0
        if world > 1:
            self.dist.broadcast(self.params, src=0)

    def step(self, x, y, *, row_idx=None, dropout=None, keep_mask=None, **kw):
        xb, yb = x[row_idx], y[row_idx]
        b = len(row_idx)
        bits = keep_mask.numpy().view(np.uint64)
        mask = torch.from_numpy(((bits[:, None] >> np.arange(64, dtype=np.uint64)) & np.uint64(1)).astype(np.float32))
        loss, grad = train_ref.loss_and_grads(self.params, self.F, xb, yb, mask)
        scale = 1.0 / (self.world * b)
        self.bucket[:self.P] = grad * b * scale            # sum over rows, pre-scaled (wx_mlp_grad's contract)
        self.bucket[self.P] = loss * b * scale
        if self.world > 1:
            self.dist.all_reduce(self.bucket)
        self.steps_done += 1
        train_ref.adam_update(self.params, self.bucket[:self.P].clone(), self.m, self.v, self.steps_done)
        return self.bucket[self.P:self.P + 1]

    def evaluate(self, x, y, *, row_idx=None, **kw):
        loss, correct = train_ref.eval_batch(self.params, self.F, x[row_idx], y[row_idx])
        tot = torch.tensor([float(loss) * len(row_idx), float(correct), float(len(row_idx))], dtype=torch.float64)
        if self.world > 1:
            self.dist.all_reduce(tot)
        return float(tot[0] / tot[2]), float(tot[1] / tot[2])


def _gloo_worker(rank, world, port, feats, labels, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from wxddp import trainer

    mine = trainer.fit(feats, labels, device="cpu", rank=rank, world=world, dist=dist, batch_size=4, max_epochs=2,
                       parity=True, verbose=False, engine_factory=_CpuDoubleEngine)
    ref = train_ref.fit(feats, labels, rank=rank, world=world, use_ddp=True, max_epochs=2)    # real torch DDP over gloo
    q.put((rank, mine["train_loss"], mine["val_loss"], mine["val_acc"], mine["params"].numpy(),
           ref["train_loss"], ref["val_loss"], ref["val_acc"], ref["params"].numpy()))
    dist.destroy_process_group()


def test_ddp_protocol_world2_gloo_matches_torch_ddp():
    import torch.multiprocessing as mp

    t = synth.synth_table(31, 203, 5)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    feats, labels = torch.from_numpy(pp["features"]), torch.from_numpy(pp["labels"])
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, 29653, feats, labels, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in range(2)], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, tl, vl, va, params, rtl, rvl, rva, rparams in res:
        assert len(tl) == len(rtl) == 2 * 21           # ceil(162/2 / 4) steps per epoch per rank
        np.testing.assert_allclose(tl, rtl, rtol=2e-5)   # rank-mean train_loss (sync_dist) per step
        np.testing.assert_allclose(vl, rvl, rtol=1e-5)
        np.testing.assert_allclose(va, rva, rtol=1e-6)
        np.testing.assert_allclose(params, rparams, rtol=1e-4, atol=1e-6)
    np.testing.assert_array_equal(res[0][4], res[1][4])      # replicas stay bit-identical across ranks
