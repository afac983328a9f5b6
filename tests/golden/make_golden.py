"""Generate tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN CLASSES (authoring container only).

Run:  python tests/golden/make_golden.py            (needs /root/reference; not run on the GPU box)

`/root/reference/jobs/train_lightning_ddp.py` imports `pytorch_lightning` and `mlflow` at module
scope (:6-11); neither is installable here.  The classes on the hot path use only a sliver of them:
`pl.LightningModule` as an `nn.Module` base with `save_hyperparameters()` and `log()`, and
`pl.seed_everything`.  This script registers stand-in modules that supply exactly that sliver, then
imports the reference file unmodified from where it lies and drives its `WeatherDataset`,
`WeatherClassifier.training_step / validation_step / configure_optimizers`.  All arithmetic in the
vectors therefore comes from the reference's code + torch, none from our oracle -- except the
dropout-mask replay (`oracle.train_ref.draw_dropout_mask`), which is asserted here to reproduce the
reference module's loss bit-for-bit before the mask is stored.

Inputs are synthetic (`oracle/synth.py`) because the reference ships no data.
"""
from __future__ import annotations

import importlib.util
import os
import random
import shutil
import sys
import tempfile
import types

import numpy as np
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import preprocess_ref, synth, train_ref  # noqa: E402

REF = "/root/reference/jobs/train_lightning_ddp.py"
OUT = os.path.dirname(os.path.abspath(__file__))


def import_reference():
    pl = types.ModuleType("pytorch_lightning")

    class LightningModule(nn.Module):
        def save_hyperparameters(self, *a, **k):
            pass

        def log(self, name, value, **k):
            self.__dict__.setdefault("logged", {})[name] = value

    def seed_everything(s):
        random.seed(s)
        np.random.seed(s)
        torch.manual_seed(s)
        os.environ["PL_GLOBAL_SEED"] = str(s)
        return s

    pl.LightningModule = LightningModule
    pl.seed_everything = seed_everything
    pl.Trainer = object
    sys.modules["pytorch_lightning"] = pl
    for name, attrs in [("pytorch_lightning.strategies", ["DDPStrategy"]),
                        ("pytorch_lightning.loggers", ["MLFlowLogger"]),
                        ("pytorch_lightning.callbacks", ["ModelCheckpoint"])]:
        m = types.ModuleType(name)
        for a in attrs:
            setattr(m, a, type(a, (), {}))
        sys.modules[name] = m
    sys.modules["mlflow"] = types.ModuleType("mlflow")
    spec = importlib.util.spec_from_file_location("ref_train_lightning_ddp", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)          # runs pl.seed_everything(42) at :14
    return mod


def make_table(seed, n, f, null_rate=0.0, const_col=None):
    t = synth.synth_table(seed, n, f, null_rate=null_rate, const_col=const_col)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    return t, pp


def gen_steps(ref, f, batch, steps, seed, tag):
    """Teacher-forcing vectors: per step params-before, batch, mask, loss, grads, params/m/v after."""
    t, pp = make_table(1234 + f, 4096, f, const_col=(f - 1 if f > 5 else None))
    feats = torch.from_numpy(pp["features"])
    labels = torch.from_numpy(pp["labels"])
    torch.manual_seed(seed)
    model = ref.WeatherClassifier(input_dim=f)
    opt = model.configure_optimizers()
    model.train()
    g = torch.Generator().manual_seed(seed + 1)
    rec = {k: [] for k in ["p_before", "x", "y", "mask", "loss", "grad", "p_after", "m_after", "v_after"]}
    for s in range(steps):
        rows = torch.randperm(feats.shape[0], generator=g)[:batch]
        x, y = feats[rows], labels[rows]
        p_before = train_ref.flatten(model)
        st = torch.get_rng_state()
        mask = train_ref.draw_dropout_mask(batch)
        torch.set_rng_state(st)
        opt.zero_grad(set_to_none=True)
        loss = model.training_step((x, y), s)               # reference :66-71
        loss.backward()
        # the replayed mask must reproduce the reference module's loss exactly
        l2, g2 = train_ref.loss_and_grads(p_before, f, x, y, mask)
        assert torch.equal(l2, loss.detach()), (float(l2), float(loss))
        grad = torch.cat([p.grad.reshape(-1) for p in model.parameters()]).clone()
        opt.step()                                          # Adam(lr=0.01) from reference :88
        rec["p_before"].append(p_before.numpy())
        rec["x"].append(x.numpy()); rec["y"].append(y.numpy())
        rec["mask"].append(train_ref.pack_mask(mask))
        rec["loss"].append(np.float32(loss.item())); rec["grad"].append(grad.numpy())
        rec["p_after"].append(train_ref.flatten(model).numpy())
        rec["m_after"].append(torch.cat([opt.state[p]["exp_avg"].reshape(-1) for p in model.parameters()]).numpy().copy())
        rec["v_after"].append(torch.cat([opt.state[p]["exp_avg_sq"].reshape(-1) for p in model.parameters()]).numpy().copy())
    np.savez_compressed(os.path.join(OUT, f"steps_{tag}.npz"), f=f, batch=batch,
                        **{k: np.stack(v) for k, v in rec.items()})
    print("wrote steps", tag, "loss[0..3] =", [float(v) for v in rec["loss"][:3]])


def gen_eval(ref, f, tag):
    t, pp = make_table(777 + f, 1024, f)
    feats = torch.from_numpy(pp["features"]); labels = torch.from_numpy(pp["labels"])
    torch.manual_seed(7)
    model = ref.WeatherClassifier(input_dim=f)
    model.eval()
    flat = train_ref.flatten(model).numpy()
    out = {"loss": [], "acc": [], "lo": [], "hi": []}
    with torch.no_grad():
        for lo, hi in [(0, 4), (4, 8), (8, 72), (72, 1024), (0, 1024), (5, 6)]:
            model.validation_step((feats[lo:hi], labels[lo:hi]), 0)          # reference :73-85
            out["loss"].append(np.float32(model.logged["val_loss"].item()))
            out["acc"].append(np.float32(model.logged["val_acc"].item()))
            out["lo"].append(lo); out["hi"].append(hi)
    np.savez_compressed(os.path.join(OUT, f"eval_{tag}.npz"), f=f, params=flat, x=pp["features"], y=pp["labels"],
                        **{k: np.array(v) for k, v in out.items()})
    print("wrote eval", tag, out["loss"][:2], out["acc"][:2])


def gen_dataset(ref):
    """WeatherDataset (:16-49) over a Spark-shaped parquet directory with nulls and two part files."""
    t = synth.synth_table(99, 600, 5, null_rate=0.02)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    tmp = tempfile.mkdtemp()
    try:
        preprocess_ref.write_parquet_dir(tmp, t["names"], pp["norm"], pp["labels32"], n_parts=3)
        ds = ref.WeatherDataset(tmp)
        feats, labels = ds.features.numpy().copy(), ds.labels.numpy().copy()
        assert len(ds) == 600 and ds[3][0].shape == (5,)
        for bad, exc in [(os.path.join(tmp, "nope"), FileNotFoundError)]:
            try:
                ref.WeatherDataset(bad)
                raise AssertionError("expected failure")
            except exc:
                pass
    finally:
        shutil.rmtree(tmp)
    lab = np.array([("\0" if s is None else s) for s in t["labels"]])
    np.savez_compressed(os.path.join(OUT, "dataset_F5.npz"), raw=t["raw"], label_strings=lab,
                        norm=pp["norm"], labels32=pp["labels32"], features=feats, labels=labels)
    print("wrote dataset_F5", feats.shape, feats.dtype, labels.dtype, "nan cells", int(np.isnan(feats).sum()))


def gen_setup(ref):
    """Seed (:14) -> random_split (:117-119) -> model init (:126): pins the RNG draw order."""
    from torch.utils.data import TensorDataset, random_split

    out = {}
    for f, n in [(5, 1000), (20, 1000)]:
        ref.pl.seed_everything(42)
        full = TensorDataset(torch.zeros(n, f), torch.zeros(n, dtype=torch.long))
        train_size = int(0.8 * len(full))
        tr, va = random_split(full, [train_size, len(full) - train_size])
        model = ref.WeatherClassifier(input_dim=f)
        out[f"train_idx_F{f}"] = np.array(tr.indices)
        out[f"val_idx_F{f}"] = np.array(va.indices)
        out[f"params_F{f}"] = train_ref.flatten(model).numpy()
    np.savez_compressed(os.path.join(OUT, "setup.npz"), **out)
    print("wrote setup")


def gen_loop(ref):
    """A hand-rolled 2-epoch loop over the reference's model/optimizer with torch's own DataLoader
    (shuffle=True, batch 4, :122-123).  Pins `oracle.train_ref.fit(world=1)`'s RNG bookkeeping against
    torch's DataLoader; Lightning's own loop is NOT exercised (not installable) -- stays unpinned."""
    from torch.utils.data import DataLoader, TensorDataset, random_split

    t, pp = make_table(4321, 203, 5)
    feats = torch.from_numpy(pp["features"]); labels = torch.from_numpy(pp["labels"])
    ref.pl.seed_everything(42)
    full = TensorDataset(feats, labels)
    train_size = int(0.8 * len(full))
    tr, va = random_split(full, [train_size, len(full) - train_size])
    train_loader = DataLoader(tr, batch_size=4, shuffle=True, num_workers=0)
    val_loader = DataLoader(va, batch_size=4, shuffle=False, num_workers=0)
    model = ref.WeatherClassifier(input_dim=5)
    opt = model.configure_optimizers()
    losses, vloss, vacc = [], [], []
    for epoch in range(2):
        model.train()
        for bi, batch in enumerate(train_loader):
            opt.zero_grad(set_to_none=True)
            loss = model.training_step(batch, bi)
            loss.backward()
            opt.step()
            losses.append(np.float32(loss.item()))
        model.eval()
        tl = ta = 0.0
        tn = 0
        with torch.no_grad():
            for bi, batch in enumerate(val_loader):
                model.validation_step(batch, bi)
                k = len(batch[1])
                tl += float(model.logged["val_loss"]) * k; ta += float(model.logged["val_acc"]) * k; tn += k
        vloss.append(tl / tn); vacc.append(ta / tn)
    np.savez_compressed(os.path.join(OUT, "loop_W1.npz"), features=pp["features"], labels=pp["labels"],
                        train_loss=np.array(losses), val_loss=np.array(vloss), val_acc=np.array(vacc),
                        params=train_ref.flatten(model).numpy())
    print("wrote loop_W1", len(losses), "steps; val", vloss, vacc)


if __name__ == "__main__":
    ref = import_reference()
    gen_setup(ref)
    gen_dataset(ref)
    gen_steps(ref, 5, 4, 24, 42, "F5_B4")
    gen_steps(ref, 20, 4, 12, 43, "F20_B4")
    gen_steps(ref, 20, 256, 12, 44, "F20_B256")
    gen_steps(ref, 5, 1000, 6, 45, "F5_B1000")
    gen_eval(ref, 5, "F5")
    gen_eval(ref, 20, "F20")
    gen_loop(ref)
