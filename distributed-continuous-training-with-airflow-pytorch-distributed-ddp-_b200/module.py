"""The reference's two Python types, kept as a drop-in surface (jobs/train_lightning_ddp.py:16-88).

`WeatherDataset` is the reference class with the same constructor, attributes and errors.
`WeatherClassifier` has the reference's constructor, `net`, `forward`, `training_step`, `validation_step`,
`configure_optimizers`, `hparams` and `load_from_checkpoint`; it derives from `pl.LightningModule` when
pytorch-lightning is importable and from a minimal stand-in otherwise (Lightning is not in this image).
On CUDA tensors its `training_step` / `validation_step` run the hand-written kernels (through a
`torch.autograd.Function`, so Lightning's `loss.backward()` + `optimizer.step()` keep working); the fast
path that bypasses autograd entirely is `wxddp.trainer.fit` / `StepEngine`.
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn
from torch.utils.data import Dataset

from . import _lib
from .surface import flat_from_state_dict

try:                                                   # pragma: no cover - Lightning absent in this image
    import pytorch_lightning as pl
    _Base = pl.LightningModule
    HAVE_LIGHTNING = True
except Exception:                                      # noqa: BLE001
    pl = None
    HAVE_LIGHTNING = False

    class _Base(nn.Module):
        """What the reference uses of LightningModule: hparams, save_hyperparameters(), log()."""

        def __init__(self):
            super().__init__()
            self.hparams = {}
            self.logged = {}

        def save_hyperparameters(self, **kw):
            self.hparams.update(kw)

        def log(self, name, value, **kw):
            self.logged[name] = value


class WeatherDataset(Dataset):
    """jobs/train_lightning_ddp.py:16-49, same behaviour: reads the Spark-written parquet DIRECTORY, picks
    every `*_norm` column in file order, casts float64 -> float32 features and int32 -> int64 labels."""

    def __init__(self, data_path):
        import pandas as pd

        parquet_path = os.path.join(data_path, "data.parquet")                              # :19
        if not os.path.exists(parquet_path):                                                # :22-26
            raise FileNotFoundError(
                f"CRITICAL ERROR: Data not found at {parquet_path}.\n"
                "Did the Spark preprocessing step finish successfully?")
        print(f"Loading data from: {parquet_path}")
        try:
            df = pd.read_parquet(parquet_path)                                              # :31
        except Exception as e:                                                              # :32-33
            raise RuntimeError(f"Failed to read Parquet file: {e}")
        feature_cols = [c for c in df.columns if c.endswith("_norm")]                       # :37
        if not feature_cols:                                                                # :39-40
            raise ValueError("CRITICAL ERROR: No columns ending with '_norm' found. Check Spark logic.")
        print(f"✓ Loaded {len(df)} rows.")
        print(f"✓ Training with {len(feature_cols)} features: {feature_cols}")
        self.feature_cols = feature_cols
        self.features = torch.tensor(df[feature_cols].to_numpy(dtype="float64", copy=True), dtype=torch.float32)  # :45
        self.labels = torch.tensor(df["label_encoded"].to_numpy(copy=True), dtype=torch.int64)                  # :46

    def __len__(self):
        return len(self.features)

    def __getitem__(self, idx):
        return self.features[idx], self.labels[idx]


class _FusedStepFn(torch.autograd.Function):
    """loss = mean CE of the MLP on a CUDA batch, gradients from wx_mlp_grad (one kernel for fwd+loss+bwd)."""

    @staticmethod
    def forward(ctx, flat, x, y, wctx, dropout, seed, step):
        out = wctx.mlp_grad(flat.detach().contiguous(), x.contiguous(), y.contiguous(), dropout=dropout, seed=seed,
                            step=step)
        ctx.save_for_backward(out)
        ctx.p = flat.numel()
        return out[ctx.p].clone()

    @staticmethod
    def backward(ctx, g):
        (out,) = ctx.saved_tensors
        return out[:ctx.p] * g, None, None, None, None, None, None


class WeatherClassifier(_Base):
    """jobs/train_lightning_ddp.py:51-88."""

    def __init__(self, input_dim):
        super().__init__()
        if HAVE_LIGHTNING:
            self.save_hyperparameters()
        else:
            self.save_hyperparameters(input_dim=input_dim)
        self.net = nn.Sequential(nn.Linear(input_dim, 64), nn.ReLU(), nn.Dropout(0.2), nn.Linear(64, 2))   # :57-62
        self.input_dim = input_dim
        self._wctx = None
        self._step = 0
        self._seed = 42

    # -- helpers -------------------------------------------------------------------------------------
    def flat_parameters(self) -> torch.Tensor:
        return torch.cat([p.reshape(-1) for p in self.net.parameters()])

    def _ctx(self, device):
        from .ops import Context
        if self._wctx is None or self._wctx.device != device:
            self._wctx = Context(self.input_dim, device=device)
        return self._wctx

    # -- reference API -------------------------------------------------------------------------------
    def forward(self, x):                                                                   # :64
        if x.is_cuda and not torch.is_grad_enabled():
            lead = x.shape[:-1]
            if self.training:
                # train-mode forward outside training_step draws torch's own dropout stream: leave it to torch
                return self.net(x)
            logits = self._ctx(x.device).mlp_logits(self.flat_parameters().detach().contiguous(),
                                                    x.reshape(-1, self.input_dim).contiguous())
            return logits.reshape(*lead, 2)                # the raw logits, as the reference's forward returns them
        return self.net(x)

    def training_step(self, batch, batch_idx):                                              # :66-71
        x, y = batch
        if x.is_cuda:
            w = self._ctx(x.device)
            mode = _lib.DROPOUT_PHILOX if self.training else _lib.DROPOUT_OFF
            loss = _FusedStepFn.apply(self.flat_parameters(), x, y, w, mode, self._seed, self._step)
            self._step += 1
        else:
            raise _lib.WxError("wxddp.WeatherClassifier.training_step needs CUDA tensors: there is no CPU path")
        self.log("train_loss", loss, sync_dist=True)
        return loss

    def validation_step(self, batch, batch_idx):                                            # :73-85
        x, y = batch
        if not x.is_cuda:
            raise _lib.WxError("wxddp.WeatherClassifier.validation_step needs CUDA tensors: there is no CPU path")
        out = self._ctx(x.device).mlp_eval(self.flat_parameters().detach().contiguous(), x.contiguous(), y.contiguous())
        n = x.shape[0]
        loss = (out[0] / n).float()
        acc = (out[1] / n).float()
        self.log("val_loss", loss, sync_dist=True, prog_bar=True)
        self.log("val_acc", acc, sync_dist=True, prog_bar=True)
        return loss

    def configure_optimizers(self):                                                         # :87-88
        return torch.optim.Adam(self.parameters(), lr=0.01)

    @classmethod
    def load_from_checkpoint(cls, checkpoint_path, map_location=None, **kwargs):
        """What score.py calls (dags/azure_auto_deploy.py:93): builds the module and loads `state_dict`."""
        if HAVE_LIGHTNING:                                                                  # pragma: no cover
            return super().load_from_checkpoint(checkpoint_path, map_location=map_location, **kwargs)
        ckpt = torch.load(checkpoint_path, map_location=map_location or "cpu", weights_only=False)
        hp = dict(ckpt.get("hyper_parameters", {}))
        hp.update(kwargs)
        model = cls(**hp)
        model.load_state_dict(ckpt["state_dict"])
        return model

    def load_flat(self, flat: torch.Tensor) -> None:
        o = 0
        with torch.no_grad():
            for p in self.net.parameters():
                k = p.numel()
                p.copy_(flat[o:o + k].view_as(p))
                o += k


__all__ = ["WeatherDataset", "WeatherClassifier", "HAVE_LIGHTNING", "flat_from_state_dict"]
