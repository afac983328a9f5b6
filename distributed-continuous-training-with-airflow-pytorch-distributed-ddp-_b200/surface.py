"""Operator surface kept from the reference: checkpoint files and metric names.

* Checkpoints: `ModelCheckpoint(dirpath, filename="weather-best-{epoch:02d}-{val_loss:.2f}", save_top_k=1,
  monitor="val_loss", mode="min", save_last=True)` (jobs/train_lightning_ddp.py:103-110) -> files
  `weather-best-epoch=NN-val_loss=X.XX.ckpt` (top-1) and `last.ckpt`, each a `torch.save` dict in
  pytorch-lightning 2.1.0's schema, loadable by the deploy DAG's `score.py`
  (`WeatherClassifier.load_from_checkpoint(path, input_dim=5)`, dags/azure_auto_deploy.py:93).
* Metrics: MLflow experiment `weather_forecasting`, metrics `train_loss` (every `log_every_n_steps`=5 steps),
  `val_loss`, `val_acc`, `epoch`; param `input_dim`; artifact dir `best_checkpoints/` (:92-96, :157-161).
  When `mlflow` is not importable the same names go to a JSON-lines file.
"""
from __future__ import annotations

import json
import os
import time
from collections import OrderedDict

import torch

PL_VERSION = "2.1.0"                       # Dockerfile.pytorch:14 pins pytorch-lightning==2.1.0
EXPERIMENT = "weather_forecasting"         # train_lightning_ddp.py:93
CKPT_TEMPLATE = "weather-best-epoch={epoch:02d}-val_loss={val_loss:.2f}.ckpt"


def state_dict_from_flat(flat: torch.Tensor, n_features: int) -> "OrderedDict[str, torch.Tensor]":
    """Flat fp32 vector -> the reference module's state_dict (keys/shapes of train_lightning_ddp.py:57-62)."""
    flat = flat.detach().to("cpu", torch.float32)
    f = n_features
    o = 0
    sd = OrderedDict()
    for key, shape in (("net.0.weight", (64, f)), ("net.0.bias", (64,)), ("net.3.weight", (2, 64)), ("net.3.bias", (2,))):
        k = 1
        for d in shape:
            k *= d
        sd[key] = flat[o:o + k].reshape(shape).clone()
        o += k
    return sd


def flat_from_state_dict(sd) -> torch.Tensor:
    return torch.cat([sd[k].reshape(-1).to(torch.float32) for k in ("net.0.weight", "net.0.bias", "net.3.weight", "net.3.bias")])


def build_checkpoint(flat, m, v, n_features: int, epoch: int, global_step: int, lr: float = 0.01,
                     callbacks: dict | None = None) -> dict:
    """A dict with the keys pytorch-lightning 2.1.0 writes (SURVEY 5 'Checkpoint / resume')."""
    sd = state_dict_from_flat(flat, n_features)
    ms = state_dict_from_flat(m, n_features)
    vs = state_dict_from_flat(v, n_features)
    opt_state = {}
    for i, key in enumerate(sd):
        opt_state[i] = {"step": torch.tensor(float(global_step)), "exp_avg": ms[key], "exp_avg_sq": vs[key]}
    optimizer_states = [{
        "state": opt_state,
        "param_groups": [{"lr": lr, "betas": (0.9, 0.999), "eps": 1e-08, "weight_decay": 0, "amsgrad": False,
                          "maximize": False, "foreach": None, "capturable": False, "differentiable": False,
                          "fused": None, "params": [0, 1, 2, 3]}],
    }]
    return {
        "epoch": epoch,
        "global_step": global_step,
        "pytorch-lightning_version": PL_VERSION,
        "state_dict": sd,
        "loops": {"fit_loop": {"state_dict": {}, "epoch_progress": {"current": {"completed": epoch + 1}}}},
        "callbacks": callbacks or {},
        "optimizer_states": optimizer_states,
        "lr_schedulers": [],
        "hparams_name": "kwargs",
        "hyper_parameters": {"input_dim": n_features},
    }


class TopOneCheckpoint:
    """The reference's ModelCheckpoint settings (:103-110): keep the best-by-val_loss file and last.ckpt."""

    def __init__(self, dirpath: str):
        self.dirpath = dirpath
        self.best_model_path = ""
        self.best_model_score = None
        os.makedirs(dirpath, exist_ok=True)               # :99-100

    def _callback_state(self):
        return {"ModelCheckpoint{'monitor': 'val_loss', 'mode': 'min', 'every_n_train_steps': 0, 'every_n_epochs': 1, "
                "'train_time_interval': None}": {
                    "monitor": "val_loss", "best_model_score": None if self.best_model_score is None
                    else torch.tensor(self.best_model_score), "best_model_path": self.best_model_path,
                    "current_score": None, "dirpath": self.dirpath, "best_k_models": {}, "kth_best_model_path": "",
                    "kth_value": None, "last_model_path": os.path.join(self.dirpath, "last.ckpt")}}

    def on_validation_end(self, make_ckpt, epoch: int, val_loss: float) -> None:
        """make_ckpt(callbacks) -> checkpoint dict.  Saves top-1 (replacing the previous best) and last.ckpt."""
        improved = self.best_model_score is None or val_loss < self.best_model_score
        if improved and val_loss == val_loss:              # NaN never becomes "best"
            path = os.path.join(self.dirpath, CKPT_TEMPLATE.format(epoch=epoch, val_loss=val_loss))
            old = self.best_model_path
            self.best_model_score = float(val_loss)
            self.best_model_path = path
            torch.save(make_ckpt(self._callback_state()), path)
            if old and old != path and os.path.exists(old):
                os.remove(old)
        torch.save(make_ckpt(self._callback_state()), os.path.join(self.dirpath, "last.ckpt"))


class MetricLogger:
    """MLflow when importable (same experiment / metric / artifact names as the reference), else JSON lines."""

    def __init__(self, tracking_uri: str | None, fallback_dir: str, params: dict, enabled: bool = True):
        self.enabled = enabled
        self.mlflow = None
        self.run_id = None
        self.path = None
        if not enabled:
            return
        try:
            import mlflow  # noqa: F401
            self.mlflow = mlflow
        except Exception:
            self.mlflow = None
        if self.mlflow is not None:
            self.mlflow.set_tracking_uri(tracking_uri or "http://mlflow-server:5000")       # :94
            self.mlflow.set_experiment(EXPERIMENT)                                          # :93
            run = self.mlflow.start_run()
            self.run_id = run.info.run_id
            for k, v in params.items():
                self.mlflow.log_param(k, v)
        else:
            os.makedirs(fallback_dir, exist_ok=True)
            self.run_id = time.strftime("run-%Y%m%d-%H%M%S")
            self.path = os.path.join(fallback_dir, f"{EXPERIMENT}-{self.run_id}.jsonl")
            self._write({"type": "params", "experiment": EXPERIMENT, "params": params})

    def _write(self, rec: dict) -> None:
        with open(self.path, "a") as fh:
            fh.write(json.dumps(rec) + "\n")

    def log_metrics(self, metrics: dict, step: int) -> None:
        if not self.enabled:
            return
        if self.mlflow is not None:
            for k, v in metrics.items():
                self.mlflow.log_metric(k, float(v), step=step)
        else:
            self._write({"type": "metrics", "step": step, **{k: float(v) for k, v in metrics.items()}})

    def log_artifact(self, local_path: str, artifact_path: str) -> None:
        if not self.enabled:
            return
        if self.mlflow is not None:
            self.mlflow.log_artifact(local_path, artifact_path=artifact_path)               # :157-161
        else:
            self._write({"type": "artifact", "local_path": local_path, "artifact_path": artifact_path})

    def close(self) -> None:
        if self.enabled and self.mlflow is not None:
            self.mlflow.end_run()
