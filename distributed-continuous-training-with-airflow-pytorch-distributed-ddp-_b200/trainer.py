"""The fit loop behind the reference's CLI (jobs/train_lightning_ddp.py:90-164), driving the CUDA step kernels.

Observable behaviour kept from the reference: zero-argument entry, env `MASTER_ADDR / MASTER_PORT / NODE_RANK /
WORLD_SIZE / MLFLOW_TRACKING_URI`, seed 42 on every rank (:14), 80/20 `random_split` drawn before model init
(:117-119,:126), batch 4 with shuffle (:122), 10 epochs (:132), `train_loss` logged every 5 steps (:139),
`val_loss`/`val_acc` per epoch (:83-84), top-1 + last checkpoints (:103-110), rank-0 artifact upload (:146-164),
exit code 0 on success.  What changes is where the arithmetic runs: the table lives in HBM, sampler indices are
uploaded once per epoch, and steps run as CUDA kernels (`wx_train_steps_small` for the reference's tiny batches,
`wx_train_step` otherwise; NCCL all-reduce of the [P+1] bucket when WORLD_SIZE > 1).

Extra env (all optional, defaults = the reference's constants): WX_DATA_DIR (/workspace/data/processed),
WX_MODEL_DIR (/workspace/data/models), WX_BATCH_SIZE (4), WX_MAX_EPOCHS (10), WX_LOG_EVERY (5), WX_PARITY (0: Philox
dropout; 1: replay torch's CPU dropout stream and loader RNG draws so the loss curve tracks the reference's),
WX_MAX_STEPS, WX_METRICS_DIR, WX_DISABLE_MLFLOW, WX_INGEST (gpu: parquet -> Arrow buffers -> pinned chunks -> HBM -> cast/pack
kernel, the default; pandas: the reference's own WeatherDataset on the host).
"""
from __future__ import annotations

import math
import os
import sys

import numpy as np
import torch

from ._lib import DROPOUT_MASK, DROPOUT_PHILOX, WxError
from .engine import StepEngine, init_params_like_reference
from .surface import MetricLogger, TopOneCheckpoint, build_checkpoint

SEED = 42                      # pl.seed_everything(42), train_lightning_ddp.py:14
SMALL_BATCH_MAX = 512          # many steps per launch pay off up to here (one CTA to 64 rows, resident multi-CTA above)
SMALL_BATCH_DDP_MAX = 64       # with the in-kernel DDP exchange: the single-CTA kernel only
STEPS_PER_LAUNCH = 4096


# ---- index streams (host logic; mirrors torch.utils.data exactly) ------------------------------------------
def split_indices(n_rows: int):
    """`random_split(full, [int(0.8 n), rest])` with the global generator (:117-119)."""
    train_size = int(0.8 * n_rows)
    perm = torch.randperm(n_rows).tolist()
    return perm[:train_size], perm[train_size:]


def distributed_order(n: int, epoch: int, rank: int, world: int, shuffle: bool, seed: int = SEED) -> list[int]:
    """DistributedSampler (torch/utils/data/distributed.py): randperm(seed+epoch) | arange, pad by wrapping to a
    multiple of world, take indices[rank::world].  Lightning injects it with seed = PL_GLOBAL_SEED = 42."""
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        idx = torch.randperm(n, generator=g).tolist()
    else:
        idx = list(range(n))
    total = math.ceil(n / world) * world
    pad = total - len(idx)
    if pad > 0:
        idx += (idx * math.ceil(pad / len(idx)))[:pad]
    return idx[rank:total:world]


def _draw_loader_base_seed() -> None:
    # every torch DataLoader iterator draws one int64 from the global generator when created
    torch.empty((), dtype=torch.int64).random_()


def random_sampler_order(n: int) -> list[int]:
    """RandomSampler under DataLoader(shuffle=True): seed drawn from the global generator, then randperm."""
    s = int(torch.empty((), dtype=torch.int64).random_().item())
    g = torch.Generator()
    g.manual_seed(s)
    return torch.randperm(n, generator=g).tolist()


def _pack_mask(mask: torch.Tensor) -> torch.Tensor:
    bits = np.packbits(mask.to(torch.bool).numpy(), axis=1, bitorder="little").view(np.int64).reshape(-1)
    return torch.from_numpy(bits.copy())


# ---- the loop ----------------------------------------------------------------------------------------------
def fit(features: torch.Tensor, labels: torch.Tensor, *, device, rank: int = 0, world: int = 1, dist=None,
        batch_size: int = 4, max_epochs: int = 10, max_steps: int | None = None, log_every: int = 5,
        parity: bool = False, checkpoint: TopOneCheckpoint | None = None, logger: MetricLogger | None = None,
        sanity_check: bool = True, verbose: bool = True, engine_factory=None) -> dict:
    """Run the reference's training loop on the GPU.  features fp32 [N,F] / labels int64 [N] are what WeatherDataset
    holds -- host tensors (uploaded once) or tensors already resident on `device` (wxddp.ingest.load_processed)."""
    n, f = features.shape
    torch.manual_seed(SEED)
    tr_idx, va_idx = split_indices(n)                       # consumes the global generator BEFORE model init
    params0 = init_params_like_reference(f)
    make = engine_factory or StepEngine        # tests inject a CPU double to exercise the N>1 host logic over gloo
    eng = make(f, device=device, world=world, rank=rank, process_group=dist, params=params0, seed=SEED)
    if world > 1 and not parity and hasattr(eng, "connect_peers"):
        eng.connect_peers()             # DDP exchange inside the step kernels (all ranks are GPUs of this node)
    dev = eng.device
    x_d = features.to(dev).contiguous()
    y_d = labels.to(dev).contiguous()
    tr_idx_t = torch.tensor(tr_idx, dtype=torch.int64)
    va_rows = torch.tensor(va_idx, dtype=torch.int64)
    hist = {"train_loss": [], "val_loss": [], "val_acc": [], "logged": []}
    global_step = 0
    if parity and sanity_check:
        _draw_loader_base_seed()                            # Lightning's sanity check opens the val loader once
    val_pos = distributed_order(len(va_idx), 0, rank, world, shuffle=False) if world > 1 else list(range(len(va_idx)))
    val_rows_d = va_rows[val_pos].to(dev)
    done = False
    for epoch in range(max_epochs):
        if parity:
            _draw_loader_base_seed()                        # train loader iterator
        if world > 1:
            order = distributed_order(len(tr_idx), epoch, rank, world, shuffle=True)
        elif parity:
            order = random_sampler_order(len(tr_idx))
        else:
            g = torch.Generator()
            g.manual_seed(SEED + epoch)
            order = torch.randperm(len(tr_idx), generator=g).tolist()
        rows = tr_idx_t[order]
        n_steps = math.ceil(len(order) / batch_size)
        if max_steps is not None:
            n_steps = min(n_steps, max_steps - global_step)
        rows_d = rows.to(dev)
        full_steps = len(order) // batch_size
        epoch_losses = _run_epoch(eng, x_d, y_d, rows_d, len(order), batch_size, n_steps, full_steps, parity, world)
        for s, l in enumerate(epoch_losses):
            global_step += 1
            hist["train_loss"].append(l)
            if global_step % log_every == 0:
                hist["logged"].append((global_step, l))
                if logger is not None and rank == 0:
                    # Lightning 2.1 logs against `_batches_that_stepped`, which is bumped only after the batch's metrics
                    # were written: the 5th batch lands at step 4, the 10th at 9, ...
                    logger.log_metrics({"train_loss": l, "epoch": epoch}, step=global_step - 1)
        if max_steps is not None and global_step >= max_steps:
            done = True
        # validation epoch (:73-85): eval mode, rank-weighted means (C4)
        if parity:
            _draw_loader_base_seed()                        # val loader iterator
        val_loss, val_acc = eng.evaluate(x_d, y_d, row_idx=val_rows_d)
        hist["val_loss"].append(val_loss)
        hist["val_acc"].append(val_acc)
        if verbose and rank == 0:
            print(f"Epoch {epoch}: train_loss={hist['train_loss'][-1]:.4f} val_loss={val_loss:.4f} val_acc={val_acc:.4f}",
                  flush=True)
        if logger is not None and rank == 0:
            logger.log_metrics({"val_loss": val_loss, "val_acc": val_acc, "epoch": epoch}, step=global_step - 1)
        if checkpoint is not None and rank == 0:
            p, m, v = eng.params.cpu(), eng.m.cpu(), eng.v.cpu()
            checkpoint.on_validation_end(lambda cb: build_checkpoint(p, m, v, f, epoch, global_step, callbacks=cb),
                                         epoch, val_loss)
        if world > 1:
            dist.barrier()                                  # C5: ranks leave the epoch together
        if done:
            break
    hist["params"] = eng.params.cpu()
    hist["steps"] = global_step
    hist["engine"] = eng
    return hist


def _run_epoch(eng: StepEngine, x_d, y_d, rows_d, n_rows: int, batch: int, n_steps: int, full_steps: int,
               parity: bool, world: int) -> list[float]:
    """All optimizer steps of one epoch.  Returns the per-step `train_loss` (rank mean when world > 1)."""
    dev = eng.device
    losses: list[float] = []
    small_max = SMALL_BATCH_MAX if world == 1 else SMALL_BATCH_DDP_MAX
    if (world == 1 or getattr(eng, "peers", False)) and batch <= small_max and not parity:
        # the reference's regime: thousands of tiny steps -> many steps per launch, one resident CTA (per rank: with
        # WORLD_SIZE > 1 every step's gradient bucket crosses NVLink inside that kernel)
        s = 0
        n_full = min(n_steps, full_steps)
        buf = torch.empty(min(STEPS_PER_LAUNCH, max(n_full, 1)), device=dev)
        while s < n_full:
            k = min(STEPS_PER_LAUNCH, n_full - s)
            eng.steps_small(x_d, y_d, rows_d[s * batch:(s + k) * batch].contiguous(), k, batch, buf)
            losses += buf[:k].cpu().tolist()
            s += k
        if world > 1:
            eng.check()
        if n_steps > full_steps:                            # ragged last batch (drop_last=False)
            l = eng.step(x_d, y_d, row_idx=rows_d[full_steps * batch:n_rows].contiguous())
            losses.append(float(l.item()))
        return losses
    loss_buf = []
    for s in range(n_steps):
        idx = rows_d[s * batch:min((s + 1) * batch, n_rows)].contiguous()
        b = idx.numel()
        if parity:
            mask = torch.empty(b, 64).bernoulli_(0.8)       # the draw nn.Dropout(0.2) makes on the CPU generator
            l = eng.step(x_d, y_d, row_idx=idx, dropout=DROPOUT_MASK, keep_mask=_pack_mask(mask).to(dev))
        else:
            l = eng.step(x_d, y_d, row_idx=idx, dropout=DROPOUT_PHILOX)
        loss_buf.append(l.clone())
        if len(loss_buf) >= 512 or s == n_steps - 1:
            losses += torch.cat([t.reshape(1) for t in loss_buf]).cpu().tolist()
            loss_buf = []
    return losses


# ---- CLI ---------------------------------------------------------------------------------------------------
def main() -> int:
    """Drop-in for `python3 /workspace/jobs/train_lightning_ddp.py` (dags/2_pytorch_training.py:55,60)."""
    from .module import WeatherDataset

    data_dir = os.environ.get("WX_DATA_DIR", "/workspace/data/processed")                  # :114
    save_path = os.environ.get("WX_MODEL_DIR", "/workspace/data/models")                   # :99
    batch = int(os.environ.get("WX_BATCH_SIZE", "4"))                                      # :122
    epochs = int(os.environ.get("WX_MAX_EPOCHS", "10"))                                    # :132
    log_every = int(os.environ.get("WX_LOG_EVERY", "5"))                                   # :139
    max_steps = os.environ.get("WX_MAX_STEPS")
    parity = os.environ.get("WX_PARITY", "0") == "1"
    world = int(os.environ.get("WORLD_SIZE", 1))                                           # :129
    rank = int(os.environ.get("NODE_RANK", os.environ.get("RANK", 0)))
    if not torch.cuda.is_available():
        raise WxError("train_lightning_ddp.py (wxddp) needs a B200: there is no CPU path")
    local = int(os.environ.get("LOCAL_RANK", rank % max(1, torch.cuda.device_count())))
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)    # staggered starts are fine
    torch.manual_seed(SEED)
    logger = None
    if rank == 0:
        logger = MetricLogger(os.environ.get("MLFLOW_TRACKING_URI", "http://mlflow-server:5000"),
                              os.environ.get("WX_METRICS_DIR", os.path.join(os.path.dirname(save_path.rstrip("/")), "mlflow_fallback")),
                              params={}, enabled=os.environ.get("WX_DISABLE_MLFLOW", "0") != "1")
    ckpt = TopOneCheckpoint(save_path) if rank == 0 else None
    if os.environ.get("WX_INGEST", "gpu") == "pandas":
        full = WeatherDataset(data_dir)                    # the reference's own loader (pandas DataFrame on the host)
        features, labels = full.features, full.labels
    else:
        # same tensors, built on the GPU: Arrow buffers -> pinned chunks -> HBM -> cast + pack kernel (wxddp/ingest.py)
        from .ingest import load_processed
        print(f"Loading data from: {os.path.join(data_dir, 'data.parquet')}")
        try:
            got = load_processed(data_dir, device=device)
        except (FileNotFoundError, ValueError):
            raise
        except Exception as e:                                                              # :32-33
            raise RuntimeError(f"Failed to read Parquet file: {e}")
        features, labels = got["features"], got["labels"]
        print(f"✓ Loaded {features.shape[0]} rows.")
        print(f"✓ Training with {len(got['feature_cols'])} features: {got['feature_cols']}")
    input_dim = features.shape[1]
    if logger is not None and logger.enabled:
        if logger.mlflow is not None:
            logger.mlflow.log_param("input_dim", input_dim)
        else:
            logger._write({"type": "params", "params": {"input_dim": input_dim}})
    hist = fit(features, labels, device=device, rank=rank, world=world, dist=dist, batch_size=batch,
               max_epochs=epochs, max_steps=int(max_steps) if max_steps else None, log_every=log_every, parity=parity,
               checkpoint=ckpt, logger=logger)
    if rank == 0:                                                                           # :146-164
        best_path = ckpt.best_model_path
        if not os.path.exists(best_path):
            print(f"⚠ Best model not found at {best_path}. Checking for 'last.ckpt'...")
            best_path = os.path.join(save_path, "last.ckpt")
        print(f"Training finished. Saving model path: {best_path}")
        if os.path.exists(best_path):
            logger.log_artifact(best_path, "best_checkpoints")
            print("✓ Model uploaded to MLflow!" if logger.mlflow is not None else "✓ Model recorded (MLflow not installed; JSONL log)")
        else:
            print("✗ CRITICAL: No model file found to upload!")
        logger.close()
    if world > 1:
        dist.destroy_process_group()
    del hist
    return 0


if __name__ == "__main__":
    sys.exit(main())
