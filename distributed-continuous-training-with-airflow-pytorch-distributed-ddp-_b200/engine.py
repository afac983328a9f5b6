"""StepEngine: the user-facing training-step object (one per process / GPU).

Holds the flat parameter vector, Adam moments and step counter on the device and drives the CUDA step
kernels.  Mirrors what Lightning's automatic optimisation does around the reference's
`training_step` (jobs/train_lightning_ddp.py:66-71) and `configure_optimizers` (:87-88): forward, loss,
backward, gradient mean over ranks (DDP, :136), Adam, and the rank-mean `train_loss` (:70).
"""
from __future__ import annotations

import torch

from ._lib import DROPOUT_MASK, DROPOUT_OFF, DROPOUT_PHILOX, IMPUTE_MEAN, IMPUTE_NONE, WxError
from .ops import Context, n_params


def init_params_like_reference(n_features: int) -> torch.Tensor:
    """Flat fp32 parameters drawn exactly as `WeatherClassifier(input_dim)` draws them (:57-62): two
    nn.Linear default inits from the CURRENT global torch generator, in module order."""
    from torch import nn

    l1 = nn.Linear(n_features, 64)
    l2 = nn.Linear(64, 2)
    return torch.cat([l1.weight.detach().reshape(-1), l1.bias.detach(), l2.weight.detach().reshape(-1),
                      l2.bias.detach()]).clone()


class StepEngine:
    def __init__(self, n_features: int, device=0, world: int = 1, rank: int = 0, process_group=None,
                 params: torch.Tensor | None = None, seed: int = 42, lr: float = 0.01, dropout_p: float = 0.2,
                 use_tc: bool = True, optimizer: str = "adam"):
        self.ctx = Context(n_features, device=device, dropout_p=dropout_p, lr=lr, optimizer=optimizer)
        self.device = self.ctx.device
        self.F = n_features
        self.P = n_params(n_features)
        self.world, self.rank = world, rank
        self.use_tc = use_tc                # prefer the tcgen05 step kernel where its preconditions hold
        self.dist = process_group           # the torch.distributed module (NCCL) when world > 1
        if world > 1 and process_group is None:
            raise WxError("world > 1 needs an initialised torch.distributed (NCCL) handle")
        if params is None:
            params = init_params_like_reference(n_features)
        self.params = params.detach().to(self.device, torch.float32).contiguous().clone()
        self.m = torch.zeros(self.P, device=self.device)
        self.v = torch.zeros(self.P, device=self.device)
        self.state = torch.zeros(2, dtype=torch.int64, device=self.device)   # [steps done, reserved]
        self.loss = torch.zeros(1, device=self.device)
        self.bucket = torch.zeros(self.P + 1, device=self.device)            # grads + loss scalar (DDP bucket)
        self.seed = seed
        self.steps_done = 0
        self.stats = None
        self._stage = None
        self._loss_host = torch.zeros(1).pin_memory() if torch.cuda.is_available() else torch.zeros(1)
        self.peers = False
        if world > 1:
            self.dist.broadcast(self.params, src=0)                          # C1: DDP's initial parameter broadcast

    def connect_peers(self) -> None:
        """Exchange CUDA IPC handles of the per-rank inboxes so the fused tensor-core steps can all-reduce the [P+1]
        bucket inside the kernel (NVLink peer stores) instead of mlp_grad -> NCCL -> adam (one launch per step).
        All ranks must be GPUs of one node.  The barrier at the end is part of the protocol: no rank may step before
        every rank has connected."""
        if self.world == 1:
            return
        mine = self.ctx.peer_export()
        handles = [None] * self.world
        self.dist.all_gather_object(handles, mine)
        self.ctx.peer_connect(self.rank, self.world, handles)
        self.dist.barrier()
        self.peers = True

    def check(self) -> None:
        """Raise if a fused step kernel recorded a protocol error since the context was created (20: a peer did not
        arrive in the in-kernel exchange -- that step was NOT applied on this rank).  Synchronises the device."""
        st = self.ctx.tc_status()
        if st != 0:
            what = "a peer did not arrive in the in-kernel gradient exchange; the step was not applied" if st == 20 \
                else "a bounded barrier wait inside the step kernel expired"
            raise WxError(f"step kernel status {st}: {what}")

    def steps_small(self, x, y, rows, n_steps: int, batch: int, losses: torch.Tensor, dropout: int = DROPOUT_PHILOX) -> None:
        """`n_steps` optimizer steps of `batch` sampler-ordered rows each in ONE launch (the reference's batch-4 regime,
        train_lightning_ddp.py:122): parameters and Adam moments stay in shared memory between the steps; with connected
        peers every step's [P+1] sums cross NVLink inside the kernel (DDP, :136)."""
        self.ctx.train_steps_small(self.params, self.m, self.v, self.state, x, y, losses, n_steps=n_steps, batch=batch,
                                   row_idx=rows, dropout=dropout, seed=self._dropout_seed())
        self.steps_done += n_steps

    # ---- statistics ------------------------------------------------------------------------------
    def set_stats(self, stats: torch.Tensor) -> None:
        self.stats = stats

    def fit_stats(self, cols) -> torch.Tensor:
        """preprocess.py:32-36 over this rank's shard, Chan-merged across ranks."""
        st = self.ctx.col_stats(cols)
        if self.world > 1:
            parts = [torch.empty_like(st) for _ in range(self.world)]
            self.dist.all_gather(parts, st)
            st = self.ctx.stats_merge(torch.stack(parts).contiguous())
        self.stats = st
        return st

    # ---- steps -----------------------------------------------------------------------------------
    def _dropout_seed(self) -> int:
        # perf-mode streams are decorrelated across ranks; parity runs pass explicit masks instead
        return (self.seed + 0x9E3779B97F4A7C15 * self.rank) & 0xFFFFFFFFFFFFFFFF

    def step(self, x, y, *, row_idx=None, row0: int = 0, batch: int | None = None, dropout: int = DROPOUT_PHILOX,
             keep_mask=None) -> torch.Tensor:
        """One optimizer step on standardised row-major features.  Returns the device scalar `train_loss`."""
        if self.world == 1:
            # both are CUDA kernels of libwxddp.so: the tcgen05 kernel when the batch is contiguous and TMA-aligned,
            # the FFMA kernel otherwise (sampler indices, F > 29, odd alignment)
            big = (batch if batch is not None else x.shape[0] - row0) >= 128 if row_idx is None else False
            mask_ok = keep_mask is None or keep_mask.data_ptr() % 16 == 0
            if self.use_tc and big and mask_ok and self.ctx.tc_ok_rowmajor(x, y, row0, row_idx):
                self.ctx.tc_train_step(self.params, self.m, self.v, self.state, x, y, self.loss, row0=row0, batch=batch,
                                       dropout=dropout, keep_mask=keep_mask, seed=self._dropout_seed())
            else:
                self.ctx.train_step(self.params, self.m, self.v, self.state, x, y, self.loss, row_idx=row_idx, row0=row0,
                                    batch=batch, dropout=dropout, keep_mask=keep_mask, seed=self._dropout_seed())
            self.steps_done += 1
            return self.loss
        b = batch if batch is not None else (row_idx.numel() if row_idx is not None else x.shape[0] - row0)
        mask_ok = keep_mask is None or keep_mask.data_ptr() % 16 == 0
        if self.peers and self.use_tc and mask_ok and self.ctx.tc_ok_rowmajor(x, y, row0, row_idx):
            # fused: forward + backward + cross-rank sum over NVLink peer memory + Adam in ONE launch
            self.ctx.tc_train_step(self.params, self.m, self.v, self.state, x, y, self.loss, row0=row0, batch=b,
                                   dropout=dropout, keep_mask=keep_mask, seed=self._dropout_seed())
            self.steps_done += 1
            return self.loss
        if self.use_tc and b >= 128 and mask_ok and self.ctx.tc_ok_rowmajor(x, y, row0, row_idx):
            self.ctx.tc_mlp_grad(self.params, x, y, row0=row0, batch=b, dropout=dropout, keep_mask=keep_mask,
                                 seed=self._dropout_seed(), step=self.steps_done, scale=1.0 / (self.world * b),
                                 out=self.bucket)
        else:
            self.ctx.mlp_grad(self.params, x, y, row_idx=row_idx, row0=row0, batch=b, dropout=dropout,
                              keep_mask=keep_mask, seed=self._dropout_seed(), step=self.steps_done,
                              scale=1.0 / (self.world * b), out=self.bucket)
        self.dist.all_reduce(self.bucket)                                    # C2 + C3 in one bucket
        self.steps_done += 1
        self.ctx.adam_step(self.params, self.bucket, self.m, self.v, self.steps_done, n=self.P)
        self.state[:1] += 1               # the device counter drives the fused kernels' Adam / dropout / exchange tags
        return self.bucket[self.P:self.P + 1]

    def step_raw(self, cols, y, *, row0: int = 0, batch: int | None = None, impute: int = IMPUTE_MEAN,
                 dropout: int = DROPOUT_PHILOX, keep_mask=None) -> torch.Tensor:
        """One optimizer step straight from raw columns (standardisation fused into the step kernel)."""
        if self.stats is None:
            raise WxError("call fit_stats()/set_stats() before step_raw()")
        if not hasattr(cols, "cp"):
            # bind once per table: the same list object comes back every step of an epoch
            cached = getattr(self, "_bound", None)
            key = (len(cols), cols[0].data_ptr(), cols[-1].data_ptr())
            if cached is None or cached[0] is not cols or cached[2] != key:
                self._bound = (cols, self.ctx.bind_columns(cols), key)
            cols = self._bound[1]
        if self.world == 1:
            mask_ok = keep_mask is None or keep_mask.data_ptr() % 16 == 0
            fn = self.ctx.train_step_raw
            if self.use_tc and mask_ok and self.ctx.tc_ok_raw(cols, y, row0):
                fn = self.ctx.tc_train_step_raw
            fn(self.params, self.m, self.v, self.state, cols, self.stats, y, self.loss, row0=row0, batch=batch,
               impute=impute, dropout=dropout, keep_mask=keep_mask, seed=self._dropout_seed())
            self.steps_done += 1
            return self.loss
        n = cols.n
        b = batch if batch is not None else n - row0
        mask_ok = keep_mask is None or keep_mask.data_ptr() % 16 == 0
        if self.peers and self.use_tc and mask_ok and self.ctx.tc_ok_raw(cols, y, row0):
            self.ctx.tc_train_step_raw(self.params, self.m, self.v, self.state, cols, self.stats, y, self.loss, row0=row0,
                                       batch=b, impute=impute, dropout=dropout, keep_mask=keep_mask,
                                       seed=self._dropout_seed())
            self.steps_done += 1
            return self.loss
        feats, _ = self.ctx.normalize_pack([c[row0:row0 + b] for c in cols], self.stats, impute=impute,
                                           want_nullmask=False)
        return self.step(feats, y[row0:row0 + b].contiguous(), dropout=dropout, keep_mask=keep_mask)

    def step_host(self, cols_host: torch.Tensor, y_host: torch.Tensor) -> float:
        """End-to-end step from HOST buffers: H2D of the raw [F,B] fp32 columns and int64 labels (pinned memory
        recommended), the fused step, and a D2H read of the loss.  Returns the loss as a Python float."""
        f, b = cols_host.shape
        if self._stage is None or self._stage[0].shape[1] != b:
            dc = torch.empty(f, b, dtype=torch.float32, device=self.device)
            self._stage = (dc, torch.empty(b, dtype=torch.int64, device=self.device),
                           self.ctx.bind_columns([dc[j] for j in range(f)]))
        dc, dy, bound = self._stage
        dc.copy_(cols_host, non_blocking=True)
        dy.copy_(y_host, non_blocking=True)
        loss = self.step_raw(bound, dy, row0=0, batch=b)
        self._loss_host.copy_(loss.reshape(1), non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return float(self._loss_host.item())

    # ---- pipelined host path --------------------------------------------------------------------------
    def submit_host(self, cols_host: torch.Tensor, y_host: torch.Tensor):
        """Pipelined form of `step_host`: the H2D copy of THIS batch runs on a copy stream while the previous step's
        kernel runs on the compute stream (two staging slots, event-ordered), and the loss comes back one call late --
        the return value is the loss of the previously submitted step (None on the first call); `drain()` returns the
        last one.  y_host may be int64 (as WeatherDataset holds it), int32 (as Spark wrote `label_encoded`,
        preprocess.py:24) or uint8 (the same 0/1 codes, one byte per row): narrow labels cross PCIe as they are and are widened on the device, which is the cast of
        train_lightning_ddp.py:46 moved behind the copy.  Same arithmetic and step order as `step_host`; host buffers should be pinned and must stay untouched
        until the next call returns."""
        f, b = cols_host.shape
        if getattr(self, "_pipe", None) is None or self._pipe["b"] != b:
            self._pipe = {
                "b": b, "k": 0, "copy": torch.cuda.Stream(device=self.device),
                "cols": [torch.empty(f, b, dtype=torch.float32, device=self.device) for _ in range(2)],
                "y": [torch.empty(b, dtype=torch.int64, device=self.device) for _ in range(2)],
                "y32": [torch.empty(b, dtype=torch.int32, device=self.device) for _ in range(2)],
                "y8": [torch.empty(b, dtype=torch.uint8, device=self.device) for _ in range(2)],
                "ready": [torch.cuda.Event() for _ in range(2)], "free": [torch.cuda.Event() for _ in range(2)],
                "done": [torch.cuda.Event() for _ in range(2)],
                "loss": [torch.zeros(1).pin_memory() for _ in range(2)], "pending": None,
            }
            self._pipe["bound"] = [self.ctx.bind_columns([c[j] for j in range(f)]) for c in self._pipe["cols"]]
        p = self._pipe
        slot = p["k"] & 1
        compute = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(p["copy"]):
            if p["k"] >= 2:
                p["copy"].wait_event(p["free"][slot])        # the step that last read this slot has finished
            p["cols"][slot].copy_(cols_host, non_blocking=True)
            if y_host.dtype == torch.int32:
                p["y32"][slot].copy_(y_host, non_blocking=True)
                p["y"][slot].copy_(p["y32"][slot])              # int32 -> int64 on the device (:46)
            elif y_host.dtype == torch.uint8:                  # the 0/1 codes of preprocess.py:23-25 in one byte each
                p["y8"][slot].copy_(y_host, non_blocking=True)
                p["y"][slot].copy_(p["y8"][slot])
            else:
                p["y"][slot].copy_(y_host, non_blocking=True)
            p["ready"][slot].record(p["copy"])
        compute.wait_event(p["ready"][slot])
        loss = self.step_raw(p["bound"][slot], p["y"][slot], row0=0, batch=b)
        p["free"][slot].record(compute)
        p["loss"][slot].copy_(loss.reshape(1), non_blocking=True)
        p["done"][slot].record(compute)
        prev, p["pending"] = p["pending"], slot
        p["k"] += 1
        if prev is None:
            return None
        p["done"][prev].synchronize()
        return float(p["loss"][prev].item())

    def drain(self):
        """Loss of the last `submit_host` call (waits for it); None if nothing is pending."""
        p = getattr(self, "_pipe", None)
        if p is None or p["pending"] is None:
            return None
        slot, p["pending"] = p["pending"], None
        p["done"][slot].synchronize()
        return float(p["loss"][slot].item())

    def evaluate(self, x, y, *, row_idx=None, row0: int = 0, batch: int | None = None):
        """validation pass (:73-85) over this rank's rows -> (val_loss, val_acc) as rank-weighted means (C4)."""
        out = self.ctx.mlp_eval(self.params, x, y, row_idx=row_idx, row0=row0, batch=batch)
        b = batch if batch is not None else (row_idx.numel() if row_idx is not None else x.shape[0] - row0)
        tot = torch.cat([out, torch.tensor([float(b)], dtype=torch.float64, device=self.device)])
        if self.world > 1:
            self.dist.all_reduce(tot)
        tot = tot.cpu()
        self.check()
        return float(tot[0] / tot[2]), float(tot[1] / tot[2])
