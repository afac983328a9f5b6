"""Plain-torch restatement of `/root/reference/jobs/train_lightning_ddp.py` (oracle, test infrastructure only).

Pinned parts (checked bit-for-bit against vectors produced by the reference's own
classes, `tests/golden/make_golden.py` -> `tests/test_oracle.py`): the network
(`:57-62`), forward (`:64`), training/validation step arithmetic (`:66-85`), the Adam
hyper-parameters (`:88`), the dataset casts (`:45-46`), the seed / split / init order
(`:14`, `:117-119`, `:126`).  Unpinned parts (pytorch-lightning 2.1.0 is not installable
here): the fit loop around them -- restated from Lightning's documented behaviour and
anchored on the call at `:131-143` (DistributedSampler injection with seed 42, metric
mean over ranks, validation once per epoch, checkpoint on `val_loss`).
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

HIDDEN = 64          # train_lightning_ddp.py:58
CLASSES = 2          # :61
P_DROP = 0.2         # :60
LR = 0.01            # :88
BETA1, BETA2, EPS = 0.9, 0.999, 1e-8      # torch.optim.Adam defaults used by :88
SEED = 42            # :14
BATCH = 4            # :122
MAX_EPOCHS = 10      # :132
LOG_EVERY = 5        # :139


def n_params(f: int) -> int:
    return HIDDEN * f + HIDDEN + CLASSES * HIDDEN + CLASSES


class WeatherNet(nn.Module):
    """Same parameter names as the reference (`net.0.*`, `net.3.*`) -- :57-62."""

    def __init__(self, input_dim: int):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(input_dim, HIDDEN), nn.ReLU(), nn.Dropout(P_DROP),
                                 nn.Linear(HIDDEN, CLASSES))

    def forward(self, x):                     # :64
        return self.net(x)


def flatten(model: nn.Module) -> torch.Tensor:
    """Flat fp32 parameter vector in state_dict order: W1[64,F], b1[64], W2[2,64], b2[2]."""
    return torch.cat([p.detach().reshape(-1) for p in model.parameters()]).clone()


def unflatten_into(model: nn.Module, flat: torch.Tensor) -> None:
    o = 0
    with torch.no_grad():
        for p in model.parameters():
            k = p.numel()
            p.copy_(flat[o:o + k].view_as(p))
            o += k


def split_flat(flat: torch.Tensor, f: int):
    o = 0
    w1 = flat[o:o + HIDDEN * f].view(HIDDEN, f); o += HIDDEN * f
    b1 = flat[o:o + HIDDEN]; o += HIDDEN
    w2 = flat[o:o + CLASSES * HIDDEN].view(CLASSES, HIDDEN); o += CLASSES * HIDDEN
    b2 = flat[o:o + CLASSES]
    return w1, b1, w2, b2


def draw_dropout_mask(b: int) -> torch.Tensor:
    """The mask `nn.Dropout(0.2)` (:60) would draw next from the global CPU generator.

    ATen's CPU dropout is `noise = empty_like(h).bernoulli_(1-p); noise.div_(1-p); h*noise`,
    so drawing the same Bernoulli tensor from the same RNG state replays it bit-for-bit
    (asserted against the reference module in tests/golden/make_golden.py).
    """
    return torch.empty(b, HIDDEN).bernoulli_(1.0 - P_DROP)


def pack_mask(mask: torch.Tensor) -> np.ndarray:
    """[B,64] 0/1 -> uint64 [B], bit j = keep(hidden unit j).  8 B/sample, the kernel's parity input."""
    m = mask.to(torch.bool).numpy()
    return np.packbits(m, axis=1, bitorder="little").view(np.uint64).reshape(-1)


def forward_with_mask(flat: torch.Tensor, f: int, x: torch.Tensor, mask: torch.Tensor | None):
    """:57-64 written out: z1 = x W1^T + b1; h = relu(z1); a = h * mask / 0.8; logits = a W2^T + b2."""
    w1, b1, w2, b2 = split_flat(flat, f)
    h = torch.relu(F.linear(x, w1, b1))
    if mask is not None:
        h = h * (mask / (1.0 - P_DROP))
    return F.linear(h, w2, b2)


def loss_and_grads(flat: torch.Tensor, f: int, x: torch.Tensor, y: torch.Tensor,
                   mask: torch.Tensor | None):
    """training_step (:66-71) + autograd backward: mean cross-entropy and d loss / d flat."""
    p = flat.clone().requires_grad_(True)
    loss = F.cross_entropy(forward_with_mask(p, f, x, mask), y)          # :69
    (g,) = torch.autograd.grad(loss, p)
    return loss.detach(), g


def eval_batch(flat: torch.Tensor, f: int, x: torch.Tensor, y: torch.Tensor):
    """validation_step (:73-85): eval-mode forward, mean CE, argmax accuracy.  Returns (loss, n_correct)."""
    logits = forward_with_mask(flat, f, x, None)
    loss = F.cross_entropy(logits, y)
    correct = int((torch.argmax(logits, dim=1) == y).sum())
    return loss, correct


def adam_update(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, step: int,
                lr: float = LR, beta1: float = BETA1, beta2: float = BETA2, eps: float = EPS) -> None:
    """In-place single-tensor Adam exactly as torch runs it on CPU for :88.

    Mirrors torch/optim/adam.py `_single_tensor_adam` (installed 2.11, same arithmetic as the
    pinned 2.1.0): lerp_ for m, mul_/addcmul_ for v, Python-double bias corrections,
    `denom = sqrt(v)/sqrt(bc2) + eps`, `p.addcdiv_(m, denom, value=-lr/bc1)`.  `step` is 1-based.
    """
    m.lerp_(g, 1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = (v.sqrt() / (bc2 ** 0.5)).add_(eps)
    p.addcdiv_(m, denom, value=-(lr / bc1))


def sgd_update(p: torch.Tensor, g: torch.Tensor, lr: float = LR) -> None:
    """In-place `torch.optim.SGD(lr)` step without momentum: p.add_(g, alpha=-lr).  The reference only uses Adam (:88);
    this is the oracle of the WX_OPT_SGD extension."""
    p.add_(g, alpha=-lr)


def seeded_setup(n_rows: int, f: int, seed: int = SEED):
    """The RNG-order-sensitive prologue of main(): seed (:14) -> random_split (:117-119) -> model init (:126).

    Returns (train_indices, val_indices, model).  `randperm` is drawn from the global generator BEFORE
    nn.Linear's kaiming-uniform init, so the initial weights depend on having made that draw.
    """
    torch.manual_seed(seed)
    train_size = int(0.8 * n_rows)
    perm = torch.randperm(n_rows).tolist()          # what random_split does with the default generator
    model = WeatherNet(f)
    return perm[:train_size], perm[train_size:], model


def load_processed(data_path: str):
    """WeatherDataset.__init__ (:17-46): parquet DIRECTORY -> (features fp32 [N,F], labels int64 [N], cols)."""
    import pandas as pd

    parquet_path = os.path.join(data_path, "data.parquet")                      # :19
    if not os.path.exists(parquet_path):                                        # :22-26
        raise FileNotFoundError(f"CRITICAL ERROR: Data not found at {parquet_path}.")
    try:
        df = pd.read_parquet(parquet_path)                                      # :31
    except Exception as e:                                                      # :32-33
        raise RuntimeError(f"Failed to read Parquet file: {e}")
    cols = [c for c in df.columns if c.endswith("_norm")]                      # :37
    if not cols:                                                                # :39-40
        raise ValueError("CRITICAL ERROR: No columns ending with '_norm' found. Check Spark logic.")
    feats = torch.FloatTensor(df[cols].values)                                  # :45
    labels = torch.LongTensor(df["label_encoded"].values)                       # :46
    return feats, labels, cols


def epoch_indices(n_train: int, epoch: int, rank: int, world: int, seed: int = SEED) -> list[int]:
    """Positions into the train split that `rank` visits in `epoch` under Lightning's injected
    DistributedSampler(shuffle=True, seed=PL_GLOBAL_SEED=42) -- torch/utils/data/distributed.py:107-141:
    randperm(n, seed+epoch) -> pad by wrapping to a multiple of W -> indices[rank::W]."""
    g = torch.Generator()
    g.manual_seed(seed + epoch)
    idx = torch.randperm(n_train, generator=g).tolist()
    total = math.ceil(n_train / world) * world
    pad = total - len(idx)
    if pad > 0:
        idx += (idx * math.ceil(pad / len(idx)))[:pad]
    return idx[rank:total:world]


def val_indices_for_rank(n_val: int, rank: int, world: int) -> list[int]:
    """DistributedSampler(shuffle=False) on the val split: arange, pad by wrapping, stride."""
    idx = list(range(n_val))
    total = math.ceil(n_val / world) * world
    pad = total - len(idx)
    if pad > 0:
        idx += (idx * math.ceil(pad / len(idx)))[:pad]
    return idx[rank:total:world]


def fit(features: torch.Tensor, labels: torch.Tensor, *, batch_size: int = BATCH,
        max_epochs: int = MAX_EPOCHS, max_steps: int | None = None, rank: int = 0, world: int = 1,
        use_ddp: bool = False, record_masks: bool = False, per_sample_loader: bool = False,
        seed: int = SEED, sanity_check: bool = True) -> dict:
    """Restatement of main() :114-143 with Lightning's loop written out.

    world == 1: DataLoader(shuffle=True) semantics (RandomSampler seeded from the global generator).
    world > 1: DistributedSampler semantics + gradient mean over ranks (DDP, :136) and a per-step
    all-reduce of the scalar loss (`self.log(..., sync_dist=True)`, :70).  `use_ddp` needs an
    initialised gloo process group.  Returns per-step losses, per-epoch val metrics and final params.
    """
    import torch.distributed as dist

    n, f = features.shape
    tr_idx, va_idx, model = seeded_setup(n, f, seed)
    tr_idx_t = torch.tensor(tr_idx)
    va_idx_t = torch.tensor(va_idx)
    net = model
    if use_ddp:
        net = nn.parallel.DistributedDataParallel(model)                        # :136 (C1 broadcast)
    opt = torch.optim.Adam(model.parameters(), lr=LR)                           # :88
    hist = {"train_loss": [], "val_loss": [], "val_acc": [], "masks": [], "batches": []}
    step = 0

    def loader_iter_draw():
        # every DataLoader iterator draws a base seed from the global generator when it is created
        # (torch/utils/data/dataloader.py `_BaseDataLoaderIter.__init__`), even with num_workers=0;
        # the draw shifts the dropout stream, so it is part of the arithmetic.
        torch.empty((), dtype=torch.int64).random_()

    if sanity_check:
        loader_iter_draw()       # Lightning's default num_sanity_val_steps=2 opens the val loader once
    for epoch in range(max_epochs):
        loader_iter_draw()                                   # train loader iterator
        if world > 1:
            order = epoch_indices(len(tr_idx), epoch, rank, world, seed)
        else:
            # RandomSampler then draws its own seed (torch/utils/data/sampler.py)
            s = int(torch.empty((), dtype=torch.int64).random_().item())
            g = torch.Generator()
            g.manual_seed(s)
            order = torch.randperm(len(tr_idx), generator=g).tolist()
        model.train()
        for b0 in range(0, len(order), batch_size):
            pos = order[b0:b0 + batch_size]
            rows = tr_idx_t[pos]
            if per_sample_loader:      # the reference's __getitem__ + default_collate path (:49,:122)
                x = torch.stack([features[i] for i in rows.tolist()])
                y = torch.stack([labels[i] for i in rows.tolist()])
            else:
                x, y = features[rows], labels[rows]
            if record_masks:
                st = torch.get_rng_state()
                hist["masks"].append(pack_mask(draw_dropout_mask(len(pos))))
                hist["batches"].append(rows.numpy().copy())
                torch.set_rng_state(st)
            opt.zero_grad(set_to_none=True)
            loss = F.cross_entropy(net(x), y)                                   # :68-69
            loss.backward()                                                     # DDP all-reduce (C2)
            opt.step()
            logged = loss.detach().clone()
            if use_ddp:
                dist.all_reduce(logged)                                         # :70 sync_dist (C3)
                logged /= world
            hist["train_loss"].append(float(logged))
            step += 1
            if max_steps is not None and step >= max_steps:
                break
        # validation epoch (:73-85), eval mode -> dropout off; epoch-level weighted mean over batches
        model.eval()
        loader_iter_draw()                                   # val loader iterator
        vpos = val_indices_for_rank(len(va_idx), rank, world) if world > 1 else list(range(len(va_idx)))
        tot_loss, tot_acc, tot_n = 0.0, 0.0, 0
        with torch.no_grad():
            for b0 in range(0, len(vpos), batch_size):
                rows = va_idx_t[vpos[b0:b0 + batch_size]]
                x, y = features[rows], labels[rows]
                logits = model(x)
                tot_loss += float(F.cross_entropy(logits, y)) * len(rows)
                tot_acc += float((torch.argmax(logits, 1) == y).float().mean()) * len(rows)
                tot_n += len(rows)
        stats = torch.tensor([tot_loss, tot_acc, float(tot_n)], dtype=torch.float64)
        if use_ddp:
            dist.all_reduce(stats)                                              # C4
        hist["val_loss"].append(float(stats[0] / stats[2]))
        hist["val_acc"].append(float(stats[1] / stats[2]))
        if max_steps is not None and step >= max_steps:
            break
    hist["params"] = flatten(model)
    hist["steps"] = step
    return hist


# ----------------------------------------------------------------------------------------------
# Timed CPU baseline: the reference's topology (gloo DDP, world_size 2, localhost) -- bench.py only.
# ----------------------------------------------------------------------------------------------
def _ddp_bench_worker(rank: int, world: int, port: int, n_rows: int, f: int, batch: int, steps: int,
                      warmup: int, threads: int, per_sample: bool, q) -> None:
    import time

    import torch.distributed as dist

    torch.set_num_threads(threads)
    # The workers may be spawned from a rank that torchrun started: drop every rendezvous variable that was
    # inherited (TORCHELASTIC_USE_AGENT_STORE makes env:// wait for the agent's store instead of hosting one)
    # and rendezvous over an explicit tcp:// address on a port the parent found free.
    for k in list(os.environ):
        if k.startswith("TORCHELASTIC_") or k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "GROUP_RANK",
                                                   "ROLE_RANK", "ROLE_WORLD_SIZE", "GROUP_WORLD_SIZE", "MASTER_ADDR",
                                                   "MASTER_PORT"):
            os.environ.pop(k, None)
    if world > 1:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    torch.manual_seed(SEED)
    g = torch.Generator().manual_seed(1234)
    feats = torch.randn(n_rows, f, generator=g)
    labels = (torch.rand(n_rows, generator=g) < 0.13).long()
    model = WeatherNet(f)
    net = nn.parallel.DistributedDataParallel(model) if world > 1 else model
    opt = torch.optim.Adam(model.parameters(), lr=LR)
    order = torch.randperm(n_rows, generator=g)
    model.train()

    def one(i):
        lo = (i * batch * world + rank * batch) % max(1, n_rows - batch)
        rows = order[lo:lo + batch]
        if per_sample:
            x = torch.stack([feats[j] for j in rows.tolist()])
            y = torch.stack([labels[j] for j in rows.tolist()])
        else:
            x, y = feats[rows], labels[rows]
        opt.zero_grad(set_to_none=True)
        loss = F.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
        logged = loss.detach().clone()
        if world > 1:
            dist.all_reduce(logged)

    for i in range(warmup):
        one(i)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        one(warmup + i)
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    if rank == 0:
        q.put(dt)
    if world > 1:
        dist.destroy_process_group()


def _free_port() -> int:
    import socket

    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return int(s.getsockname()[1])


def time_cpu_ddp(f: int, batch: int, steps: int, warmup: int = 20, world: int = 2, threads: int = 1,
                 n_rows: int = 145_460, per_sample: bool = True, port: int | None = None) -> dict:
    """Time `steps` optimizer steps of the restated reference step on CPU with gloo DDP.

    `per_sample=True` keeps the reference's per-row `__getitem__` + collate (:49,:122).
    Returns {"seconds", "samples_per_sec", "cores"} where cores = world * threads.
    """
    import torch.multiprocessing as mp

    if port is None:
        port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ddp_bench_worker,
                         args=(r, world, port, n_rows, f, batch, steps, warmup, threads, per_sample, q))
             for r in range(world)]
    for p in procs:
        p.start()
    import queue as _queue
    import time as _time

    dt, deadline = None, _time.time() + 1800
    while dt is None:
        try:
            dt = q.get(timeout=1.0)
        except _queue.Empty:
            dead = [p.exitcode for p in procs if p.exitcode not in (None, 0)]
            if dead or _time.time() > deadline:
                for p in procs:
                    if p.is_alive():
                        p.terminate()
                raise RuntimeError(f"time_cpu_ddp: worker exit codes {dead or 'timeout'}")
    for p in procs:
        p.join(timeout=60)
    return {"seconds": dt, "samples_per_sec": steps * batch * world / dt, "cores": world * threads}
