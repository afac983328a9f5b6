"""CPU oracle of the wide tabular MLP step (test infrastructure only -- never imported by the product path).

The reference has no wide model: its net is Linear(F,64)-ReLU-Dropout(0.2)-Linear(64,2)
(/root/reference/jobs/train_lightning_ddp.py:57-62) with CE loss (:69) and Adam(lr=0.01) (:88).  BASELINE.json
configs[3] widens that same pattern (n_hidden x [Linear, ReLU, Dropout], Linear head); this file restates it in plain
torch fp32 and, with `emulate_bf16=True`, rounds to bfloat16 at exactly the points where the CUDA path stores bf16
(inputs, hidden weights, hidden activations, dZ), so the only remaining difference is fp32 summation order.
Parity of the layer pattern itself is pinned through the narrow model (tests/golden/*, oracle/train_ref.py): with
hidden=64, n_hidden=1 and emulate_bf16=False `step()` is that model (checked in tests/test_oracle.py).
"""
from __future__ import annotations

import numpy as np
import torch

from . import philox_ref


def n_params(f: int, h: int, n_hidden: int, c: int = 2) -> int:
    return h * f + h + (n_hidden - 1) * (h * h + h) + c * h + c


def split(flat: torch.Tensor, f: int, h: int, n_hidden: int, c: int = 2):
    """[(W_1, b_1), ..., (W_L, b_L), (W_o, b_o)] views of the flat vector (torch module order)."""
    out, off = [], 0
    for l in range(n_hidden + 1):
        rows = h if l < n_hidden else c
        cols = f if l == 0 else h
        w = flat[off:off + rows * cols].view(rows, cols)
        off += rows * cols
        b = flat[off:off + rows]
        off += rows
        out.append((w, b))
    assert off == flat.numel()
    return out


def init_params(f: int, h: int, n_hidden: int, c: int = 2) -> torch.Tensor:
    """nn.Linear default init of every layer from the current global generator, in module order (:57-62)."""
    from torch import nn
    parts = []
    for l in range(n_hidden + 1):
        lin = nn.Linear(f if l == 0 else h, h if l < n_hidden else c)
        parts += [lin.weight.detach().reshape(-1), lin.bias.detach()]
    return torch.cat(parts).clone()


def keep_mask(seed: int, step: int, layer: int, rows: int, cols: int, p_drop: float) -> np.ndarray:
    """[rows, cols] bool: the kernels' dropout stream.  Philox4x32-10, counter (row, col // 8, layer, step), key seed;
    element col % 8 = e takes 16 bits: word e // 2, half e % 2 (low half first); keep <=> bits < round((1-p) * 65536)."""
    thresh = int(np.rint(np.float32(1.0 - p_drop) * np.float32(65536.0)))
    r = np.arange(rows, dtype=np.uint32)[:, None]
    cb = np.arange(cols // 8, dtype=np.uint32)[None, :]
    w = philox_ref.philox4x32_10(r, cb, np.uint32(layer), np.uint32(step & 0xFFFFFFFF), seed & 0xFFFFFFFF,
                                 (seed >> 32) & 0xFFFFFFFF)
    out = np.empty((rows, cols // 8, 8), dtype=bool)
    for e in range(8):
        bits = (w[e >> 1] >> np.uint32(16 * (e & 1))) & np.uint32(0xFFFF)
        out[:, :, e] = bits < thresh
    return out.reshape(rows, cols)


def _r(t: torch.Tensor, on: bool) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32) if on else t


def loss_and_grads(flat: torch.Tensor, x: torch.Tensor, y: torch.Tensor, f: int, h: int, n_hidden: int, *,
                   p_drop: float = 0.2, dropout: bool = True, seed: int = 42, step: int = 0, grad_scale=None,
                   emulate_bf16: bool = True, masks=None, intermediates: dict | None = None):
    """Manual forward/backward in the CUDA path's operation order.  Returns (loss*scale*B... see below, grads [P+1]).

    grads[P] = sum_rows(CE) * grad_scale, like the kernels' bucket (grad_scale defaults to 1/B -> mean loss)."""
    B = x.shape[0]
    scale = (1.0 / B) if grad_scale is None else grad_scale
    layers = split(flat, f, h, n_hidden)
    e = emulate_bf16
    inv_keep = np.float32(1.0) / np.float32(1.0 - p_drop) if dropout and p_drop > 0 else np.float32(1.0)
    hs = [_r(x.float(), e)]
    for l in range(n_hidden):
        w, b = layers[l]
        z = hs[-1] @ _r(w, e).t() + b
        a = torch.relu(z)
        if dropout and p_drop > 0:
            if masks is not None:
                keep = masks[l]
            else:
                keep = torch.from_numpy(keep_mask(seed, step, l + 1, B, h, p_drop))
            a = torch.where(keep, a * float(inv_keep), torch.zeros_like(a))
        hs.append(_r(a, e))
    wo, bo = layers[n_hidden]
    logits = hs[-1] @ wo.t() + bo
    lse = torch.logsumexp(logits, dim=1)
    ce = lse - logits.gather(1, y.view(-1, 1)).squeeze(1)
    dlog = (torch.softmax(logits, dim=1) - torch.nn.functional.one_hot(y, 2).float()) * scale
    grads = torch.zeros(flat.numel() + 1)
    g_layers = split(grads[:-1], f, h, n_hidden)
    g_layers[n_hidden][0].copy_(dlog.t() @ hs[-1])
    g_layers[n_hidden][1].copy_(dlog.sum(0))
    grads[-1] = ce.sum() * scale
    dz = _r((dlog @ wo) * (hs[-1] > 0) * float(inv_keep), e)
    if intermediates is not None:
        intermediates.update(hs=hs, dlog=dlog, dz={n_hidden: dz})
    for l in range(n_hidden - 1, -1, -1):
        g_layers[l][0].copy_(dz.t() @ hs[l])
        g_layers[l][1].copy_(dz.sum(0))
        if l > 0:
            dz = _r((dz @ _r(layers[l][0], e)) * (hs[l] > 0) * float(inv_keep), e)
            if intermediates is not None:
                intermediates["dz"][l] = dz
    return grads


def torch_autograd_grads(flat: torch.Tensor, x: torch.Tensor, y: torch.Tensor, f: int, h: int, n_hidden: int, masks,
                         p_drop: float = 0.2):
    """The same step through nn-style autograd in pure fp32 (no bf16 rounding): mean CE, grads [P+1]."""
    p = flat.clone().requires_grad_(True)
    layers = split(p, f, h, n_hidden)
    a = x.float()
    for l in range(n_hidden):
        a = torch.relu(a @ layers[l][0].t() + layers[l][1])
        if masks is not None:
            a = a * masks[l].float() / (1.0 - p_drop)
    logits = a @ layers[n_hidden][0].t() + layers[n_hidden][1]
    loss = torch.nn.functional.cross_entropy(logits, y)
    loss.backward()
    return torch.cat([p.grad, loss.detach().view(1)])


def time_cpu_steps(f: int, h: int, n_hidden: int, batch: int, steps: int, warmup: int = 1, threads: int | None = None):
    """The same net as plain torch modules (nn.Sequential of Linear/ReLU/Dropout(0.2) + CE + torch.optim.Adam(lr=0.01),
    the reference's recipe at train_lightning_ddp.py:57-62,:69,:88 widened) timed on the host cores, fp32."""
    import time
    from torch import nn
    if threads:
        torch.set_num_threads(threads)
    torch.manual_seed(42)
    mods, d = [], f
    for _ in range(n_hidden):
        mods += [nn.Linear(d, h), nn.ReLU(), nn.Dropout(0.2)]
        d = h
    mods.append(nn.Linear(d, 2))
    net = nn.Sequential(*mods)
    opt = torch.optim.Adam(net.parameters(), lr=0.01)
    x = torch.randn(batch, f)
    y = (torch.rand(batch) < 0.3).long()
    def one():
        opt.zero_grad(set_to_none=True)
        loss = torch.nn.functional.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
        return float(loss.detach())
    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        last = one()
    dt = time.perf_counter() - t0
    return {"samples_per_sec": steps * batch / dt, "seconds": dt, "cores": torch.get_num_threads(), "last_loss": last}
