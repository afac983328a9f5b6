"""Torch-tensor front end of the C ABI.  torch is used for device memory and streams only; every
operation below is one or more kernels of libwxddp.so, launched on torch's current stream."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import (DROPOUT_MASK, DROPOUT_OFF, DROPOUT_PHILOX, IMPUTE_MEAN, IMPUTE_NONE, WX_F32, WX_F64,
                   ModelDesc, WxError, check)

HIDDEN, CLASSES = 64, 2


def n_params(n_features: int) -> int:
    return HIDDEN * n_features + HIDDEN + CLASSES * HIDDEN + CLASSES


def _p(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


class Columns:
    """A table's columns bound once: the checked tensors, their device pointers as the HOST pointer array the C ABI takes,
    dtype and length.  Passing this instead of a list of tensors to the raw-column calls skips the per-call validation and
    pointer gathering (about 20 us of Python per step for 20 columns -- the whole step at small batch sizes)."""

    def __init__(self, tensors, cp, vp, dt, n):
        self.tensors, self.cp, self.vp, self.dt, self.n = list(tensors), cp, vp, dt, n
        self.ptrs = [t.data_ptr() for t in self.tensors]
        self.dtype = self.tensors[0].dtype

    def __len__(self):
        return len(self.tensors)

    def __getitem__(self, j):
        return self.tensors[j]


class Context:
    """One per process / GPU (wx_ctx).  Owns scratch only; all tensors stay caller-owned."""

    def __init__(self, n_features: int, device: int | torch.device = 0, dropout_p: float = 0.2,
                 lr: float = 0.01, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, optimizer: str = "adam"):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise WxError("wxddp needs a CUDA device (sm_100a); there is no CPU path")
        dev = torch.device(device) if not isinstance(device, int) else torch.device("cuda", device)
        self.device = dev
        self.n_features = int(n_features)
        self.P = n_params(self.n_features)
        torch.cuda.init()
        with torch.cuda.device(dev):
            torch.cuda.current_stream()      # make sure torch's primary context exists before ours attaches
        if optimizer not in ("adam", "sgd"):
            raise WxError(f"optimizer={optimizer!r}: 'adam' (the reference's, train_lightning_ddp.py:88) or 'sgd'")
        desc = ModelDesc(self.n_features, HIDDEN, CLASSES, dropout_p, lr, beta1, beta2, eps, _lib.OPT_SGD if optimizer == "sgd" else _lib.OPT_ADAM)
        h = C.c_void_p()
        check(self.lib.wx_create(dev.index or 0, C.byref(desc), C.byref(h)))
        self._h = h
        self.peer_world = 1

    # -- bookkeeping ---------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None):
            self.lib.wx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_sms(self) -> int:
        return self.lib.wx_num_sms(self._h)

    @property
    def launch_count(self) -> int:
        return self.lib.wx_launch_count(self._h)

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def _chk(self, t: torch.Tensor | None, dtype, name: str, numel: int | None = None) -> None:
        if t is None:
            return
        if not t.is_cuda or t.device != self.device:
            raise WxError(f"{name} must live on {self.device}")
        if t.dtype != dtype:
            raise WxError(f"{name} must be {dtype}, got {t.dtype}")
        if not t.is_contiguous():
            raise WxError(f"{name} must be contiguous")
        if numel is not None and t.numel() < numel:
            raise WxError(f"{name} has {t.numel()} elements, needs {numel}")

    def bind_columns(self, cols, validity=None) -> Columns:
        """Validate a table's columns once and keep the pointer array for repeated raw-column steps."""
        if isinstance(cols, Columns):
            return cols
        cp, vp, dt, n = self._cols(cols, validity)
        return Columns(cols, cp, vp, dt, n)

    def _cols(self, cols, validity):
        if isinstance(cols, Columns):
            if len(cols) != self.n_features:
                raise WxError(f"expected {self.n_features} columns, got {len(cols)}")
            return cols.cp, cols.vp, cols.dt, cols.n
        if len(cols) != self.n_features:
            raise WxError(f"expected {self.n_features} columns, got {len(cols)}")
        dt = cols[0].dtype
        if dt not in (torch.float32, torch.float64):
            raise WxError("columns must be float32 or float64")
        n = cols[0].numel()
        for j, c in enumerate(cols):
            self._chk(c, dt, f"column {j}", n)
            if c.numel() != n:
                raise WxError("columns differ in length")
        cp = _lib.ptr_array([c.data_ptr() for c in cols])
        vp = None
        if validity is not None:
            for j, v in enumerate(validity):
                self._chk(v, torch.uint8, f"validity {j}", (n + 7) // 8)
            vp = _lib.ptr_array([None if v is None else v.data_ptr() for v in validity])
        return cp, vp, (WX_F32 if dt == torch.float32 else WX_F64), n

    # -- K1 -------------------------------------------------------------------------------------
    def label_encode(self, offsets: torch.Tensor, data: torch.Tensor, validity: torch.Tensor | None = None,
                     want_int32: bool = False):
        """preprocess.py:23-25 over an Arrow utf8 column.  Returns int64 labels (and int32 if asked)."""
        self._chk(offsets, torch.int32, "offsets")
        self._chk(data, torch.uint8, "data")
        n = offsets.numel() - 1
        self._chk(validity, torch.uint8, "validity", (n + 7) // 8)
        out = torch.empty(n, dtype=torch.int64, device=self.device)
        out32 = torch.empty(n, dtype=torch.int32, device=self.device) if want_int32 else None
        if data.numel() == 0:
            data = torch.zeros(1, dtype=torch.uint8, device=self.device)
        check(self.lib.wx_label_encode(self._h, _p(offsets), _p(data), _p(validity), n, _p(out), _p(out32),
                                       self._stream()))
        return (out, out32) if want_int32 else out

    def col_stats(self, cols, validity=None) -> torch.Tensor:
        """preprocess.py:32-36: [F,4] float64 = count, mean, M2, std_val (zero-guarded)."""
        cp, vp, dt, n = self._cols(cols, validity)
        out = torch.empty(self.n_features, 4, dtype=torch.float64, device=self.device)
        check(self.lib.wx_col_stats(self._h, cp, dt, vp, n, _p(out), self._stream()))
        return out

    def stats_merge(self, parts: torch.Tensor) -> torch.Tensor:
        self._chk(parts, torch.float64, "parts")
        w = parts.shape[0]
        out = torch.empty(self.n_features, 4, dtype=torch.float64, device=self.device)
        check(self.lib.wx_stats_merge(self._h, _p(parts), w, _p(out), self._stream()))
        return out

    def normalize_pack(self, cols, stats: torch.Tensor, impute: int = IMPUTE_NONE, validity=None,
                       want_nullmask: bool = True, out: torch.Tensor | None = None):
        """preprocess.py:38-41 + train_lightning_ddp.py:45: row-major fp32 [N,F] (+ per-row null mask)."""
        cp, vp, dt, n = self._cols(cols, validity)
        self._chk(stats, torch.float64, "stats", self.n_features * 4)
        if out is None:
            out = torch.empty(n, self.n_features, dtype=torch.float32, device=self.device)
        self._chk(out, torch.float32, "out", n * self.n_features)
        nm = torch.empty(n, dtype=torch.int32, device=self.device) if want_nullmask else None
        check(self.lib.wx_normalize_pack(self._h, cp, dt, vp, n, _p(stats), impute, _p(out), _p(nm), self._stream()))
        return out, nm

    def zscore_f64(self, cols, stats: torch.Tensor, validity=None, out: torch.Tensor | None = None) -> torch.Tensor:
        """preprocess.py:38-41 as Spark writes it: columnar float64 `(col - mean) / std_val`, [F,N]; nulls -> NaN."""
        cp, vp, dt, n = self._cols(cols, validity)
        self._chk(stats, torch.float64, "stats", self.n_features * 4)
        if out is None:
            out = torch.empty(self.n_features, n, dtype=torch.float64, device=self.device)
        self._chk(out, torch.float64, "out", n * self.n_features)
        op = _lib.ptr_array([out[j].data_ptr() for j in range(self.n_features)])
        check(self.lib.wx_zscore_f64(self._h, cp, dt, vp, n, _p(stats), op, self._stream()))
        return out

    def impute_index_list(self, nullmask: torch.Tensor, capacity: int | None = None):
        """Ascending flat indices row*F+col of the null cells; returns (idx[:min(count,capacity)], count)."""
        self._chk(nullmask, torch.int32, "nullmask")
        n = nullmask.numel()
        if capacity is None:
            capacity = n * self.n_features
        idx = torch.empty(max(capacity, 1), dtype=torch.int64, device=self.device)
        cnt = torch.zeros(1, dtype=torch.int64, device=self.device)
        check(self.lib.wx_impute_index_list(self._h, _p(nullmask), n, _p(idx), capacity, _p(cnt), self._stream()))
        c = int(cnt.item())
        return idx[:min(c, capacity)], c

    # -- K2 / K3 / K5 ---------------------------------------------------------------------------
    def _batch_args(self, x, y, row_idx, row0, batch, keep_mask, dropout):
        self._chk(x, torch.float32, "x")
        self._chk(y, torch.int64, "y")
        self._chk(row_idx, torch.int64, "row_idx")
        n_rows = x.numel() // self.n_features
        if batch is None:
            batch = row_idx.numel() if row_idx is not None else n_rows - row0
        if row_idx is not None and row_idx.numel() < batch:
            raise WxError("row_idx shorter than batch")
        if row_idx is None and (row0 < 0 or row0 + batch > n_rows):
            raise WxError(f"rows {row0}..{row0 + batch} outside the table of {n_rows}")
        if dropout == DROPOUT_MASK:
            self._chk(keep_mask, torch.int64, "keep_mask", batch)
        return batch

    def mlp_grad(self, params, x, y, *, row_idx=None, row0: int = 0, batch: int | None = None,
                 dropout: int = DROPOUT_OFF, keep_mask=None, seed: int = 0, step: int = 0,
                 scale: float | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
        """train_lightning_ddp.py:57-71: out[0:P] = scale * sum_rows grad, out[P] = scale * sum_rows loss."""
        self._chk(params, torch.float32, "params", self.P)
        batch = self._batch_args(x, y, row_idx, row0, batch, keep_mask, dropout)
        if scale is None:
            scale = 1.0 / batch
        if out is None:
            out = torch.empty(self.P + 1, dtype=torch.float32, device=self.device)
        self._chk(out, torch.float32, "out", self.P + 1)
        check(self.lib.wx_mlp_grad(self._h, _p(params), _p(x), _p(y), _p(row_idx), row0, batch, dropout,
                                   _p(keep_mask), seed, step, scale, _p(out), self._stream()))
        return out

    def tc_mlp_grad(self, params, x, y, *, row0: int = 0, batch: int | None = None, dropout: int = DROPOUT_OFF,
                    keep_mask=None, seed: int = 0, step: int = 0, scale: float | None = None,
                    out: torch.Tensor | None = None) -> torch.Tensor:
        """wx_mlp_grad on the tensor cores (tcgen05 tf32x3): contiguous rows, F <= 29."""
        self._chk(params, torch.float32, "params", self.P)
        batch = self._batch_args(x, y, None, row0, batch, keep_mask, dropout)
        if scale is None:
            scale = 1.0 / batch
        if out is None:
            out = torch.empty(self.P + 1, dtype=torch.float32, device=self.device)
        self._chk(out, torch.float32, "out", self.P + 1)
        check(self.lib.wx_tc_mlp_grad(self._h, _p(params), _p(x), _p(y), row0, batch, dropout, _p(keep_mask), seed, step,
                                      scale, _p(out), self._stream()))
        return out

    def tc_train_step(self, params, m, v, state, x, y, loss_out, *, row0: int = 0, batch: int | None = None,
                      dropout: int = DROPOUT_PHILOX, keep_mask=None, seed: int = 0) -> None:
        """wx_train_step on the tensor cores: one launch per optimizer step."""
        self._opt_args(params, m, v, state, loss_out)
        batch = self._batch_args(x, y, None, row0, batch, keep_mask, dropout)
        check(self.lib.wx_tc_train_step(self._h, _p(params), _p(m), _p(v), _p(state), _p(x), _p(y), row0, batch, dropout,
                                        _p(keep_mask), seed, _p(loss_out), self._stream()))

    def tc_train_step_raw(self, params, m, v, state, cols, stats, y, loss_out, *, row0: int = 0,
                          batch: int | None = None, impute: int = IMPUTE_NONE, dropout: int = DROPOUT_PHILOX,
                          keep_mask=None, seed: int = 0) -> None:
        """wx_train_step_raw on the tensor cores: raw fp32 columns in, standardised in the kernel."""
        self._opt_args(params, m, v, state, loss_out)
        cp, _, dt, n = self._cols(cols, None)
        if dt != WX_F32:
            raise WxError("tc_train_step_raw takes float32 columns (use train_step_raw for float64)")
        self._chk(stats, torch.float64, "stats", self.n_features * 4)
        self._chk(y, torch.int64, "y", n)
        if batch is None:
            batch = n - row0
        if row0 < 0 or row0 + batch > n:
            raise WxError(f"rows {row0}..{row0 + batch} outside the table of {n}")
        if dropout == DROPOUT_MASK:
            self._chk(keep_mask, torch.int64, "keep_mask", batch)
        check(self.lib.wx_tc_train_step_raw(self._h, _p(params), _p(m), _p(v), _p(state), cp, _p(stats), impute, _p(y),
                                            row0, batch, dropout, _p(keep_mask), seed, _p(loss_out), self._stream()))

    # -- multi-GPU: in-kernel gradient exchange over NVLink peer memory ---------------------------------
    def peer_export(self) -> bytes:
        """This rank's inbox as a 64-byte CUDA IPC handle (allocates the inbox on first use)."""
        h = (C.c_ubyte * 64)()
        check(self.lib.wx_peer_export(self._h, h))
        return bytes(h)

    def peer_connect(self, rank: int, world: int, handles: list[bytes]) -> None:
        """Map every rank's inbox; afterwards tc_train_step / tc_train_step_raw all-reduce inside the kernel."""
        blob = b"".join(handles)
        if len(blob) != 64 * world:
            raise WxError("peer_connect needs one 64-byte handle per rank")
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        check(self.lib.wx_peer_connect(self._h, rank, world, buf))
        self.peer_world = world

    # -- kernel choice --------------------------------------------------------------------------------
    TC_MAX_FEATURES = 31

    def tc_ok_rowmajor(self, x, y, row0: int, row_idx=None) -> bool:
        """True when the tensor-core step kernel can take this batch (contiguous rows, F <= 31, TMA alignment)."""
        return (row_idx is None and self.n_features <= self.TC_MAX_FEATURES
                and (x.data_ptr() + row0 * self.n_features * 4) % 16 == 0 and (y.data_ptr() + row0 * 8) % 16 == 0)

    def tc_ok_raw(self, cols, y, row0: int) -> bool:
        if isinstance(cols, Columns):
            if row0 % 4 or cols.dtype != torch.float32 or self.n_features > self.TC_MAX_FEATURES:
                return False
            if getattr(cols, "_al16", None) is None:
                cols._al16 = all(p % 16 == 0 for p in cols.ptrs)
            return cols._al16 and (y.data_ptr() + row0 * 8) % 16 == 0
        return (self.n_features <= self.TC_MAX_FEATURES and cols[0].dtype == torch.float32
                and all((c.data_ptr() + row0 * 4) % 16 == 0 for c in cols) and (y.data_ptr() + row0 * 8) % 16 == 0)

    def tc_status(self) -> int:
        out = C.c_int(0)
        check(self.lib.wx_tc_last_status(self._h, C.byref(out)))
        return out.value

    def adam_step(self, params, grads, m, v, step: int, n: int | None = None) -> None:
        """train_lightning_ddp.py:88, torch.optim.Adam single-tensor arithmetic; step is 1-based."""
        n = params.numel() if n is None else n
        for t, name in ((params, "params"), (grads, "grads"), (m, "m"), (v, "v")):
            self._chk(t, torch.float32, name, n)
        check(self.lib.wx_adam_step(self._h, _p(params), _p(grads), _p(m), _p(v), n, step, self._stream()))

    def _opt_args(self, params, m, v, state, loss_out, n_loss=1):
        for t, name in ((params, "params"), (m, "m"), (v, "v")):
            self._chk(t, torch.float32, name, self.P)
        self._chk(state, torch.int64, "state", 2)
        self._chk(loss_out, torch.float32, "loss_out", n_loss)

    def train_step(self, params, m, v, state, x, y, loss_out, *, row_idx=None, row0: int = 0,
                   batch: int | None = None, dropout: int = DROPOUT_PHILOX, keep_mask=None, seed: int = 0) -> None:
        """One optimizer step (forward, loss, backward, Adam) in one launch."""
        self._opt_args(params, m, v, state, loss_out)
        batch = self._batch_args(x, y, row_idx, row0, batch, keep_mask, dropout)
        check(self.lib.wx_train_step(self._h, _p(params), _p(m), _p(v), _p(state), _p(x), _p(y), _p(row_idx), row0,
                                     batch, dropout, _p(keep_mask), seed, _p(loss_out), self._stream()))

    def train_step_raw(self, params, m, v, state, cols, stats, y, loss_out, *, row0: int = 0,
                       batch: int | None = None, impute: int = IMPUTE_NONE, dropout: int = DROPOUT_PHILOX,
                       keep_mask=None, seed: int = 0) -> None:
        """Same step from raw columns: standardise in registers, then train (preprocess.py:38-41 fused in)."""
        self._opt_args(params, m, v, state, loss_out)
        cp, _, dt, n = self._cols(cols, None)
        self._chk(stats, torch.float64, "stats", self.n_features * 4)
        self._chk(y, torch.int64, "y", n)
        if batch is None:
            batch = n - row0
        if row0 < 0 or row0 + batch > n:
            raise WxError(f"rows {row0}..{row0 + batch} outside the table of {n}")
        if dropout == DROPOUT_MASK:
            self._chk(keep_mask, torch.int64, "keep_mask", batch)
        check(self.lib.wx_train_step_raw(self._h, _p(params), _p(m), _p(v), _p(state), cp, dt, _p(stats), impute,
                                         _p(y), row0, batch, dropout, _p(keep_mask), seed, _p(loss_out),
                                         self._stream()))

    def train_steps_small(self, params, m, v, state, x, y, loss_out, *, n_steps: int, batch: int, row_idx=None,
                          row0: int = 0, dropout: int = DROPOUT_PHILOX, keep_mask=None, seed: int = 0) -> None:
        """n_steps optimizer steps of `batch` (<= 64) rows each in ONE launch (the reference's batch-4 regime)."""
        self._opt_args(params, m, v, state, loss_out, n_steps)
        self._chk(x, torch.float32, "x")
        self._chk(y, torch.int64, "y")
        self._chk(row_idx, torch.int64, "row_idx", n_steps * batch)
        if dropout == DROPOUT_MASK:
            self._chk(keep_mask, torch.int64, "keep_mask", n_steps * batch)
        if row_idx is None and row0 + n_steps * batch > x.numel() // self.n_features:
            raise WxError("row window runs past the table")
        check(self.lib.wx_train_steps_small(self._h, _p(params), _p(m), _p(v), _p(state), _p(x), _p(y), _p(row_idx),
                                            row0, batch, n_steps, dropout, _p(keep_mask), seed, _p(loss_out),
                                            self._stream()))

    def mlp_eval(self, params, x, y, *, row_idx=None, row0: int = 0, batch: int | None = None,
                 out: torch.Tensor | None = None) -> torch.Tensor:
        """validation_step (:73-85): device float64[2] = (sum of row losses, number correct)."""
        self._chk(params, torch.float32, "params", self.P)
        batch = self._batch_args(x, y, row_idx, row0, batch, None, DROPOUT_OFF)
        if out is None:
            out = torch.empty(2, dtype=torch.float64, device=self.device)
        self._chk(out, torch.float64, "out", 2)
        check(self.lib.wx_mlp_eval(self._h, _p(params), _p(x), _p(y), _p(row_idx), row0, batch, _p(out),
                                   self._stream()))
        return out

    def mlp_predict(self, params, x) -> torch.Tensor:
        """score.py run(): softmax probabilities [N,2]."""
        self._chk(params, torch.float32, "params", self.P)
        self._chk(x, torch.float32, "x")
        n = x.numel() // self.n_features
        out = torch.empty(n, 2, dtype=torch.float32, device=self.device)
        check(self.lib.wx_mlp_predict(self._h, _p(params), _p(x), n, _p(out), self._stream()))
        return out

    def mlp_logits(self, params, x) -> torch.Tensor:
        """forward() in eval mode (train_lightning_ddp.py:64): the two raw logits per row, [N,2]."""
        self._chk(params, torch.float32, "params", self.P)
        self._chk(x, torch.float32, "x")
        n = x.numel() // self.n_features
        out = torch.empty(n, 2, dtype=torch.float32, device=self.device)
        check(self.lib.wx_mlp_logits(self._h, _p(params), _p(x), n, _p(out), self._stream()))
        return out
