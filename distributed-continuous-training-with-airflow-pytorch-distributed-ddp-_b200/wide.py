"""Wide tabular MLP (BASELINE.json configs[3]: 4 layers, hidden 4096) on the bf16 tcgen05 GEMMs of libwxddp.so.

Same operator shape as the narrow StepEngine: one object per process / GPU holding flat fp32 parameters, Adam moments and
the workspace; `step(x, y)` = forward, CE loss, backward, DDP gradient mean over ranks, Adam -- the loop Lightning runs
around jobs/train_lightning_ddp.py:66-71,:88,:136, for a net of the same pattern (:57-62) widened.  The DDP exchange is
bucketed by layer: the all-reduce of layer l's slice runs on a side stream while layer l-1's backward GEMMs run.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from ._lib import DROPOUT_OFF, DROPOUT_PHILOX, ModelDesc, WideDesc, WxError, check, load


class WideEngine:
    def __init__(self, n_features: int, hidden: int = 4096, n_hidden: int = 3, *, max_batch: int, device=0, world: int = 1,
                 rank: int = 0, process_group=None, params: torch.Tensor | None = None, seed: int = 42, lr: float = 0.01,
                 dropout_p: float = 0.2):
        if not torch.cuda.is_available():
            raise WxError("wxddp needs a CUDA device (B200, sm_100a): there is no CPU fallback")
        self.device = torch.device("cuda", device) if not isinstance(device, torch.device) else device
        self.lib = load()
        self.desc = WideDesc(n_features, hidden, n_hidden, 2, dropout_p, lr, 0.9, 0.999, 1e-8)
        self.F, self.H, self.L = n_features, hidden, n_hidden
        self.P = int(self.lib.wx_wide_n_params(C.byref(self.desc)))
        if self.P < 0:
            check(-1)
        # a context of the narrow model carries the device, scratch and launch counter (wx_create is the only allocator)
        h = C.c_void_p()
        nd = ModelDesc(min(n_features, 32), 64, 2, 0.2, lr, 0.9, 0.999, 1e-8)
        check(self.lib.wx_create(self.device.index, C.byref(nd), C.byref(h)))
        self._h = h
        self.max_batch = max_batch
        nbytes = int(self.lib.wx_wide_workspace_bytes(C.byref(self.desc), max_batch))
        if nbytes < 0:
            check(-1)
        self._ws_raw = torch.empty(nbytes + 1024, dtype=torch.uint8, device=self.device)
        skew = (-self._ws_raw.data_ptr()) % 1024                     # the C ABI wants the workspace 1024-byte aligned
        self.ws = self._ws_raw[skew:skew + nbytes]
        self.world, self.rank, self.dist = world, rank, process_group
        if world > 1 and process_group is None:
            raise WxError("world > 1 needs an initialised torch.distributed (NCCL) handle")
        if params is None:
            from torch import nn
            parts = []
            for l in range(n_hidden + 1):
                lin = nn.Linear(n_features if l == 0 else hidden, hidden if l < n_hidden else 2)
                parts += [lin.weight.detach().reshape(-1), lin.bias.detach()]
            params = torch.cat(parts)
        if params.numel() != self.P:
            raise WxError(f"params has {params.numel()} elements, the model has {self.P}")
        self.params = params.detach().to(self.device, torch.float32).contiguous().clone()
        if world > 1:
            self.dist.broadcast(self.params, src=0)
        self.m = torch.zeros(self.P, device=self.device)
        self.v = torch.zeros(self.P, device=self.device)
        self.grads = torch.zeros(self.P + 1, device=self.device)
        self.seed, self.steps_done = seed, 0
        self._side = torch.cuda.Stream(device=self.device) if world > 1 else None
        # slices of the flat gradient per hidden layer (W_l, b_l contiguous), head last (+ the loss element)
        self.slices, off = [], 0
        for l in range(n_hidden):
            n = hidden * (n_features if l == 0 else hidden) + hidden
            self.slices.append((off, off + n))
            off += n
        self.head_slice = (off, self.P + 1)
        # under DDP the GEMM grids leave 16 SMs to NCCL, whose kernels otherwise queue behind the one-CTA-per-SM persistent
        # GEMMs (8 GPUs: 3.10 -> 2.92 ms/step; profiles/r1_batch_sweep_and_wide.md)
        self._reserve = int(os.environ.get("WX_WIDE_RESERVE_SMS", "16" if world > 1 else "0"))
        self.cast_weights()

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self.lib.wx_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @property
    def launch_count(self) -> int:
        return int(self.lib.wx_launch_count(self._h))

    def status(self) -> int:
        out = C.c_int(0)
        check(self.lib.wx_tc_last_status(self._h, C.byref(out)))
        return out.value

    def cast_weights(self) -> None:
        check(self.lib.wx_wide_cast_weights(self._h, C.byref(self.desc), self.params.data_ptr(), self.ws.data_ptr(),
                                            self._stream()))

    def _check_batch(self, x, y):
        b = x.shape[0]
        if x.dtype != torch.float32 or not x.is_contiguous() or x.shape[1] != self.F or x.device != self.device:
            raise WxError("x must be a contiguous fp32 [batch, F] tensor on the engine's device")
        if y.dtype != torch.int64 or y.shape[0] != b or y.device != self.device:
            raise WxError("y must be int64 [batch] on the engine's device")
        if b > self.max_batch:
            raise WxError(f"batch {b} exceeds max_batch {self.max_batch}")
        return b

    def loss_and_grads(self, x, y, *, dropout: int = DROPOUT_PHILOX, step: int | None = None, grad_scale=None,
                       overlap_reduce: bool = False) -> torch.Tensor:
        """Forward + backward of one batch into self.grads ([P+1], loss at [P]); returns self.grads.  With
        overlap_reduce the gradient slices are SUM-all-reduced across ranks on a side stream, layer by layer."""
        self._run(x, y, dropout, step, grad_scale, reduce=overlap_reduce, adam_step=None)
        return self.grads

    def _run(self, x, y, dropout, step, grad_scale, *, reduce: bool, adam_step: int | None) -> None:
        """The step's schedule.  Main stream: forward, head, then per layer dW and dZ-below GEMMs.  Side stream: bias
        gradients (column sums of dZ_l, which only read dZ_l), the layer's all-reduce, and -- when adam_step is given --
        that layer's Adam, all running under the GEMMs of the layers below.  Events keep three hazards in order:
        dZ buffers ping-pong (the dZ product below may not overwrite a buffer the side stream still sums), Adam rewrites
        the bf16 weights (only after the layer's own dZ product has read them), and the next step's forward needs all
        of it (the streams join at the end)."""
        b = self._check_batch(x, y)
        step = self.steps_done if step is None else step
        scale = 1.0 / (self.world * b) if grad_scale is None else grad_scale
        d = C.byref(self.desc)
        lib, h, ws, g = self.lib, self._h, self.ws.data_ptr(), self.grads.data_ptr()
        main = torch.cuda.current_stream(self.device)
        if getattr(self, "_aux", None) is None:
            self._aux = torch.cuda.Stream(device=self.device)
        side = self._aux
        mst, sst = C.c_void_p(main.cuda_stream), C.c_void_p(side.cuda_stream)

        def after_main():
            ev = torch.cuda.Event()
            ev.record(main)
            side.wait_event(ev)

        def reduce_slice(lo, hi):
            if reduce:                                         # DDP SUM of the slice, on the side stream
                with torch.cuda.stream(side):
                    self.dist.all_reduce(self.grads[lo:hi])

        def adam(layer):
            if adam_step is not None:
                check(lib.wx_wide_adam_layer(h, d, self.params.data_ptr(), g, self.m.data_ptr(), self.v.data_ptr(),
                                             adam_step, ws, layer, sst))

        reserve = self._reserve if reduce else 0
        if self._reserve:
            check(lib.wx_wide_reserve_sms(h, 0))               # nothing else runs during the forward: all SMs
        check(lib.wx_wide_forward(h, d, self.params.data_ptr(), ws, x.data_ptr(), y.data_ptr(), b, dropout, self.seed, step,
                                  scale, g, mst))
        after_main()                                           # dZ_L, db_o and the loss are enqueued
        check(lib.wx_wide_backward_layer(h, d, ws, b, self.L, dropout, 16, g, sst))            # dW_o (side stream)
        reduce_slice(*self.head_slice)
        adam(self.L + 1)
        db_done = {}
        for layer in range(self.L, 0, -1):
            # side: bias gradient of this layer (dZ_layer is ready: the side stream already waited for its producer)
            # (below the top hidden layer it is a sum of the 32-row partials the dZ product above wrote beside dZ_layer)
            check(lib.wx_wide_backward_layer(h, d, ws, b, layer, dropout, 2 if layer == self.L else 32, g, sst))
            db_done[layer] = torch.cuda.Event()
            db_done[layer].record(side)
            check(lib.wx_wide_backward_layer(h, d, ws, b, layer, dropout, 1, g, mst))          # dW
            if reserve and layer == self.L:
                check(lib.wx_wide_reserve_sms(h, reserve))     # from here on a big all-reduce is always in flight
            after_main()
            check(lib.wx_wide_backward_layer(h, d, ws, b, layer, dropout, 8, g, sst))          # fold split-K slab (side)
            reduce_slice(*self.slices[layer - 1])             # starts as soon as dW is final, under the dZ product
            if layer > 1:
                if layer + 1 in db_done:
                    main.wait_event(db_done[layer + 1])       # dZ_{layer-1} reuses the buffer of dZ_{layer+1}
                check(lib.wx_wide_backward_layer(h, d, ws, b, layer, dropout, 4, g, mst))      # dZ below
                after_main()
            adam(layer)                                        # rewrites the bf16 weights the dZ product has just read
        main.wait_stream(side)

    def step(self, x, y, *, dropout: int = DROPOUT_PHILOX) -> torch.Tensor:
        """One optimizer step; returns the device scalar train_loss (rank mean)."""
        self._run(x, y, dropout, None, None, reduce=self.world > 1, adam_step=self.steps_done + 1)
        self.steps_done += 1
        return self.grads[self.P:self.P + 1]

    def evaluate(self, x, y):
        """validation_step (:73-85) for this rank's rows: eval-mode forward -> (val_loss, val_acc), rank-weighted means."""
        b = self._check_batch(x, y)
        out = torch.zeros(2, device=self.device)
        check(self.lib.wx_wide_eval(self._h, C.byref(self.desc), self.params.data_ptr(), self.ws.data_ptr(), x.data_ptr(),
                                    y.data_ptr(), b, out.data_ptr(), self._stream()))
        tot = torch.cat([out.double(), torch.tensor([float(b)], dtype=torch.float64, device=self.device)])
        if self.world > 1:
            self.dist.all_reduce(tot)
        tot = tot.cpu()
        return float(tot[0] / tot[2]), float(tot[1] / tot[2])

    def step_host(self, x_host: torch.Tensor, y_host: torch.Tensor) -> float:
        """End-to-end step from HOST buffers (pinned memory recommended): H2D of the fp32 [batch, F] features and int64
        labels, the step, and a D2H read of the loss."""
        b = x_host.shape[0]
        if getattr(self, "_stage", None) is None or self._stage[0].shape[0] != b:
            self._stage = (torch.empty(b, self.F, dtype=torch.float32, device=self.device),
                           torch.empty(b, dtype=torch.int64, device=self.device), torch.zeros(1).pin_memory())
        dx, dy, hl = self._stage
        dx.copy_(x_host, non_blocking=True)
        dy.copy_(y_host, non_blocking=True)
        loss = self.step(dx, dy)
        hl.copy_(loss, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return float(hl.item())

    def gemm(self, a: torch.Tensor, b: torch.Tensor, a_kmajor: bool, b_kmajor: bool, bias=None) -> torch.Tensor:
        """Test hook for wx_wide_gemm: bf16 row-major operands, fp32 result [M, N]."""
        m, k = (a.shape if a_kmajor else (a.shape[1], a.shape[0]))
        n = b.shape[0] if b_kmajor else b.shape[1]
        out = torch.empty(m, n, dtype=torch.float32, device=self.device)
        check(self.lib.wx_wide_gemm(self._h, int(a_kmajor), int(b_kmajor), a.data_ptr(), a.stride(0), b.data_ptr(),
                                    b.stride(0), m, n, k, bias.data_ptr() if bias is not None else None, out.data_ptr(),
                                    self._stream()))
        return out
