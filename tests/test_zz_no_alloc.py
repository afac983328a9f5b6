"""GPU test of wx_mlp_grad_static, the gradient entry point for callers that are not allowed to allocate
device memory (plugins under a framework-owned memory budget, latency-critical loops, no-allocation-after-
start policies).  It keeps its partial-sum scratch in a fixed array inside the library instead of in a
context, so this file checks that, called straight through the C ABI, it returns what wx_mlp_grad returns
-- on two calls in a row, the second of which finds the first call's leftovers in the shared array (and
uses a different model width, so the leftovers are laid out differently).

Collected last on purpose (file name): if it fails under `pytest -x`, every other test has already run.
It does not depend on them: it passes alone in a fresh process as well as after the rest of the suite.
"""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL, ATOL_G = 1e-5, 1e-7          # the tolerances of tests/test_gpu_parity.py


def _call_static(lib, f, p, x, y, batch, row0=0, row_idx=None, mode=0, mask=None, seed=0, step=0, scale=None):
    out = torch.full((64 * f + 64 + 128 + 2 + 1,), float("nan"), device="cuda")
    rc = lib.wx_mlp_grad_static(f, p.data_ptr(), x.data_ptr(), y.data_ptr(),
                                None if row_idx is None else row_idx.data_ptr(), row0, batch, mode,
                                None if mask is None else mask.data_ptr(), seed, step,
                                (1.0 / batch) if scale is None else scale, out.data_ptr(),
                                torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.wx_last_error()
    return out


def test_static_scratch_entry_matches_context_entry_twice_in_a_row():
    import wxddp

    lib = wxddp.load()
    g = torch.Generator().manual_seed(2024)
    cases = [  # (features, rows, batch, row0, dropout mode)
        (20, 70_000, 65_537, 1_001, wxddp.DROPOUT_PHILOX),    # first call: wide model, many CTAs, ragged tail
        (5, 9_000, 8_999, 1, wxddp.DROPOUT_MASK),             # second call: other width, reuses the dirty array
        (20, 70_000, 13, 0, wxddp.DROPOUT_OFF),               # third: tiny batch, single CTA
    ]
    for f, n, batch, row0, mode in cases:
        x = torch.randn(n, f, generator=g).cuda()
        y = (torch.rand(n, generator=g) < 0.3).long().cuda()
        p = (torch.randn(64 * f + 194, generator=g) * 0.2).cuda()
        mask = None
        if mode == wxddp.DROPOUT_MASK:
            keep = torch.rand(batch, 64, generator=g) < 0.8
            bits = np.packbits(keep.numpy(), axis=1, bitorder="little").view(np.int64).reshape(-1)
            mask = torch.from_numpy(bits).cuda()
        got = _call_static(lib, f, p, x, y, batch, row0=row0, mode=mode, mask=mask, seed=77, step=3)
        ctx = wxddp.Context(f, device=0)
        want = ctx.mlp_grad(p, x, y, row0=row0, batch=batch, dropout=mode, keep_mask=mask, seed=77, step=3)
        torch.cuda.synchronize()
        got, want = got.cpu().numpy(), want.cpu().numpy()
        assert np.isfinite(got).all()
        np.testing.assert_allclose(got[-1], want[-1], rtol=RTOL)
        np.testing.assert_allclose(got[:-1], want[:-1], rtol=RTOL, atol=ATOL_G)
        ctx.close()
    # the sampler-index form (row_idx) goes through the same scratch
    f, n, batch = 5, 9_000, 4_000
    x = torch.randn(n, f, generator=g).cuda()
    y = (torch.rand(n, generator=g) < 0.3).long().cuda()
    p = (torch.randn(64 * f + 194, generator=g) * 0.2).cuda()
    idx = torch.randperm(n, generator=g)[:batch].cuda()
    got = _call_static(lib, f, p, x, y, batch, row_idx=idx).cpu().numpy()
    ctx = wxddp.Context(f, device=0)
    want = ctx.mlp_grad(p, x, y, row_idx=idx).cpu().numpy()
    np.testing.assert_allclose(got[-1], want[-1], rtol=RTOL)
    np.testing.assert_allclose(got[:-1], want[:-1], rtol=RTOL, atol=ATOL_G)
    ctx.close()


@pytest.mark.parametrize("f", [1, 8, 16, 17, 32])
def test_static_scratch_entry_large_batch_every_width(f):
    """1 Mi rows at narrow and wide models: the launch must size its grid to the slabs the fixed array holds (a narrow
    model reaches several CTAs per SM, a wide one has large slabs) and still return wx_mlp_grad's sums."""
    import wxddp

    lib = wxddp.load()
    n = 1 << 20
    g = torch.Generator(device="cuda").manual_seed(5 + f)
    x = torch.randn(n, f, device="cuda", generator=g)
    y = (torch.rand(n, device="cuda", generator=g) < 0.3).long()
    p = torch.randn(64 * f + 194, device="cuda", generator=g) * 0.2
    got = _call_static(lib, f, p, x, y, n, mode=wxddp.DROPOUT_PHILOX, seed=9, step=2).cpu().numpy()
    ctx = wxddp.Context(f, device=0)
    want = ctx.mlp_grad(p, x, y, dropout=wxddp.DROPOUT_PHILOX, seed=9, step=2).cpu().numpy()
    ctx.close()
    assert np.isfinite(got).all()
    np.testing.assert_allclose(got[-1], want[-1], rtol=RTOL)
    # the two launches may use different grids (different partial-sum grouping): one part in 1e5 of the largest entry
    np.testing.assert_allclose(got[:-1], want[:-1], rtol=RTOL, atol=1e-5 * np.abs(want[:-1]).max())


def test_static_scratch_entry_rejects_requests_beyond_its_limits():
    import wxddp

    lib = wxddp.load()
    dummy = torch.zeros(8, device="cuda")
    yy = torch.zeros(8, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    too_many = (1 << 24) + 1
    rc = lib.wx_mlp_grad_static(5, dummy.data_ptr(), dummy.data_ptr(), yy.data_ptr(), None, 0, too_many, 0, None, 0, 0,
                                1.0, dummy.data_ptr(), st)
    assert rc == -1 and b"16777216" in lib.wx_last_error()
    rc = lib.wx_mlp_grad_static(33, dummy.data_ptr(), dummy.data_ptr(), yy.data_ptr(), None, 0, 8, 0, None, 0, 0,
                                1.0, dummy.data_ptr(), st)
    assert rc == -1 and b"n_features" in lib.wx_last_error()
    rc = lib.wx_mlp_grad_static(5, dummy.data_ptr(), dummy.data_ptr(), yy.data_ptr(), None, 0, 0, 0, None, 0, 0,
                                1.0, dummy.data_ptr(), st)
    assert rc == -1
