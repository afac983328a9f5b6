"""numpy Philox4x32-10 and the dropout-stream contract of WX_DROPOUT_PHILOX (oracle, test infrastructure only).

No counterpart in the reference: torch's CPU dropout draws from the mt19937 global generator
(train_lightning_ddp.py:60), which a massively parallel kernel cannot replay; parity with the
reference uses WX_DROPOUT_MASK instead.  This file pins the kernel's own counter-based stream
(csrc/mlp_step.cuh) so the perf-mode dropout is testable bit-for-bit.
Known-answer test (Random123 kat_vectors, philox4x32-10): ctr=0,key=0 -> 6627e8d5 e169c58d bc57ac4c 9b00dbd8.
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
KEEP_THRESH = 52429


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint32 arrays (broadcastable).  Returns four uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(v, dtype=np.uint32) for v in (c0, c1, c2, c3))
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = c0.astype(np.uint64) * M0
        p1 = c2.astype(np.uint64) * M1
        hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
        hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
        c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint32(k0), lo1, hi0 ^ c3 ^ np.uint32(k1), lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def dropout_keep_mask(batch: int, seed: int, step: int) -> np.ndarray:
    """[batch, 64] bool keep mask the kernel uses for (seed, step): batch position b, hidden unit j.

    counter = (b >> 2, j & 31, step_lo, step_hi ^ ((b >> 2) >> 32)); word b & 3; low half -> j < 32, high -> j >= 32.
    """
    b = np.arange(batch, dtype=np.uint64)
    grp = b >> np.uint64(2)
    lane = np.arange(32, dtype=np.uint32)
    c0 = (grp & np.uint64(0xFFFFFFFF)).astype(np.uint32)[:, None]
    c3 = (np.uint32((step >> 32) & 0xFFFFFFFF) ^ (grp >> np.uint64(32)).astype(np.uint32))[:, None]
    out = philox4x32_10(c0, lane[None, :], np.uint32(step & 0xFFFFFFFF), c3, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    words = np.stack(out, axis=0)                        # [4, batch, 32]
    w = words[(b & np.uint64(3)).astype(np.int64), np.arange(batch)]   # [batch, 32]
    keep = np.empty((batch, 64), dtype=bool)
    keep[:, :32] = (w & np.uint32(0xFFFF)) < KEEP_THRESH
    keep[:, 32:] = (w >> np.uint32(16)) < KEEP_THRESH
    return keep
