"""numpy Philox4x32-10 and the dropout-stream contract of WX_DROPOUT_PHILOX (oracle, test infrastructure only).

No counterpart in the reference: torch's CPU dropout draws from the mt19937 global generator
(train_lightning_ddp.py:60), which a massively parallel kernel cannot replay; parity with the
reference uses WX_DROPOUT_MASK instead.  This file pins the kernels' own counter-based stream
(csrc/mlp_step.cuh) so the perf-mode dropout is testable bit-for-bit.
Known-answer test (Random123 kat_vectors, philox4x32-10): ctr=0,key=0 -> 6627e8d5 e169c58d bc57ac4c 9b00dbd8.
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
KEEP_THRESH = 52429        # 0xCCCD; P(keep) = 52429/65536


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
    """[batch, 64] bool keep mask the kernels use for (seed, step): batch position b, hidden unit j.

    For q = 0..7: counter (b_lo, b_hi, step_lo, (step_hi << 3) | q) -> words W[4q..4q+3] (calls 0..3 serve units
    0..31, calls 4..7 units 32..63); digit plane U_d = W[d] | W[16 + d] << 32 (d = 0 most significant); bit j of U_d
    is binary digit d of a 16-bit uniform u_j; keep_j <=> u_j < 0xCCCD.
    """
    b = np.arange(batch, dtype=np.uint64)
    b_lo = (b & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    b_hi = (b >> np.uint64(32)).astype(np.uint32)
    words = np.empty((32, batch), dtype=np.uint32)
    for q in range(8):
        c3 = np.uint32((((step >> 32) << 3) | q) & 0xFFFFFFFF)
        out = philox4x32_10(b_lo, b_hi, np.uint32(step & 0xFFFFFFFF), c3, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        for i in range(4):
            words[4 * q + i] = out[i]
    u = np.zeros((batch, 64), dtype=np.uint32)                      # the 16-bit uniform of every (row, unit)
    for d in range(16):
        plane = words[d].astype(np.uint64) | (words[16 + d].astype(np.uint64) << np.uint64(32))
        bits = ((plane[:, None] >> np.arange(64, dtype=np.uint64)) & np.uint64(1)).astype(np.uint32)
        u |= bits << np.uint32(15 - d)
    return u < KEEP_THRESH
