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

    Two Philox calls per batch position, q = 0, 1: counter (b_lo, b_hi, step_lo, (step_hi << 3) | q), key = seed
    -> words W[4q .. 4q+3].  Word W[k] serves hidden units 8k .. 8k+7 through its first eight base-5 digits:
        s_0 = W[k];   s_{t+1} = (5 * s_t) mod 2^32;   digit_t = floor(5 * s_t / 2^32)  in {0..4}
    and keep_{8k+t} <=> digit_t != 0, i.e. P(keep) = 1 - ceil(2^32 / 5) / 2^32 = 0.8 to within 1e-10 (the reference's
    nn.Dropout(0.2), train_lightning_ddp.py:60).  One 32x32->64 multiply per unit in the kernels.
    """
    b = np.arange(batch, dtype=np.uint64)
    b_lo = (b & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    b_hi = (b >> np.uint64(32)).astype(np.uint32)
    keep = np.empty((batch, 64), dtype=bool)
    for q in range(2):
        c3 = np.uint32((((step >> 32) << 3) | q) & 0xFFFFFFFF)
        out = philox4x32_10(b_lo, b_hi, np.uint32(step & 0xFFFFFFFF), c3, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        for i in range(4):
            s = out[i].astype(np.uint64)
            for t in range(8):
                prod = s * np.uint64(5)
                keep[:, 8 * (4 * q + i) + t] = (prod >> np.uint64(32)) != 0
                s = prod & np.uint64(0xFFFFFFFF)
    return keep
