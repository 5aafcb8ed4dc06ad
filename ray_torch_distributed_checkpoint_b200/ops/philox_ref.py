"""NumPy restatement of the dropout mask the GEMM epilogue draws (csrc/ptx.cuh philox4x32_10 +
csrc/gemm_tf32.cu rand16), used by the tests to check the fused dropout bit for bit and by CPU
code that needs to replay a mask."""
from __future__ import annotations

import numpy as np

_M0, _M1, _W0, _W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
_MASK = 0xFFFFFFFF


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & _MASK for c in (c0, c1, c2, c3))
    k0 = np.uint64(k0 & _MASK)
    k1 = np.uint64(k1 & _MASK)
    for _ in range(10):
        p0 = np.uint64(_M0) * c0
        p1 = np.uint64(_M1) * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & np.uint64(_MASK)
        hi1, lo1 = p1 >> np.uint64(32), p1 & np.uint64(_MASK)
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0 = (k0 + np.uint64(_W0)) & np.uint64(_MASK)
        k1 = (k1 + np.uint64(_W1)) & np.uint64(_MASK)
    return c0, c1, c2, c3


def dropout_keep_mask(rows: int, cols: int, p: float, seed: int, step: int = 0, salt: int = 0) -> np.ndarray:
    """Boolean [rows, cols] keep mask: element e = row*cols + col uses 16 bits (e % 8) of the Philox
    block with counter (e // 8, step, salt) and key = seed."""
    thr = int(p * 65536.0 + 0.5)
    thr = min(max(thr, 1), 65535) if p > 0 else 0
    e = np.arange(rows * cols, dtype=np.uint64)
    grp = e >> np.uint64(3)
    x, y, z, w = philox4x32_10(grp & np.uint64(_MASK), grp >> np.uint64(32), np.full_like(grp, step & _MASK),
                               np.full_like(grp, salt & _MASK), seed & _MASK, (seed >> 32) & _MASK)
    j = (e & np.uint64(7)).astype(np.int64)
    words = np.stack([x, y, z, w], axis=0)
    word = words[j // 2, np.arange(e.size)]
    r16 = np.where(j % 2 == 1, word >> np.uint64(16), word & np.uint64(0xFFFF))
    return (r16 >= np.uint64(thr)).reshape(rows, cols)
