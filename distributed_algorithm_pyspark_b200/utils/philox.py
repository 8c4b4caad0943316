"""Philox4x32-10 counter-based RNG, host restatement (NumPy and torch).

The device kernels (``csrc/philox.cuh``) use the same function, so a Bernoulli row
mask or a Monte-Carlo sample can be reproduced bit-for-bit on the host for any
(seed, iteration, index).  The reference samples with CPython's Mersenne Twister
(``optimization/ssgd.py:97``, ``randomized_algorithm/monte_carlo.py:18-19``), which is
sequential and cannot be evaluated per row on a GPU; Philox replaces it in the
benchmark configurations, MT19937 masks remain available for parity runs.
"""
from __future__ import annotations

import numpy as np

PHILOX_M0 = 0xD2511F53
PHILOX_M1 = 0xCD9E8D57
PHILOX_W0 = 0x9E3779B9
PHILOX_W1 = 0xBB67AE85
_MASK = 0xFFFFFFFF


def philox4x32_10_np(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised over the counter words (uint64 arrays holding 32-bit values)."""
    c0 = np.asarray(c0, dtype=np.uint64) & _MASK
    c1 = np.broadcast_to(np.asarray(c1, dtype=np.uint64) & _MASK, c0.shape)
    c2 = np.broadcast_to(np.asarray(c2, dtype=np.uint64) & _MASK, c0.shape)
    c3 = np.broadcast_to(np.asarray(c3, dtype=np.uint64) & _MASK, c0.shape)
    k0 = np.uint64(k0 & _MASK)
    k1 = np.uint64(k1 & _MASK)
    m0, m1 = np.uint64(PHILOX_M0), np.uint64(PHILOX_M1)
    mask = np.uint64(_MASK)
    sh = np.uint64(32)
    for _ in range(10):
        p0 = m0 * c0
        p1 = m1 * c2
        hi0, lo0 = p0 >> sh, p0 & mask
        hi1, lo1 = p1 >> sh, p1 & mask
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & mask, lo1, (hi0 ^ c3 ^ k1) & mask, lo0
        k0 = (k0 + np.uint64(PHILOX_W0)) & mask
        k1 = (k1 + np.uint64(PHILOX_W1)) & mask
    return c0, c1, c2, c3


def bernoulli_threshold(fraction: float) -> int:
    """Rows are kept iff ``(philox_word0 >> 8) < threshold`` (24-bit compare, exact on every backend)."""
    return int(min(max(fraction, 0.0), 1.0) * (1 << 24))


def philox_row_mask_np(n_rows: int, fraction: float, seed: int, iteration: int,
                       row_offset: int = 0) -> np.ndarray:
    """Keep-mask for global rows [row_offset, row_offset+n_rows) at ``iteration``.
    counter = (row_lo, row_hi, iteration, 0x5347 'SG'), key = (seed_lo, seed_hi)."""
    rows = np.arange(row_offset, row_offset + n_rows, dtype=np.uint64)
    x0, _, _, _ = philox4x32_10_np(rows & _MASK, rows >> np.uint64(32), iteration, 0x5347,
                                   seed & _MASK, (seed >> 32) & _MASK)
    return (x0 >> np.uint64(8)) < np.uint64(bernoulli_threshold(fraction))


def philox_row_mask_torch(n_rows: int, fraction: float, seed: int, iteration: int,
                          row_offset: int = 0, device="cpu"):
    """Same mask with torch int64 arithmetic (used by the NCCL baseline on device)."""
    import torch
    M = 0xFFFFFFFF
    rows = torch.arange(row_offset, row_offset + n_rows, dtype=torch.int64, device=device)
    c0, c1 = rows & M, (rows >> 32) & M
    c2 = torch.full_like(rows, iteration & M)
    c3 = torch.full_like(rows, 0x5347)
    k0, k1 = seed & M, (seed >> 32) & M

    def mulhilo(a: int, b):
        # a < 2^32, b < 2^32 : split b to stay inside signed 64-bit
        b_lo, b_hi = b & 0xFFFF, b >> 16
        t_lo = a * b_lo                   # < 2^48
        t_hi = a * b_hi                   # < 2^48
        full_lo = (t_lo + ((t_hi & 0xFFFF) << 16))
        lo = full_lo & M
        hi = ((t_hi >> 16) + (full_lo >> 32)) & M
        return hi, lo

    for _ in range(10):
        hi0, lo0 = mulhilo(PHILOX_M0, c0)
        hi1, lo1 = mulhilo(PHILOX_M1, c2)
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & M, lo1, (hi0 ^ c3 ^ k1) & M, lo0
        k0 = (k0 + PHILOX_W0) & M
        k1 = (k1 + PHILOX_W1) & M
    return (c0 >> 8) < bernoulli_threshold(fraction)


def philox_uniform_pairs_np(n: int, seed: int, offset: int = 0):
    """Two (x, y) points in [-1, 1)^2 per counter, as the Monte-Carlo kernel draws them
    (``csrc/monte_carlo.cu``): u = word * 2^-32, coordinate = 2u - 1."""
    idx = np.arange(offset, offset + n, dtype=np.uint64)
    x0, x1, x2, x3 = philox4x32_10_np(idx & _MASK, idx >> np.uint64(32), 0, 0x4D43,
                                      seed & _MASK, (seed >> 32) & _MASK)
    f = lambda v: v.astype(np.float64) * (2.0 ** -31) - 1.0
    return f(x0), f(x1), f(x2), f(x3)
