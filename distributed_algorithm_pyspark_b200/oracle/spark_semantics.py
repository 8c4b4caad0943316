"""Exact host-side restatement of the PySpark RDD rules the reference relies on.

Nothing here touches a GPU.  The reference (``/root/reference``) never states
these rules -- they are Spark internals -- but its
results depend on them: which rows fall in which partition
(``optimization/ssgd.py:86``), which rows a ``sample(False, f, seed)`` keeps
(``optimization/ssgd.py:97``), and which rows ``takeSample`` returns
(``machine_learning/k-means.py:53``).
"""
from __future__ import annotations

import math
import random
import sys
from typing import List, Sequence, Tuple

import numpy as np


# --------------------------------------------------------------------------- A.1
def parallelize_slices(n_items: int, num_slices: int) -> List[Tuple[int, int]]:
    """[start, end) row ranges of ``sc.parallelize(collection, num_slices)``.

    PySpark pickles the collection in batches of
    ``max(1, min(len // num_slices, 1024))`` and the JVM then slices the
    *batches* with ``start = i*B // n``.  398 rows / 4 slices -> 99,99,99,101.
    """
    if num_slices < 1:
        raise ValueError("num_slices must be >= 1")
    if n_items == 0:
        return [(0, 0)] * num_slices
    batch = max(1, min(n_items // num_slices, 1024))
    n_batches = (n_items + batch - 1) // batch
    out = []
    for i in range(num_slices):
        b0 = (i * n_batches) // num_slices
        b1 = ((i + 1) * n_batches) // num_slices
        out.append((min(b0 * batch, n_items), min(b1 * batch, n_items)))
    return out


def parallelize_range_slices(n: int, num_slices: int) -> List[Tuple[int, int]]:
    """Slices of ``sc.parallelize(range(n), num_slices)`` (``monte_carlo.py:28``,
    ``matrix_decomposition.py:52``): partition i = [int(i*n/s), int((i+1)*n/s))."""
    return [(int(i * n / num_slices), int((i + 1) * n / num_slices))
            for i in range(num_slices)]


def partition_offsets(n_items: int, num_slices: int) -> np.ndarray:
    """Partition boundaries as an int64 array of length num_slices+1."""
    sl = parallelize_slices(n_items, num_slices)
    return np.array([sl[0][0]] + [e for _, e in sl], dtype=np.int64)


# --------------------------------------------------------------------------- A.2
def bernoulli_sample_mask(part_sizes: Sequence[int], fraction: float, seed: int) -> np.ndarray:
    """Boolean keep-mask of ``rdd.sample(False, fraction, seed)`` over all rows.

    Per partition ``split``: ``random.Random(seed ^ split)``, ten warm-up
    ``randint(0, 1)`` calls, then one ``random()`` per element in order; an
    element is kept iff the draw is ``< fraction``.
    """
    total = int(sum(part_sizes))
    mask = np.zeros(total, dtype=bool)
    pos = 0
    for split, size in enumerate(part_sizes):
        rng = random.Random(seed ^ split)
        for _ in range(10):
            rng.randint(0, 1)
        draw = rng.random
        for i in range(size):
            mask[pos + i] = draw() < fraction
        pos += size
    return mask


def bernoulli_sample_masks(part_sizes: Sequence[int], fraction: float, seed0: int,
                           n_iterations: int) -> np.ndarray:
    """Masks for seeds ``seed0 + t`` (``ssgd.py:97`` uses ``42 + t``): bool [T, N]."""
    return np.stack([bernoulli_sample_mask(part_sizes, fraction, seed0 + t)
                     for t in range(n_iterations)])


# --------------------------------------------------------------------------- A.5
def _fraction_for_sample_size(num: int, total: int, with_replacement: bool = False) -> float:
    fraction = float(num) / total
    if with_replacement:
        num_std = 9 if num < 12 else 5
        return fraction + num_std * math.sqrt(fraction / total)
    delta = 0.00005
    gamma = -math.log(delta) / total
    return min(1.0, fraction + gamma + math.sqrt(gamma * gamma + 2 * gamma * fraction))


def take_sample_indices(part_sizes: Sequence[int], num: int, seed: int) -> List[int]:
    """Row indices returned by ``rdd.takeSample(False, num, seed)`` (``k-means.py:53``)."""
    total = int(sum(part_sizes))
    rand = random.Random(seed)
    if num == 0 or total == 0:
        return []
    if num >= total:
        idx = list(range(total))
        rand.shuffle(idx)
        return idx
    fraction = _fraction_for_sample_size(num, total)
    idx = list(np.flatnonzero(bernoulli_sample_mask(part_sizes, fraction, seed)))
    while len(idx) < num:
        seed = rand.randint(0, sys.maxsize)
        idx = list(np.flatnonzero(bernoulli_sample_mask(part_sizes, fraction, seed)))
    rand.shuffle(idx)
    return [int(i) for i in idx[:num]]


# --------------------------------------------------------------------------- A.3 / A.6
def left_fold_sum(rows: np.ndarray, zero=0.0):
    """Spark's per-partition ``seqOp`` fold: ((zero + r0) + r1) + ...  (order matters in fp)."""
    acc = zero
    for r in rows:
        acc = acc + r
    return acc
