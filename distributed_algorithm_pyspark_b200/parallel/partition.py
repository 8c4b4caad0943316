"""Logical partitions (Spark ``n_slices``) decoupled from physical GPUs.

The reference's results depend on ``n_slices`` (partition-wise sampling streams, per-partition
means in ``optimization/ma.py:39-43``, one replica per partition).  Here a *logical* partition
``p`` always contains the rows Spark would give it (``oracle.spark_semantics.parallelize_slices``)
and is placed on GPU ``p * world // n_slices``; results are therefore independent of the GPU count
(SURVEY.md section 7.1 L3).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

import numpy as np

from ..oracle.spark_semantics import partition_offsets


def owner_rank(p: int, n_slices: int, world: int) -> int:
    return (p * world) // n_slices if world <= n_slices else p


def partitions_of_rank(rank: int, n_slices: int, world: int) -> List[int]:
    return [p for p in range(n_slices) if owner_rank(p, n_slices, world) == rank]


@dataclass
class RowSharding:
    """Row ranges of a logically partitioned [N, ...] array and their owning ranks."""
    n_rows: int
    n_slices: int
    world: int
    offsets: np.ndarray            # int64 [n_slices + 1]

    @staticmethod
    def spark(n_rows: int, n_slices: int, world: int) -> "RowSharding":
        return RowSharding(n_rows, n_slices, world, partition_offsets(n_rows, n_slices))

    @staticmethod
    def even(n_rows: int, n_slices: int, world: int) -> "RowSharding":
        offs = np.array([(i * n_rows) // n_slices for i in range(n_slices + 1)], dtype=np.int64)
        return RowSharding(n_rows, n_slices, world, offs)

    def local_partitions(self, rank: int) -> List[int]:
        return partitions_of_rank(rank, self.n_slices, self.world)

    def local_row_range(self, rank: int) -> Tuple[int, int]:
        ps = self.local_partitions(rank)
        if not ps:
            return (0, 0)
        return int(self.offsets[ps[0]]), int(self.offsets[ps[-1] + 1])

    def local_offsets(self, rank: int) -> np.ndarray:
        """Partition boundaries of this rank's partitions, relative to its first row."""
        ps = self.local_partitions(rank)
        if not ps:
            return np.zeros(1, dtype=np.int64)
        base = self.offsets[ps[0]]
        return (self.offsets[ps[0]:ps[-1] + 2] - base).astype(np.int64)

    def sizes(self) -> np.ndarray:
        return np.diff(self.offsets)


def vertex_range(n_vertices: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous dst-vertex range owned by ``rank`` (PageRank dst-range CSR)."""
    return (rank * n_vertices) // world, ((rank + 1) * n_vertices) // world


def hash_owner_u64(keys, world: int):
    """Owner rank of a packed 64-bit key (transitive-closure all-to-all).
    splitmix64 finaliser; identical to ``dapb_hash64`` in csrc/common.cuh."""
    import torch
    x = keys.to(torch.int64)
    # arithmetic on int64 with wraparound == uint64 arithmetic modulo 2^64
    x = (x ^ ((x >> 30) & 0x3FFFFFFFF)) * -0x40A7B892E31B1A47            # 0xBF58476D1CE4E5B9
    x = (x ^ ((x >> 27) & 0x1FFFFFFFFF)) * -0x6B2FB644ECCEEE15           # 0x94D049BB133111EB
    x = x ^ ((x >> 31) & 0x1FFFFFFFF)
    return ((x >> 33) & 0x7FFFFFFF) % world
