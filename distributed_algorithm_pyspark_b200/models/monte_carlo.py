"""Monte-Carlo estimate of pi (reference: ``randomized_algorithm/monte_carlo.py``).

``n`` points uniform in [-1, 1)^2, accepted iff ``x^2 + y^2 <= 1`` (:17-20), counts summed over the
partitions (:28), ``pi ~ 4 * count / n`` (:31).  Sample ``i`` is a pure function of ``(seed, i)``
(Philox4x32-10, two points per counter), so the sample range can be cut across any number of GPUs --
or recomputed on the host -- with identical totals.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch

from .. import _native
from ..config import MonteCarloConfig
from ..oracle.algorithms import monte_carlo_oracle
from ..parallel.dist import Dist, current
from ..utils.philox import philox4x32_10_np


def accept_count_np(first: int, n: int, seed: int, chunk: int = 1 << 20) -> int:
    """Host twin of ``dapb_mc_pi`` (exact integer acceptance test)."""
    total = 0
    c0, c1 = first >> 1, (first + n + 1) >> 1
    for a in range(c0, c1, chunk):
        b = min(c1, a + chunk)
        k = np.arange(a, b, dtype=np.uint64)
        x0, x1, x2, x3 = philox4x32_10_np(k & 0xFFFFFFFF, k >> np.uint64(32), 0, 0x4D43, seed & 0xFFFFFFFF,
                                          (seed >> 32) & 0xFFFFFFFF)

        def acc(ux, uy):
            sx = ux.astype(np.uint32).view(np.int32).astype(np.int64)
            sy = uy.astype(np.uint32).view(np.int32).astype(np.int64)
            return (sx * sx).astype(np.uint64) + (sy * sy).astype(np.uint64) <= np.uint64(1 << 62)
        s0 = 2 * k.astype(np.int64)
        in0 = (s0 >= first) & (s0 < first + n)
        in1 = (s0 + 1 >= first) & (s0 + 1 < first + n)
        total += int(np.count_nonzero(acc(x0, x1) & in0)) + int(np.count_nonzero(acc(x2, x3) & in1))
    return total


def mc_count_fused(first: int, n_local: int, seed: int, dist: Dist, heap=None, return_tensor: bool = False):
    """Accepted samples over ALL ranks: sampler + grid reduction + cross-GPU total in one launch per rank
    (csrc/monte_carlo.cu); the sum arrives through peer memory, ``torch.distributed`` is not involved."""
    lib = _native.load()
    dev = dist.device
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    if dist.world == 1:
        _native.check(lib.dapb_mc_pi(first, n_local, seed, cnt.data_ptr(), _native.stream_ptr()), "mc_pi")
        return cnt if return_tensor else int(cnt.item())
    from ..parallel.symm import default_heap
    heap = heap or default_heap()
    reg = heap.region("mc_pi_total", 2 * _native.MAX_RANKS * 8, _native.MAX_RANKS)
    total = torch.zeros(1, dtype=torch.int64, device=dev)
    ticket = torch.zeros(1, dtype=torch.int32, device=dev)
    _native.check(lib.dapb_mc_pi_allreduce(heap.comm_ref(), first, n_local, seed, cnt.data_ptr(), total.data_ptr(),
                                           ticket.data_ptr(), reg.payload_off, reg.flag_base, reg.next_epoch(),
                                           _native.stream_ptr()), "mc_pi_allreduce")
    return total if return_tensor else int(total.item())


def run_monte_carlo(cfg: MonteCarloConfig, backend: str = "auto", dist: Optional[Dist] = None) -> Tuple[float, int]:
    """Returns (pi estimate, accepted count)."""
    if backend == "auto":
        backend = "fused" if _native.available() else "torch"
    if backend == "oracle":
        return monte_carlo_oracle(cfg)
    dist = dist or current()
    seed = 0 if cfg.seed is None else int(cfg.seed)
    a = (dist.rank * cfg.n) // dist.world
    b = ((dist.rank + 1) * cfg.n) // dist.world
    if backend == "fused":
        count = mc_count_fused(a, b - a, seed, dist)
        return 4.0 * count / cfg.n, count
    else:   # torch baseline: torch.rand + all_reduce (different stream of random numbers, same estimator)
        g = torch.Generator(device=dist.device).manual_seed(seed + dist.rank)
        local = torch.zeros(1, dtype=torch.int64, device=dist.device)
        done = 0
        while done < b - a:
            m = min(1 << 24, b - a - done)
            xy = torch.rand(2, m, device=dist.device, generator=g) * 2 - 1
            local += ((xy[0] * xy[0] + xy[1] * xy[1]) <= 1).sum()
            done += m
    dist.all_reduce_sum(local)
    count = int(local.item())
    return 4.0 * count / cfg.n, count
