"""Transitive closure by repeated join (reference: ``graph_computation/transitive_closure.py``).

The reference joins all current paths with the (reversed) edges each round, unions, de-duplicates and
stops when the pair count no longer grows (:29-40); the printed result is the number of closure pairs
(:42).  Here each round joins only the pairs found in the previous round (same closure, same number of
rounds), de-duplicates against a device hash set of packed 64-bit ``(x, z)`` keys and, on several GPUs,
routes every candidate pair to its hash owner through queues in peer memory (csrc/transitive_closure.cu).

Backends: ``oracle`` (Python sets), ``torch`` (sort/unique based, ``all_to_all`` -- the NCCL baseline),
``fused`` (hash-join kernels).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Iterable, Optional, Tuple

import numpy as np
import torch

from .. import _native
from ..config import TransitiveClosureConfig
from ..oracle.algorithms import transitive_closure_oracle
from ..parallel.dist import Dist, current
from ..parallel.partition import hash_owner_u64


@dataclass
class ClosureResult:
    n_pairs: int
    rounds: int
    pairs: Optional[np.ndarray] = None      # [n_pairs, 2] original vertex ids (small problems only)


def _pack(x: torch.Tensor, z: torch.Tensor) -> torch.Tensor:
    return (x.to(torch.int64) << 32) | z.to(torch.int64)


def _in_csr(src: torch.Tensor, dst: torch.Tensor, V: int):
    """CSR over the join key y = dst: in-neighbours x of y (ref transitive_closure.py:27 reversed edges)."""
    order = torch.argsort(dst, stable=True)
    indeg = torch.bincount(dst, minlength=V)
    ptr = torch.zeros(V + 1, dtype=torch.int64, device=src.device)
    ptr[1:] = torch.cumsum(indeg, 0)
    return ptr, src[order].to(torch.int32).contiguous(), indeg


def closure_device(src: torch.Tensor, dst: torch.Tensor, V: int, backend: str = "fused",
                   dist: Optional[Dist] = None, max_rounds: int = 1 << 20, keep_pairs: bool = False,
                   heap=None, max_pairs: int = 1 << 28, persistent: Optional[bool] = None
                   ) -> Tuple[int, int, Optional[torch.Tensor]]:
    """Closure of a graph with vertex ids 0..V-1 (the FULL edge list on every rank, int64 on device).
    Returns (global pair count, rounds, local keys or None)."""
    dist = dist or current()
    dev = src.device
    if persistent is None:
        persistent = os.environ.get("DAPB_TC_ROUNDS") is None
    ekeys = torch.unique(_pack(src, dst))
    src_u, dst_u = (ekeys >> 32), (ekeys & 0xFFFFFFFF)
    in_ptr, in_src, indeg = _in_csr(src_u, dst_u, V)
    W, me = dist.world, dist.rank

    if backend != "fused":
        # ---------------- torch / NCCL baseline: sort-unique sets, all_to_all by hash owner
        own = hash_owner_u64(ekeys, W) == me if W > 1 else torch.ones_like(ekeys, dtype=torch.bool)
        known = ekeys[own]
        frontier = known
        total = int(dist.sum_over_ranks(known.numel()))
        rounds = 0
        while rounds < max_rounds:
            old = total
            y, z = frontier >> 32, frontier & 0xFFFFFFFF
            deg = indeg[y]
            rep = torch.repeat_interleave(torch.arange(frontier.numel(), device=dev), deg)
            start = in_ptr[y]
            offs = torch.arange(rep.numel(), device=dev) - torch.repeat_interleave(torch.cumsum(deg, 0) - deg, deg)
            x = in_src[(start[rep] + offs)].to(torch.int64)
            cand = torch.unique((x << 32) | z[rep])
            if W > 1:
                o = hash_owner_u64(cand, W)
                recv = dist.all_to_all([cand[o == r] for r in range(W)])
                cand = torch.unique(torch.cat(recv))
            pos = torch.searchsorted(known, cand)
            pos = pos.clamp(max=max(known.numel() - 1, 0))
            is_new = known[pos] != cand if known.numel() else torch.ones_like(cand, dtype=torch.bool)
            frontier = cand[is_new]
            known = torch.sort(torch.cat([known, frontier])).values
            total = int(dist.sum_over_ranks(known.numel()))
            rounds += 1
            if total == old:
                break
        return total, rounds, known if keep_pairs else None

    # ---------------- fused hash-join kernels
    from ..parallel.symm import default_heap
    lib = _native.load()
    heap = heap or default_heap()
    st = _native.stream_ptr()
    u64 = torch.int64                                            # bit patterns of uint64 keys
    cap = 1 << max(10, int(np.ceil(np.log2(max(4 * ekeys.numel() // max(W, 1), 1024)))))
    table = torch.full((cap,), -1, dtype=u64, device=dev)
    counters = torch.zeros(1 + _native.MAX_RANKS, dtype=u64, device=dev)
    frontier = torch.empty(_pow2(ekeys.numel()) if persistent else max(ekeys.numel(), 1), dtype=u64, device=dev)
    _native.check(lib.dapb_tc_seed(ekeys.data_ptr(), ekeys.numel(), table.data_ptr(), cap, frontier.data_ptr(),
                                   counters.data_ptr(), W, me, st), "tc_seed")
    if persistent:
        return _closure_persistent(lib, heap, dist, in_ptr, in_src, ekeys, table, cap, frontier, counters, max_rounds,
                                   max_pairs, keep_pairs)
    n_front = int(counters[0].item())
    n_known = n_front
    total = int(dist.sum_over_ranks(n_known))
    rounds = 0
    region = None
    stat = torch.zeros(2, dtype=torch.int64, device=dev)
    while rounds < max_rounds:
        old = total
        fr = frontier[:n_front]
        # one host sync + one tiny collective: how many candidates can this round emit (locally / at most per rank)?
        stat[0] = indeg[(fr >> 32)].sum() if n_front else 0
        stat[1] = stat[0]
        if W > 1:
            torch.distributed.all_reduce(stat[1:], op=torch.distributed.ReduceOp.MAX)
        bound_local, bound_max = (int(v) for v in stat.tolist())
        bound_new = bound_local if W == 1 else bound_max * W          # worst case: every rank sends everything to me
        if n_known + bound_new > max_pairs:
            raise MemoryError(f"transitive closure would exceed max_pairs={max_pairs}")
        need = 2 * (n_known + bound_new)
        if need > cap:                                            # grow + rehash (keys re-seeded, no frontier output)
            new_cap = 1 << int(np.ceil(np.log2(need)))
            old_keys = table[table != -1]
            table = torch.full((new_cap,), -1, dtype=u64, device=dev)
            scratch = torch.empty(max(old_keys.numel(), 1), dtype=u64, device=dev)
            cnt = torch.zeros(1, dtype=u64, device=dev)
            _native.check(lib.dapb_tc_seed(old_keys.data_ptr(), old_keys.numel(), table.data_ptr(), new_cap,
                                           scratch.data_ptr(), cnt.data_ptr(), 1, 0, st), "tc_seed(rehash)")
            cap = new_cap
        nxt = torch.empty(max(bound_new, 1), dtype=u64, device=dev)
        counters.zero_()
        q_off, q_cap = 0, 0
        if W > 1:
            q_cap = 1 << max(12, int(np.ceil(np.log2(max(bound_max, 1)))))     # power of two: few distinct regions
            region = heap.region(f"tc_queues_{q_cap}", W * (1 + q_cap) * 8, _native.MAX_RANKS)
            q_off = region.payload_off
        _native.check(lib.dapb_tc_expand(heap.comm_ref(), in_ptr.data_ptr(), in_src.data_ptr(), fr.data_ptr(), n_front,
                                         table.data_ptr(), cap, nxt.data_ptr(), counters.data_ptr(), q_off, q_cap, st),
                      "tc_expand")
        if W > 1:
            _native.check(lib.dapb_tc_exchange(heap.comm_ref(), counters.data_ptr(), q_off, q_cap, region.flag_base,
                                               region.next_epoch(), st), "tc_exchange")
            _native.check(lib.dapb_tc_drain(heap.comm_ref(), table.data_ptr(), cap, nxt.data_ptr(), counters.data_ptr(),
                                            q_off, q_cap, st), "tc_drain")
        # second (and last) host sync of the round: local and global number of new pairs; it also orders this
        # round's drain before the next round's pushes (ref :38-40: the termination test must agree on all ranks)
        stat[0] = counters[0]
        stat[1] = counters[0]
        if W > 1:
            torch.distributed.all_reduce(stat[1:], op=torch.distributed.ReduceOp.SUM)
        n_front, new_global = (int(v) for v in stat.tolist())
        frontier = nxt
        n_known += n_front
        total += new_global
        rounds += 1
        if total == old:
            break
    keys = table[table != -1] if keep_pairs else None
    return total, rounds, keys


def _pow2(n: int, lo: int = 4096) -> int:
    return 1 << max(int(np.ceil(np.log2(max(n, 1)))), int(np.log2(lo)))


def _closure_persistent(lib, heap, dist, in_ptr, in_src, ekeys, table, cap, frontier0, counters, max_rounds,
                        max_pairs, keep_pairs):
    """All rounds in one cooperative launch (``dapb_tc_closure``): the kernel returns only when the closure is complete
    or a buffer must grow.  Buffer growth is decided from the same in-kernel message matrix on every rank, so the
    host loop needs no collective of its own."""
    dev, W = in_ptr.device, dist.world
    st = _native.stream_ptr()
    u64 = torch.int64
    n_e = ekeys.numel()
    state = torch.zeros(_native.TC_STATE_WORDS, dtype=u64, device=dev)
    state[0] = counters[0]
    state[3] = counters[0]
    front_cap = frontier0.numel()                       # >= number of seeds on any rank
    fronts = [frontier0, torch.empty_like(frontier0)]
    q_cap = _pow2(n_e // W)
    msg = heap.region("tc_msg", 2 * _native.MAX_RANKS * 16 * 8, _native.MAX_RANKS)
    torch.cuda.synchronize()
    dist.barrier()                       # nobody may still be using the queues / messages of a previous closure
    while True:
        q = heap.region(f"tc_pq_{q_cap}", 2 * W * q_cap * 8, 0)
        state[7] = msg.epoch
        _native.check(lib.dapb_tc_closure(heap.comm_ref(), in_ptr.data_ptr(), in_src.data_ptr(), table.data_ptr(), cap,
                                          fronts[0].data_ptr(), fronts[1].data_ptr(), front_cap, state.data_ptr(),
                                          q.payload_off, q_cap, msg.payload_off, msg.flag_base, max_rounds, st),
                      "tc_closure")
        h = state.tolist()
        msg.epoch = h[7]
        if h[6] == _native.TC_DONE:
            break
        assert h[6] == _native.TC_GROW, f"tc_closure: unexpected status {h[6]}"
        need_q, need_table, need_front = h[8], h[9], h[10]
        if need_table // 2 > max_pairs:
            raise MemoryError(f"transitive closure would exceed max_pairs={max_pairs}")
        if need_q > q_cap:
            q_cap = _pow2(2 * need_q)
        if need_front > front_cap:
            new_cap = _pow2(2 * need_front)
            grown = [torch.empty(new_cap, dtype=u64, device=dev) for _ in range(2)]
            cur = h[2]
            grown[cur][:h[cur]] = fronts[cur][:h[cur]]
            fronts, front_cap = grown, new_cap
        if need_table > cap:
            new_cap = _pow2(2 * need_table)
            old_keys = table[table != -1]
            table = torch.full((new_cap,), -1, dtype=u64, device=dev)
            scratch = torch.empty(max(old_keys.numel(), 1), dtype=u64, device=dev)
            cnt = torch.zeros(1, dtype=u64, device=dev)
            _native.check(lib.dapb_tc_seed(old_keys.data_ptr(), old_keys.numel(), table.data_ptr(), new_cap,
                                           scratch.data_ptr(), cnt.data_ptr(), 1, 0, st), "tc_seed(rehash)")
            cap = new_cap
    keys = table[table != -1] if keep_pairs else None
    return int(h[4]), int(h[5]), keys


def run_transitive_closure(edges: Iterable[Tuple[int, int]], cfg: Optional[TransitiveClosureConfig] = None,
                           backend: str = "auto", dist: Optional[Dist] = None) -> ClosureResult:
    """Closure of an edge list with arbitrary integer vertex ids (ref script semantics)."""
    cfg = cfg or TransitiveClosureConfig()
    if backend == "auto":
        backend = "fused" if _native.available() else "torch"
    edges = [(int(a), int(b)) for a, b in edges]
    if backend == "oracle":
        n, r, pairs = transitive_closure_oracle(edges)
        return ClosureResult(n, r, np.array(sorted(pairs), dtype=np.int64).reshape(-1, 2))
    dist = dist or current()
    ids = sorted(set(v for e in edges for v in e))
    remap = {v: i for i, v in enumerate(ids)}
    src = torch.tensor([remap[a] for a, _ in edges], dtype=torch.int64, device=dist.device)
    dst = torch.tensor([remap[b] for _, b in edges], dtype=torch.int64, device=dist.device)
    total, rounds, keys = closure_device(src, dst, len(ids), backend, dist, cfg.max_rounds, keep_pairs=True)
    parts = dist.all_gather_object(keys.cpu().numpy())
    allk = np.unique(np.concatenate(parts))
    idarr = np.array(ids, dtype=np.int64)
    pairs = np.stack([idarr[allk >> 32], idarr[allk & 0xFFFFFFFF]], axis=1)
    return ClosureResult(total, rounds, pairs)
