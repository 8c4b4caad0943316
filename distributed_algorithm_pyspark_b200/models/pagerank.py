"""PageRank (reference: ``graph_computation/pagerank.py``).

Graph preparation mirrors ``links.distinct().groupByKey()`` (:41): duplicate edges are dropped,
self-loops kept; ``n`` is the number of distinct SOURCE vertices (:44).  The iteration is the vector
form of :52-57 including the reference's "absent vertex" rule.

Backends: ``oracle`` (Python floats, bit-exact with the values recorded in pagerank.py:66-68), ``torch``
(CSR SpMV via ``index_add_`` + ``all_gather`` -- the NCCL baseline), ``fused`` (csrc/pagerank.cu: SpMV +
damping + all-gather push; the default warp layout runs ALL iterations of ``run(n)`` in one persistent
cooperative launch, fp32 or fp64, contributions travelling by ``multimem.st`` when the heap has a multicast
mapping).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Dict, Iterable, Optional, Tuple

import numpy as np
import torch

from .. import _native
from ..config import PageRankConfig
from ..oracle.algorithms import pagerank_oracle
from ..parallel.dist import Dist, current
from ..parallel.partition import vertex_range
from ..utils import checkpoint as ck
from ..utils.trace import span


@dataclass
class ShardedGraph:
    """dst-range shard of the de-duplicated graph in CSR(A^T) form."""
    V: int
    n_sources: int
    n_edges: int                 # global, after de-duplication
    v0: int
    v1: int
    bounds: tuple                # vertex-range boundaries of every rank (len world + 1), nnz-balanced
    row_ptr: torch.Tensor        # int64 [nv_local + 1]
    col_idx: torch.Tensor        # int32 [nnz_local] source ids
    inv_outdeg: torch.Tensor     # float32 [nv_local]
    outdeg_local: torch.Tensor   # int64 [nv_local]
    blocks: Optional[torch.Tensor] = None    # int32 [n_blocks + 1] row boundaries of ~block_nnz non-zeros each
    block_nnz: int = 2048
    hub: Optional["HubRows"] = None          # warp-granular layout: rows too long for one warp (see split_hub_rows)


@dataclass
class HubRows:
    """Rows with more than ``threshold`` in-edges, taken out of the main CSR (csrc/pagerank.cu, warp-granular variant)."""
    rows: torch.Tensor            # int32 [n_hub] local row ids, ascending
    col: torch.Tensor             # int32 source ids of all hub rows, row after row
    chunk_start: torch.Tensor     # int64 [n_chunks] offsets into col
    chunk_end: torch.Tensor       # int64 [n_chunks]
    chunk_ptr: torch.Tensor       # int32 [n_hub + 1] chunks of hub row i
    partial: torch.Tensor         # float64-sized scratch [n_chunks] (viewed as the run's value type)
    ptr: Optional[torch.Tensor] = None   # int64 [n_hub + 1] offsets of the hub rows in col
    chunk_row: Optional[torch.Tensor] = None   # int32 [n_chunks] hub row index of each chunk
    row_done: Optional[torch.Tensor] = None    # int32 [n_hub] completion counters (zero between iterations)

    def fill(self, h: "_native.DapbPrHub"):
        h.hub_rows, h.hub_col = self.rows.data_ptr(), self.col.data_ptr()
        h.chunk_start, h.chunk_end = self.chunk_start.data_ptr(), self.chunk_end.data_ptr()
        h.hub_chunk_ptr, h.chunk_partial = self.chunk_ptr.data_ptr(), self.partial.data_ptr()
        h.n_hub, h.n_chunks = int(self.rows.numel()), int(self.chunk_start.numel())
        h.chunk_row, h.row_done = self.chunk_row.data_ptr(), self.row_done.data_ptr()


def split_hub_rows(row_ptr: torch.Tensor, col_idx: torch.Tensor, threshold: int, chunk: int = 2048):
    """Returns (row_ptr', col_idx', HubRows): the CSR without the non-zeros of rows longer than ``threshold`` (those rows
    have length 0 in row_ptr') and the description of the removed rows cut into ``chunk``-element pieces."""
    dev = row_ptr.device
    lens = row_ptr[1:] - row_ptr[:-1]
    rows = torch.nonzero(lens > threshold).flatten()
    i64, i32 = torch.int64, torch.int32
    if rows.numel() == 0:
        z = torch.zeros(0, dtype=i64, device=dev)
        return row_ptr, col_idx, HubRows(z.to(i32), z.to(i32), z, z, torch.zeros(1, dtype=i32, device=dev),
                                         torch.zeros(1, dtype=torch.float64, device=dev), torch.zeros(1, dtype=i64, device=dev),
                                         z.to(i32), torch.zeros(1, dtype=i32, device=dev))
    hs, he = row_ptr[rows], row_ptr[rows + 1]
    is_hub = torch.zeros(lens.numel(), dtype=torch.bool, device=dev)
    is_hub[rows] = True
    edge_in_hub = torch.repeat_interleave(is_hub, lens)            # [nnz] bool, rows are contiguous in the CSR
    hub_col = col_idx[edge_in_hub].contiguous()
    main_col = col_idx[~edge_in_hub].contiguous()
    del edge_in_hub
    lens_main = lens.clone()
    lens_main[rows] = 0
    rp = torch.zeros_like(row_ptr)
    rp[1:] = torch.cumsum(lens_main, 0)
    hub_len = he - hs
    hub_ptr = torch.zeros(rows.numel() + 1, dtype=i64, device=dev)
    hub_ptr[1:] = torch.cumsum(hub_len, 0)
    nch = (hub_len + chunk - 1) // chunk
    chunk_ptr = torch.zeros(rows.numel() + 1, dtype=i64, device=dev)
    chunk_ptr[1:] = torch.cumsum(nch, 0)
    owner = torch.repeat_interleave(torch.arange(rows.numel(), device=dev), nch)
    within = torch.arange(owner.numel(), device=dev) - chunk_ptr[owner]
    c_start = hub_ptr[owner] + within * chunk
    c_end = torch.minimum(c_start + chunk, hub_ptr[owner + 1])
    hub = HubRows(rows.to(i32).contiguous(), hub_col, c_start.contiguous(), c_end.contiguous(),
                  chunk_ptr.to(i32).contiguous(), torch.zeros(owner.numel(), dtype=torch.float64, device=dev), hub_ptr,
                  owner.to(i32).contiguous(), torch.zeros(rows.numel(), dtype=i32, device=dev))
    return rp, main_col, hub


def balanced_vertex_bounds(indeg_all: torch.Tensor, world: int, vertex_cost: Optional[float] = None,
                           hub_threshold: int = 64, hub_edge_cost: Optional[float] = None,
                           outdeg_all: Optional[torch.Tensor] = None, push_cost: Optional[float] = None):
    """Contiguous destination-vertex ranges of (almost) equal modelled cost per rank.  A vertex costs 1, plus
    ``vertex_cost`` if it has in-edges (row bookkeeping), plus 1 per in-edge -- ``hub_edge_cost`` per in-edge for hub
    rows (> hub_threshold in-edges), whose non-zeros are streamed by the cheaper chunk path -- plus ``push_cost`` if it
    has in-edges AND out-edges: only those vertices ever hold a non-zero contribution, so only they are pushed over
    NVLink every iteration (pass ``outdeg_all`` to enable; the caller scales ``push_cost`` by the number of copies a
    push makes: 1 with NVLS multicast, world - 1 with unicast stores).  R-MAT graphs concentrate edges on low vertex
    ids, so equal-size ranges would give rank 0 most of the work and the last rank most of the vertices
    (SURVEY.md 7.3-6).  ``DAPB_PR_VERTEX_COST`` / ``DAPB_PR_HUB_EDGE_COST`` / ``DAPB_PR_PUSH_COST`` override."""
    if vertex_cost is None:
        vertex_cost = float(os.environ.get("DAPB_PR_VERTEX_COST", 3))      # 8-GPU sweep: 3 beats 8, 1.5 and 1 (profiles/pagerank_n8_sweep_r2.md)
    if hub_edge_cost is None:
        hub_edge_cost = float(os.environ.get("DAPB_PR_HUB_EDGE_COST", 0.65))    # hub non-zeros stream through the chunk path: cheaper per edge (8-GPU sweep)
    if push_cost is None:
        push_cost = float(os.environ.get("DAPB_PR_PUSH_COST", 2.0))
    V = indeg_all.numel()
    deg = indeg_all.to(torch.float64)
    edge = torch.where(indeg_all > hub_threshold, deg * hub_edge_cost, deg)
    per_vertex = edge + vertex_cost * (indeg_all > 0) + 1
    if outdeg_all is not None and push_cost > 0:
        per_vertex = per_vertex + push_cost * ((indeg_all > 0) & (outdeg_all > 0))
    cost = torch.cumsum(per_vertex, 0)
    total = float(cost[-1].item())
    targets = torch.tensor([total * r / world for r in range(1, world)], device=indeg_all.device, dtype=cost.dtype)
    cuts = torch.searchsorted(cost, targets, right=False).tolist() if world > 1 else []
    b = [0] + [min(int(c) + 1, V) for c in cuts] + [V]
    for i in range(1, len(b)):
        b[i] = max(b[i], b[i - 1])
    return tuple(b)


def build_sharded_graph(src: torch.Tensor, dst: torch.Tensor, V: int, dist: Optional[Dist] = None,
                        block_nnz: Optional[int] = None, balance: bool = True, layout: Optional[str] = None,
                        hub_threshold: Optional[int] = None, hub_chunk: Optional[int] = None,
                        push_copies: Optional[int] = None) -> ShardedGraph:
    """De-duplicate (ref pagerank.py:41) and build this rank's CSR(A^T) rows.  ``src``/``dst`` hold the FULL
    edge list on every rank (the generators are deterministic), int64 on the target device.  ``push_copies``: NVLink
    copies one pushed contribution costs (1 = multicast, world - 1 = unicast) for the partitioner's egress term."""
    dist = dist or current()
    if layout is None:                  # "warp": warp-granular kernel + hub rows (default); "cta": CTA-wide row blocks
        layout = os.environ.get("DAPB_PR_LAYOUT", "warp")
    auto_block = block_nnz is None and layout == "warp" and os.environ.get("DAPB_PR_BLOCK_NNZ") is None
    if block_nnz is None:
        block_nnz = int(os.environ.get("DAPB_PR_BLOCK_NNZ", 512)) if layout == "warp" else 2048
    if hub_threshold is None:
        hub_threshold = int(os.environ.get("DAPB_PR_HUB_THRESHOLD", 64))       # measured: 32..256 within 5 %, 64 best
    if hub_chunk is None:
        hub_chunk = int(os.environ.get("DAPB_PR_HUB_CHUNK", 2048))
    key = torch.unique(dst.to(torch.int64) * V + src.to(torch.int64))        # sorted by (dst, src)
    d_all = torch.div(key, V, rounding_mode="floor")
    s_all = key - d_all * V
    n_edges = int(key.numel())
    outdeg = torch.bincount(s_all, minlength=V)
    n_sources = int((outdeg > 0).sum().item())
    if balance and dist.world > 1:
        copies = 1 if (push_copies is None or push_copies <= 0) else push_copies
        bounds = balanced_vertex_bounds(torch.bincount(d_all, minlength=V), dist.world,
                                        hub_threshold=hub_threshold if layout == "warp" else (1 << 62),
                                        outdeg_all=outdeg,
                                        push_cost=float(os.environ.get("DAPB_PR_PUSH_COST", 2.0)) * copies)
    else:
        bounds = tuple(vertex_range(V, r, dist.world)[0] for r in range(dist.world)) + (V,)
    v0, v1 = bounds[dist.rank], bounds[dist.rank + 1]
    lo = int(torch.searchsorted(d_all, torch.tensor([v0], device=d_all.device)).item())
    hi = int(torch.searchsorted(d_all, torch.tensor([v1], device=d_all.device)).item())
    d_loc, s_loc = d_all[lo:hi] - v0, s_all[lo:hi]
    indeg = torch.bincount(d_loc, minlength=v1 - v0)
    row_ptr = torch.zeros(v1 - v0 + 1, dtype=torch.int64, device=src.device)
    row_ptr[1:] = torch.cumsum(indeg, 0)
    od = outdeg[v0:v1]
    inv = torch.where(od > 0, 1.0 / od.clamp(min=1).to(torch.float32), torch.zeros((), device=src.device))
    # work-balanced row blocks for the CSR-stream kernel: a row costs (in-degree + 1) units, block k starts at the
    # first row whose cumulative cost is >= k * block_nnz, so a block holds <= block_nnz rows AND (unless it
    # contains one very long row) <= block_nnz non-zeros -- the skewed R-MAT degrees are balanced by construction
    nv = v1 - v0
    col_local = s_loc.to(torch.int32).contiguous()
    hub = None
    if layout == "warp":
        row_ptr, col_local, hub = split_hub_rows(row_ptr, col_local, hub_threshold, hub_chunk)
    nnz = int(row_ptr[-1].item())
    if auto_block:
        # A work item keeps a warp busy for tens of microseconds (48 warps share an SM's gather bandwidth), and the
        # end-of-iteration tail of the persistent kernel is one item long: keep >= ~32 items per resident warp so the tail
        # stays a few per cent of the iteration (measured at 2 ranks, scale 22: 512 -> 256 per item = -16 %)
        warps = 148 * 6 * 8
        per_warp = (nnz + nv) / warps
        block_nnz = 512 if per_warp >= 512 * 32 else (256 if per_warp >= 256 * 32 else 128)
    cumw = (row_ptr[:-1] + torch.arange(nv, device=src.device, dtype=torch.int64)).contiguous()
    n_blocks = max(1, (nnz + nv + block_nnz - 1) // block_nnz)
    targets = torch.arange(n_blocks, device=src.device, dtype=torch.int64) * block_nnz
    starts = torch.searchsorted(cumw, targets, right=False).clamp(max=nv)
    if starts.numel():
        starts[0] = 0
    blocks = torch.cat([starts, torch.tensor([nv], device=src.device)]).to(torch.int32).contiguous()
    return ShardedGraph(V=V, n_sources=n_sources, n_edges=n_edges, v0=v0, v1=v1, bounds=bounds, row_ptr=row_ptr,
                        col_idx=col_local, inv_outdeg=inv.contiguous(), outdeg_local=od,
                        blocks=blocks, block_nnz=block_nnz, hub=hub)


class ShardedPageRank:
    """Iterates PageRank on a :class:`ShardedGraph`; ``ranks_local`` holds the owned slice.

    ``run(n)`` performs ``n`` iterations; with the fused backend and the warp layout that is ONE persistent cooperative
    launch (csrc/pagerank.cu::pagerank_run_kernel), ``step()`` is ``run(1)``.  ``dtype`` float32 or float64 is honoured
    by every backend (the CTA layout of the fused backend is float32 only)."""

    def __init__(self, graph: ShardedGraph, cfg: PageRankConfig, dist: Optional[Dist] = None,
                 backend: str = "fused", heap=None, grid: int = 0, dtype=torch.float32, collect_stats: bool = False):
        self.g, self.cfg, self.backend = graph, cfg, backend
        self.dist = dist or current()
        if backend == "fused" and graph.hub is None and dtype != torch.float32:
            raise ValueError("the CTA-layout fused kernel is float32 only; use layout='warp' for float64")
        self.dtype = dtype
        dev = graph.row_ptr.device
        self.nv = graph.v1 - graph.v0
        self.ranks_local = torch.zeros(self.nv, dtype=self.dtype, device=dev)
        self.n = max(graph.n_sources, 1)
        self.grid = grid
        self.it = 0
        self.skip_empty = os.environ.get("DAPB_PR_NOSKIP") is None
        self.stats = None
        self._collect_stats = collect_stats
        if backend == "fused":
            from ..parallel.symm import default_heap
            self.lib = _native.load()
            esz = 8 if dtype == torch.float64 else 4
            self.heap = heap or default_heap(max(1 << 30, 2 * graph.V * esz + (64 << 20)))
            self.region = self.heap.region(f"pagerank_V{graph.V}_b{esz}", 2 * graph.V * esz, _native.MAX_RANKS)
            self.sync = torch.zeros(4, dtype=torch.int32, device=dev)
            # float: 1/outdeg (multiply); double: outdeg itself (the kernel divides exactly like pagerank.py:24)
            self.inv_outdeg = graph.inv_outdeg if dtype == torch.float32 else graph.outdeg_local.to(torch.float64).contiguous()
        else:
            self.c = None
        self.reset()

    def reset(self):
        g = self.g
        self.it = 0
        self._restored_at = 0             # iteration at which both parities were last rebuilt (init or restore)
        r0 = 1.0 / self.n
        if self.backend == "fused":
            if self.region.epoch % 2:
                self.region.epoch += 1            # init writes parity 0 == (epoch_of_first_step - 1) & 1
            torch.cuda.synchronize()
            self.dist.barrier()
            _native.check(self.lib.dapb_pagerank_init(self.heap.comm_ref(), self.inv_outdeg.data_ptr(), self.nv, g.v0, g.V,
                                                      r0, self.region.payload_off, int(self.dtype == torch.float64),
                                                      _native.stream_ptr()), "pagerank_init")
            if g.hub is not None and g.hub.row_done is not None:
                g.hub.row_done.zero_()
            torch.cuda.synchronize()
            self.dist.barrier()
            self._first = True
            io = self.inv_outdeg
        else:
            self.inv = torch.where(g.outdeg_local > 0, 1.0 / g.outdeg_local.clamp(min=1).to(self.dtype),
                                   torch.zeros((), dtype=self.dtype, device=g.row_ptr.device))
            self.c = self._gather(r0 * self.inv)
            io = self.inv
        self.ranks_local = torch.where(io > 0, torch.full_like(io, r0), torch.zeros_like(io))

    def state_dict(self) -> dict:
        """Owned slice of the rank vector + iteration counter (rank-local: one file per rank, or gather first)."""
        self.finish()
        return {"ranks_local": self.ranks_local.detach().cpu().numpy(), "it": np.int64(self.it),
                "v0": np.int64(self.g.v0), "v1": np.int64(self.g.v1)}

    def load_state_dict(self, st: dict):
        """Restore after :meth:`state_dict`: the per-source contributions r/outdeg every rank reads are rebuilt from
        the restored slices (one all-gather; collective -- every rank must call it)."""
        if int(st["v0"]) != self.g.v0 or int(st["v1"]) != self.g.v1:
            raise ValueError("checkpoint was written for a different vertex range / world size")
        dev = self.ranks_local.device
        self.ranks_local = torch.as_tensor(np.asarray(st["ranks_local"]), dtype=self.ranks_local.dtype, device=dev)
        self.it = int(st["it"])
        self._restored_at = self.it
        if self.backend == "fused":
            self.finish()
            if self.region.epoch % 2:
                self.region.epoch += 1
            if self.dtype == torch.float32:         # same arithmetic as the kernel: r * (1/outdeg) in fp32, r / outdeg in fp64
                c_loc = self.ranks_local * self.g.inv_outdeg
            else:
                deg = self.g.outdeg_local.to(self.dtype)
                c_loc = torch.where(deg > 0, self.ranks_local / deg.clamp(min=1), torch.zeros_like(deg))
            c = self._gather(c_loc)
            buf = self.heap.tensor(self.region, (2, self.g.V), self.dtype)
            buf[0].copy_(c)                     # parity read by the next step
            # the other parity must read 0 wherever nothing is pushed any more (sinks, rows without in-edges)
            buf[1].zero_()
            self._first = True
            self.finish()
        else:
            self.c = self._gather(self.ranks_local.to(self.dtype) * self.inv)

    def _gather(self, loc: torch.Tensor) -> torch.Tensor:
        if self.dist.world == 1:
            return loc.clone()
        W = self.dist.world
        sizes = [(self.g.bounds[r], self.g.bounds[r + 1]) for r in range(W)]
        m = max(b - a for a, b in sizes)
        if getattr(self, "_ag_pad", None) is None or self._ag_pad.numel() != m or self._ag_pad.dtype != loc.dtype:
            self._ag_pad = torch.zeros(m, dtype=loc.dtype, device=loc.device)
            self._ag_out = torch.empty(W * m, dtype=loc.dtype, device=loc.device)
        self._ag_pad[:loc.numel()] = loc
        self.dist.all_gather_into(self._ag_out, self._ag_pad)        # ONE NCCL all-gather of equal (padded) slices
        return torch.cat([self._ag_out[r * m:r * m + (b - a)] for r, (a, b) in enumerate(sizes)])

    def step(self):
        self.run(1)

    def run(self, n_iters: int):
        """``n_iters`` more iterations (ref pagerank.py:50-57)."""
        if n_iters <= 0:
            return
        with span("pagerank.run"):
            if self.backend == "fused" and self.g.hub is not None:
                self._run_persistent(n_iters)
                self.it += n_iters
            else:
                for _ in range(n_iters):
                    self._step()
                    self.it += 1

    def _run_persistent(self, n_iters: int):
        g, cfg = self.g, self.cfg
        a = _native.DapbPrRunArgs()
        a.row_ptr, a.col_idx, a.inv_outdeg = g.row_ptr.data_ptr(), g.col_idx.data_ptr(), self.inv_outdeg.data_ptr()
        a.blocks, a.n_blocks, a.f64 = g.blocks.data_ptr(), g.blocks.numel() - 1, int(self.dtype == torch.float64)
        a.nv_local, a.v0, a.V = self.nv, g.v0, g.V
        a.base, a.damp = cfg.q / self.n, 1.0 - cfg.q
        a.ranks_local, a.c_off, a.flag_base = self.ranks_local.data_ptr(), self.region.payload_off, self.region.flag_base
        a.epoch0 = self.region.epoch + 1
        self.region.epoch += n_iters
        # after a restore the zero rows of the read parity are rebuilt, so the same "third iteration" rule applies
        a.n_iters, a.it0 = n_iters, self.it - self._restored_at
        a.wait_first, a.skip_empty_ok, a.grid = 0 if self._first else 1, int(self.skip_empty), self.grid
        a.sync = self.sync.data_ptr()
        if self._collect_stats:
            self.stats = torch.zeros(n_iters, 4, dtype=torch.int64, device=self.sync.device)
            a.stats = self.stats.data_ptr()
        g.hub.fill(a.hub)
        _native.check(self.lib.dapb_pagerank_run(C.byref(a), self.heap.comm_ref(), _native.stream_ptr()), "pagerank_run")
        self._first = False

    def _step(self):
        g, cfg = self.g, self.cfg
        base, damp = cfg.q / self.n, 1.0 - cfg.q
        if self.backend == "fused":             # CTA layout: one launch per iteration
            wait_prev, skip = 0 if self._first else 1, int(self.it - self._restored_at >= 2 and self.skip_empty)
            _native.check(self.lib.dapb_pagerank_step(
                self.heap.comm_ref(), g.row_ptr.data_ptr(), g.col_idx.data_ptr(), g.inv_outdeg.data_ptr(),
                g.blocks.data_ptr(), g.blocks.numel() - 1, self.nv, g.v0, g.V,
                base, damp, self.ranks_local.data_ptr(), self.region.payload_off, self.region.flag_base,
                self.region.next_epoch(), self.sync.data_ptr(), wait_prev, skip, self.grid,
                _native.stream_ptr()), "pagerank_step")
            self._first = False
        else:
            s = self._spmv_baseline()
            r = torch.where(s > 0, base + damp * s, torch.zeros_like(s))
            self.ranks_local = r
            self.c = self._gather(r * self.inv)

    def _spmv_gather(self) -> torch.Tensor:
        g = self.g
        vals = self.c[g.col_idx.long()]
        if getattr(self, "_rows_idx", None) is None:
            self._rows_idx = torch.repeat_interleave(torch.arange(self.nv, device=vals.device), g.row_ptr[1:] - g.row_ptr[:-1])
            self._hub_rows_idx = None
            if g.hub is not None and g.hub.rows.numel():
                self._hub_rows_idx = torch.repeat_interleave(g.hub.rows.long(), g.hub.ptr[1:] - g.hub.ptr[:-1])
        s = torch.zeros(self.nv, dtype=self.dtype, device=vals.device)
        s.index_add_(0, self._rows_idx, vals)
        if self._hub_rows_idx is not None:                    # long rows live outside the main CSR
            s.index_add_(0, self._hub_rows_idx, self.c[g.hub.col.long()])
        return s

    def _spmv_csr(self) -> torch.Tensor:
        g = self.g
        dev = g.row_ptr.device
        if getattr(self, "_csr", None) is None:
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                ones = torch.ones(g.col_idx.numel(), dtype=self.dtype, device=dev)
                self._csr = torch.sparse_csr_tensor(g.row_ptr, g.col_idx.long(), ones, size=(self.nv, g.V))
                self._csr_hub = None
                if g.hub is not None and g.hub.rows.numel():
                    h = g.hub
                    self._csr_hub = torch.sparse_csr_tensor(h.ptr, h.col.long(),
                                                            torch.ones(h.col.numel(), dtype=self.dtype, device=dev),
                                                            size=(h.rows.numel(), g.V))
        s = torch.mv(self._csr, self.c)
        if self._csr_hub is not None:
            s.index_add_(0, g.hub.rows.long(), torch.mv(self._csr_hub, self.c))
        return s

    def _spmv_baseline(self) -> torch.Tensor:
        """s = A^T c for the owned rows, baseline arm: gather + ``index_add_`` with precomputed row indices.  (cuSPARSE CSR
        SpMV through ``torch.sparse`` was measured as the alternative -- ``_spmv_csr`` -- and is the slower one here:
        21 ms against 11 ms per iteration at scale 24 on one B200.)"""
        if os.environ.get("DAPB_PR_BASELINE_SPMV") == "csr" and self.g.row_ptr.device.type == "cuda":
            return self._spmv_csr()
        return self._spmv_gather()

    def push_bytes_per_iter(self) -> int:
        """NVLink egress of this rank per iteration: vertices with in- AND out-edges x value size x copies."""
        g = self.g
        lens = g.row_ptr[1:] - g.row_ptr[:-1]
        has_in = lens > 0
        if g.hub is not None and g.hub.rows.numel():
            has_in = has_in.clone()
            has_in[g.hub.rows.long()] = True
        n_push = int((has_in & (g.outdeg_local > 0)).sum().item())
        esz = 8 if self.dtype == torch.float64 else 4
        if self.dist.world == 1 or self.backend != "fused":
            return 0
        copies = 1 if self.heap.multicast else self.dist.world - 1
        return n_push * esz * copies

    def finish(self):
        """Make sure every rank's last pushes landed before buffers are reused (fused backend)."""
        if self.backend == "fused":
            torch.cuda.synchronize()
            self.dist.barrier()

    def ranks(self) -> torch.Tensor:
        """Full rank vector on every rank (the reference's ``ranks.collect()``, pagerank.py:60)."""
        self.finish()
        return self._gather(self.ranks_local)


def run_pagerank(edges: Iterable[Tuple[int, int]], cfg: PageRankConfig, backend: str = "auto",
                 dist: Optional[Dist] = None, checkpoint: Optional[str] = None, checkpoint_every: int = 0,
                 resume: bool = False) -> Dict[int, float]:
    """Rank the vertices of an edge list with arbitrary integer ids; returns {vertex: rank} for the
    vertices the reference would print (those that still hold a rank).  With ``checkpoint`` every rank writes its
    slice to ``<checkpoint>.rank<r>`` every ``checkpoint_every`` iterations and ``resume`` continues from there."""
    if backend == "auto":
        backend = "fused" if _native.available() else "torch"
    edges = [(int(a), int(b)) for a, b in edges]
    if backend == "oracle":
        return pagerank_oracle(edges, cfg)
    dist = dist or current()
    ids = sorted(set(v for e in edges for v in e))
    remap = {v: i for i, v in enumerate(ids)}
    src = torch.tensor([remap[a] for a, _ in edges], dtype=torch.int64, device=dist.device)
    dst = torch.tensor([remap[b] for _, b in edges], dtype=torch.int64, device=dist.device)
    g = build_sharded_graph(src, dst, len(ids), dist)
    pr = ShardedPageRank(g, cfg, dist, backend="fused" if backend == "fused" else "nccl", dtype=torch.float64)
    every = checkpoint_every if (path_wanted := checkpoint is not None) and checkpoint_every > 0 else 0
    path = None if checkpoint is None else ck.rank_path(checkpoint, dist.rank)
    meta = dict(kind="pagerank", V=len(ids), n_edges=int(g.n_edges), world=int(dist.world))
    if path and resume:
        have = all(dist.all_gather_object(os.path.exists(path)))
        if have:
            st, m = ck.load_checkpoint(path)
            ck.check_meta(m, **meta)
            pr.load_state_dict(st)
    while pr.it < cfg.n_iterations:
        left = cfg.n_iterations - pr.it
        pr.run(min(left, every - pr.it % every) if every else left)      # one persistent launch per checkpoint interval
        if path and every and (pr.it % every == 0 or pr.it == cfg.n_iterations):
            ck.save_checkpoint(path, pr.state_dict(), dict(meta, iteration=pr.it))
    r = pr.ranks().double().cpu().numpy()
    return {ids[i]: float(r[i]) for i in range(len(ids)) if r[i] > 0}
