"""PageRank (reference: ``graph_computation/pagerank.py``).

Graph preparation mirrors ``links.distinct().groupByKey()`` (:41): duplicate edges are dropped,
self-loops kept; ``n`` is the number of distinct SOURCE vertices (:44).  The iteration is the vector
form of :52-57 including the reference's "absent vertex" rule.

Backends: ``oracle`` (Python floats, bit-exact with the values recorded in pagerank.py:66-68), ``torch``
(CSR SpMV via ``index_add_`` + ``all_gather`` -- the NCCL baseline), ``fused`` (csrc/pagerank.cu: SpMV +
damping + all-gather push in one launch per iteration).
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
    partial: torch.Tensor         # float32 [n_chunks] scratch
    ptr: Optional[torch.Tensor] = None   # int64 [n_hub + 1] offsets of the hub rows in col

    def c_struct(self) -> "_native.DapbPrHub":
        h = _native.DapbPrHub()
        h.hub_rows, h.hub_col = self.rows.data_ptr(), self.col.data_ptr()
        h.chunk_start, h.chunk_end = self.chunk_start.data_ptr(), self.chunk_end.data_ptr()
        h.hub_chunk_ptr, h.chunk_partial = self.chunk_ptr.data_ptr(), self.partial.data_ptr()
        h.n_hub, h.n_chunks = int(self.rows.numel()), int(self.chunk_start.numel())
        return h


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
                                         torch.zeros(0, dtype=torch.float32, device=dev), torch.zeros(1, dtype=i64, device=dev))
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
                  chunk_ptr.to(i32).contiguous(), torch.zeros(owner.numel(), dtype=torch.float32, device=dev), hub_ptr)
    return rp, main_col, hub


def balanced_vertex_bounds(indeg_all: torch.Tensor, world: int, vertex_cost: Optional[float] = None,
                           hub_threshold: int = 64, hub_edge_cost: Optional[float] = None):
    """Contiguous destination-vertex ranges of (almost) equal modelled cost per rank.  A vertex costs 1, plus
    ``vertex_cost`` if it has in-edges (row bookkeeping + its new contribution is pushed to every peer each iteration),
    plus 1 per in-edge -- ``hub_edge_cost`` per in-edge for hub rows (> hub_threshold in-edges), whose non-zeros are
    streamed by the cheaper chunk kernel.  R-MAT graphs concentrate edges on low vertex ids, so equal-size ranges would
    give rank 0 most of the work.  ``DAPB_PR_VERTEX_COST`` / ``DAPB_PR_HUB_EDGE_COST`` override."""
    if vertex_cost is None:
        vertex_cost = float(os.environ.get("DAPB_PR_VERTEX_COST", 8))
    if hub_edge_cost is None:
        hub_edge_cost = float(os.environ.get("DAPB_PR_HUB_EDGE_COST", 1.0))
    V = indeg_all.numel()
    deg = indeg_all.to(torch.float64)
    edge = torch.where(indeg_all > hub_threshold, deg * hub_edge_cost, deg)
    cost = torch.cumsum(edge + vertex_cost * (indeg_all > 0) + 1, 0)
    total = float(cost[-1].item())
    targets = torch.tensor([total * r / world for r in range(1, world)], device=indeg_all.device, dtype=cost.dtype)
    cuts = torch.searchsorted(cost, targets, right=False).tolist() if world > 1 else []
    b = [0] + [min(int(c) + 1, V) for c in cuts] + [V]
    for i in range(1, len(b)):
        b[i] = max(b[i], b[i - 1])
    return tuple(b)


def build_sharded_graph(src: torch.Tensor, dst: torch.Tensor, V: int, dist: Optional[Dist] = None,
                        block_nnz: Optional[int] = None, balance: bool = True, layout: Optional[str] = None,
                        hub_threshold: Optional[int] = None, hub_chunk: Optional[int] = None) -> ShardedGraph:
    """De-duplicate (ref pagerank.py:41) and build this rank's CSR(A^T) rows.  ``src``/``dst`` hold the FULL
    edge list on every rank (the generators are deterministic), int64 on the target device."""
    dist = dist or current()
    if layout is None:                  # "warp": warp-granular kernel + hub rows (default); "cta": CTA-wide row blocks
        layout = os.environ.get("DAPB_PR_LAYOUT", "warp")
    if block_nnz is None:
        block_nnz = 512 if layout == "warp" else 2048
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
        bounds = balanced_vertex_bounds(torch.bincount(d_all, minlength=V), dist.world,
                                        hub_threshold=hub_threshold if layout == "warp" else (1 << 62))
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
    """Iterates PageRank on a :class:`ShardedGraph`; ``ranks_local`` holds the owned slice."""

    def __init__(self, graph: ShardedGraph, cfg: PageRankConfig, dist: Optional[Dist] = None,
                 backend: str = "fused", heap=None, grid: int = 0, dtype=torch.float32):
        self.g, self.cfg, self.backend = graph, cfg, backend
        self.dtype = torch.float32 if backend == "fused" else dtype
        self.dist = dist or current()
        dev = graph.row_ptr.device
        self.nv = graph.v1 - graph.v0
        self.ranks_local = torch.zeros(self.nv, dtype=torch.float32, device=dev)
        self.n = max(graph.n_sources, 1)
        self.grid = grid
        self.it = 0
        import os
        self.skip_empty = os.environ.get("DAPB_PR_NOSKIP") is None
        if backend == "fused":
            from ..parallel.symm import default_heap
            self.lib = _native.load()
            self.heap = heap or default_heap(max(1 << 30, 2 * graph.V * 4 + (64 << 20)))
            self.region = self.heap.region(f"pagerank_V{graph.V}", 2 * graph.V * 4, _native.MAX_RANKS)
            self.ticket = torch.zeros(1, dtype=torch.int32, device=dev)
            self._hub = graph.hub.c_struct() if graph.hub is not None else None
        else:
            self.c = None
        self.reset()

    def reset(self):
        g = self.g
        self.it = 0
        self._skip_from = 2               # rows without in-edges may be skipped once both parities hold their zeros
        r0 = 1.0 / self.n
        if self.backend == "fused":
            if self.region.epoch % 2:
                self.region.epoch += 1            # init writes parity 0 == (epoch_of_first_step - 1) & 1
            torch.cuda.synchronize()
            self.dist.barrier()
            _native.check(self.lib.dapb_pagerank_init(self.heap.comm_ref(), g.inv_outdeg.data_ptr(), self.nv, g.v0, g.V,
                                                      r0, self.region.payload_off, _native.stream_ptr()), "pagerank_init")
            torch.cuda.synchronize()
            self.dist.barrier()
            self._first = True
        else:
            self.inv = torch.where(g.outdeg_local > 0, 1.0 / g.outdeg_local.clamp(min=1).to(self.dtype),
                                   torch.zeros((), dtype=self.dtype, device=g.row_ptr.device))
            self.c = self._gather(r0 * self.inv)
        io = g.inv_outdeg.to(self.dtype)
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
        self._skip_from = self.it + 2
        if self.backend == "fused":
            self.finish()
            if self.region.epoch % 2:
                self.region.epoch += 1
            c = self._gather(self.ranks_local.float() * self.g.inv_outdeg)
            self.heap.tensor(self.region, (2, self.g.V), torch.float32)[0].copy_(c)     # parity read by the next step
            self._first = True
            self.finish()
        else:
            self.c = self._gather(self.ranks_local.to(self.dtype) * self.inv)

    def _gather(self, loc: torch.Tensor) -> torch.Tensor:
        if self.dist.world == 1:
            return loc.clone()
        sizes = [(self.g.bounds[r], self.g.bounds[r + 1]) for r in range(self.dist.world)]
        m = max(b - a for a, b in sizes)
        pad = torch.zeros(m, dtype=loc.dtype, device=loc.device)
        pad[:loc.numel()] = loc
        parts = self.dist.all_gather(pad)
        return torch.cat([p[:b - a] for p, (a, b) in zip(parts, sizes)])

    def step(self):
        with span("pagerank.step"):
            self._step()
        self.it += 1

    def _step(self):
        g, cfg = self.g, self.cfg
        base, damp = cfg.q / self.n, 1.0 - cfg.q
        if self.backend == "fused":
            wait_prev, skip = 0 if self._first else 1, int(self.it >= self._skip_from and self.skip_empty)
            if g.hub is not None:               # warp-granular kernel + hub rows (csrc/pagerank.cu)
                _native.check(self.lib.dapb_pagerank_step_v2(
                    self.heap.comm_ref(), g.row_ptr.data_ptr(), g.col_idx.data_ptr(), g.inv_outdeg.data_ptr(),
                    g.blocks.data_ptr(), g.blocks.numel() - 1, self.nv, g.v0, g.V, C.byref(self._hub),
                    base, damp, self.ranks_local.data_ptr(), self.region.payload_off, self.region.flag_base,
                    self.region.next_epoch(), self.ticket.data_ptr(), wait_prev, skip, self.grid,
                    _native.stream_ptr()), "pagerank_step_v2")
            else:
                _native.check(self.lib.dapb_pagerank_step(
                    self.heap.comm_ref(), g.row_ptr.data_ptr(), g.col_idx.data_ptr(), g.inv_outdeg.data_ptr(),
                    g.blocks.data_ptr(), g.blocks.numel() - 1, self.nv, g.v0, g.V,
                    base, damp, self.ranks_local.data_ptr(), self.region.payload_off, self.region.flag_base,
                    self.region.next_epoch(), self.ticket.data_ptr(), wait_prev, skip, self.grid,
                    _native.stream_ptr()), "pagerank_step")
            self._first = False
        else:
            vals = self.c[g.col_idx.long()]
            rows = torch.repeat_interleave(torch.arange(self.nv, device=vals.device), g.row_ptr[1:] - g.row_ptr[:-1])
            s = torch.zeros(self.nv, dtype=self.dtype, device=vals.device)
            s.index_add_(0, rows, vals)
            if g.hub is not None and g.hub.rows.numel():          # long rows live outside the main CSR
                h = g.hub
                hrows = torch.repeat_interleave(h.rows.long(), h.ptr[1:] - h.ptr[:-1])
                s.index_add_(0, hrows, self.c[h.col.long()])
            r = torch.where(s > 0, base + damp * s, torch.zeros_like(s))
            self.ranks_local = r
            self.c = self._gather(r * self.inv)

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
    path = None if checkpoint is None else ck.rank_path(checkpoint, dist.rank)
    meta = dict(kind="pagerank", V=len(ids), n_edges=int(g.n_edges), world=int(dist.world))
    if path and resume:
        have = all(dist.all_gather_object(os.path.exists(path)))
        if have:
            st, m = ck.load_checkpoint(path)
            ck.check_meta(m, **meta)
            pr.load_state_dict(st)
    while pr.it < cfg.n_iterations:
        pr.step()
        if path and checkpoint_every > 0 and (pr.it % checkpoint_every == 0 or pr.it == cfg.n_iterations):
            ck.save_checkpoint(path, pr.state_dict(), dict(meta, iteration=pr.it))
    r = pr.ranks().double().cpu().numpy()
    return {ids[i]: float(r[i]) for i in range(len(ids)) if r[i] > 0}
