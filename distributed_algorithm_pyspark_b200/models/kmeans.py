"""Lloyd's k-means (reference: ``machine_learning/k-means.py``).

Semantics kept from the reference: initial centres = ``takeSample(False, k, 42)`` (:53), ties go to the
lowest centre index (:25), a cluster that receives no point keeps its centre (:70-71), a fixed number
of iterations (``convergeDist`` is defined but unused, :16).

Backends: ``oracle`` (NumPy), ``torch`` (``cdist``-style matmul + ``index_add_`` + ``all_reduce`` -- the
NCCL baseline), ``fused`` (sm_100a: assignment kernel -- tcgen05 tensor-core version for the benchmark
shape, CUDA-core version otherwise -- accumulating per-centroid sums, then ONE finalize kernel that
fuses reduce-scatter, mean update and centroid all-gather over NVLink peer memory).
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch

from .. import _native
from ..config import KMeansConfig
from ..oracle import spark_semantics as sp
from ..oracle.algorithms import KMeansResult, kmeans_oracle
from ..parallel.dist import Dist, current
from ..parallel.partition import RowSharding
from ..utils import checkpoint as ck
from ..utils.trace import span

_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float64: 2}


def initial_center_indices(n_points: int, cfg: KMeansConfig):
    sizes = np.diff(sp.partition_offsets(n_points, cfg.n_slices))
    return sp.take_sample_indices(sizes, cfg.k, cfg.seed)


def run_kmeans(points: np.ndarray, cfg: KMeansConfig, backend: str = "auto", dist: Optional[Dist] = None,
               init_centers: Optional[np.ndarray] = None, checkpoint: Optional[str] = None,
               checkpoint_every: int = 0, resume: bool = False) -> KMeansResult:
    """Cluster a host array the way the reference script does (small / parity problems).
    ``checkpoint`` / ``checkpoint_every`` / ``resume``: centres + iteration counter (utils/checkpoint.py)."""
    if backend == "auto":
        backend = "fused" if _native.available() else "torch"
    if backend == "oracle":
        return kmeans_oracle(points, cfg, init_centers)
    dist = dist or current()
    pts = np.asarray(points, dtype=np.float64)
    n = pts.shape[0]
    idx = [] if init_centers is not None else initial_center_indices(n, cfg)
    c0 = np.asarray(init_centers, dtype=np.float64) if init_centers is not None else pts[idx]
    sh = RowSharding.spark(n, cfg.n_slices, dist.world)
    r0, r1 = sh.local_row_range(dist.rank)
    X = torch.as_tensor(np.ascontiguousarray(pts[r0:r1]), dtype=torch.float64, device=dist.device)
    eng = ShardedKMeans(X, cfg.k, torch.as_tensor(c0, dtype=torch.float64, device=dist.device), dist=dist,
                        backend="fused" if backend == "fused" else "nccl")
    history, t0 = [], 0
    meta = dict(kind="kmeans", k=int(cfg.k), dim=int(pts.shape[1]), n_points=int(n))
    if checkpoint and resume and os.path.exists(checkpoint):
        st, m = ck.load_checkpoint(checkpoint)
        ck.check_meta(m, **meta)
        t0, history = int(m["iteration"]), list(st["history"])
        eng.centers.copy_(torch.as_tensor(st["centers"], dtype=eng.centers.dtype))
    for t in range(t0, cfg.n_iterations):
        eng.step()
        history.append(eng.centers.cpu().numpy().copy())
        if checkpoint and checkpoint_every > 0 and ((t + 1) % checkpoint_every == 0 or t + 1 == cfg.n_iterations) \
                and dist.is_main:
            ck.save_checkpoint(checkpoint, dict(centers=history[-1], history=np.stack(history)), dict(meta, iteration=t + 1))
    assign_local = eng.assign()                                    # final centres (ref k-means.py:76, A.7)
    parts = dist.all_gather_object(assign_local.cpu().numpy())
    return KMeansResult(centers=eng.centers.cpu().numpy(), assignment=np.concatenate(parts).astype(np.int64),
                        init_indices=list(idx), history=history)


class ShardedKMeans:
    """Per-rank k-means over a resident ``[n_local, d]`` shard; centres replicated on every rank."""

    def __init__(self, X: torch.Tensor, k: int, centers: torch.Tensor, dist: Optional[Dist] = None,
                 backend: str = "fused", use_tc: Optional[bool] = None, heap=None):
        self.dist = dist or current()
        self.backend = backend
        assert X.dim() == 2 and X.is_contiguous() and X.dtype in _DT
        self.X, self.k = X, int(k)
        self.n, self.d = X.shape
        self.acc_dtype = torch.float64 if X.dtype == torch.float64 else torch.float32
        self.centers = centers.to(device=X.device, dtype=self.acc_dtype).clone().contiguous()
        self.counts_out = torch.zeros(self.k, dtype=torch.int32, device=X.device)
        self.iteration = 0
        if backend == "fused":
            from ..parallel.symm import default_heap
            self.lib = _native.load()
            self.heap = heap or default_heap()
            self.sums = torch.zeros(self.k, self.d, dtype=self.acc_dtype, device=X.device)
            self.counts = torch.zeros(self.k, dtype=torch.int32, device=X.device)
            eb = 8 if self.acc_dtype == torch.float64 else 4
            nbytes = int(self.lib.dapb_kmeans_payload_bytes(self.dist.world, self.k, self.d, eb))
            self.region = self.heap.region(f"kmeans_k{self.k}_d{self.d}_{eb}", max(nbytes, 256), 2 * _native.MAX_RANKS)
            from ..ops import kmeans_tc
            self.tc = None
            if use_tc is None:
                use_tc = kmeans_tc.supported(self.X, self.k)
            if use_tc:
                self.tc = kmeans_tc.KMeansTC(self.X, self.k)
            self._operand_ready = False         # tensor-core operands match self.centers (emitted by the finalize kernel)
            # launches per Lloyd iteration in steady state: assignment + finalize (which emits the next operands)
            self.launches_per_iter = 2

    def state_dict(self) -> dict:
        """Everything a resumed run needs (utils/checkpoint.py); replicated, so rank 0's copy is enough."""
        return {"centers": self.centers.detach().cpu().numpy(), "iteration": np.int64(self.iteration)}

    def load_state_dict(self, st: dict):
        self.centers.copy_(torch.as_tensor(np.asarray(st["centers"]), dtype=self.acc_dtype))
        self.iteration = int(st["iteration"])
        self._operand_ready = False

    def step(self):
        """One Lloyd iteration (ref k-means.py:55-71)."""
        with span("kmeans.step"):
            self._step()
        self.iteration += 1

    def _step(self):
        if self.backend == "fused":
            st = _native.stream_ptr()
            if self.tc is not None:
                self.tc.assign_accumulate(self.centers, self.sums, self.counts, prepared=self._operand_ready)
                _native.check(self.lib.dapb_kmeans_finalize_tc(
                    self.heap.comm_ref(), self.sums.data_ptr(), self.counts.data_ptr(), self.centers.data_ptr(), self.k,
                    self.d, self.region.payload_off, self.region.flag_base, self.region.next_epoch(),
                    self.counts_out.data_ptr(), 1, self.tc.b_operand.data_ptr(), self.tc.cnorm.data_ptr(),
                    self.tc.xnorm_max, st), "kmeans_finalize_tc")
                self._operand_ready = True
                return
            else:
                _native.check(self.lib.dapb_kmeans_assign(self.X.data_ptr(), self.n, self.d, _DT[self.X.dtype],
                                                          self.centers.data_ptr(), self.k, 0, self.sums.data_ptr(),
                                                          self.counts.data_ptr(), st), "kmeans_assign")
            _native.check(self.lib.dapb_kmeans_finalize(self.heap.comm_ref(), self.sums.data_ptr(), self.counts.data_ptr(),
                                                        self.centers.data_ptr(), self.k, self.d,
                                                        int(self.acc_dtype == torch.float64), self.region.payload_off,
                                                        self.region.flag_base, self.region.next_epoch(),
                                                        self.counts_out.data_ptr(), 1, st), "kmeans_finalize")
        else:
            # baseline: chunked cuBLAS distances + arg-min + index_add_ (chunks of ~0.5 M rows keep the launches few
            # and the [chunk, k] distance tile at a few GB), then one NCCL all-reduce
            sums = torch.zeros(self.k, self.d, dtype=self.acc_dtype, device=self.X.device)
            cnt = torch.zeros(self.k, dtype=torch.int64, device=self.X.device)
            chunk = self._chunk_rows()
            onehot_gemm = self.X.is_cuda and self.X.dtype in (torch.bfloat16, torch.float32) and self.k >= 64
            for i in range(0, self.n, chunk):
                a = self._assign_torch(i, min(self.n, i + chunk))
                x = self.X[i:i + chunk]
                if onehot_gemm:
                    # per-cluster sums as a second cuBLAS GEMM (one-hot^T @ X, fp32 accumulate and output): far faster
                    # on a GPU than index_add_, whose atomics serialise on the k hot rows
                    if getattr(self, "_oh", None) is None or self._oh.shape[0] != x.shape[0]:
                        self._oh = torch.empty(x.shape[0], self.k, dtype=x.dtype, device=x.device)
                    oh = self._oh
                    oh.zero_()
                    oh.scatter_(1, a[:, None], 1.0)
                    try:
                        sums += torch.mm(oh.T, x, out_dtype=torch.float32) if x.dtype != torch.float32 else oh.T @ x
                    except (TypeError, RuntimeError):
                        sums += (oh.T @ x).float()
                else:
                    sums.index_add_(0, a, x.to(self.acc_dtype))
                cnt += torch.bincount(a, minlength=self.k)
            cnt = cnt.to(self.acc_dtype)
            buf = torch.cat([sums, cnt[:, None]], dim=1)
            self.dist.all_reduce_sum(buf)
            sums, cnt = buf[:, :-1], buf[:, -1]
            has = cnt > 0
            self.centers[has] = sums[has] / cnt[has][:, None]
            self.counts_out.copy_(cnt.to(torch.int32))

    def _chunk_rows(self) -> int:
        if self.X.dtype == torch.float64:
            return max(1, (64 << 20) // max(self.k * self.d * 8, 1))      # [chunk, k, d] broadcast tile
        return max(1, (2 << 30) // max(self.k * 4, 1))                     # [chunk, k] fp32 distance tile of 2 GB

    def _assign_torch(self, lo: int = 0, hi: Optional[int] = None) -> torch.Tensor:
        """Assignment of rows [lo, hi) (default: all rows, chunked)."""
        if hi is None:
            hi = self.n
        Cc = self.centers
        chunk = self._chunk_rows()
        outs = []
        if self.X.dtype == torch.float64:
            for i in range(lo, hi, chunk):                         # exact (x-c)^2 form for parity data
                x = self.X[i:min(hi, i + chunk)]
                outs.append(((x[:, None, :] - Cc[None, :, :]) ** 2).sum(-1).argmin(dim=1))
        else:
            cn = (Cc * Cc).sum(1)
            Cm = Cc.to(self.X.dtype)
            CmT = Cm.T.contiguous()
            for i in range(lo, hi, chunk):
                x = self.X[i:min(hi, i + chunk)]
                try:        # fp32 output straight from cuBLAS, then two in-place passes and the arg-min
                    dm = torch.mm(x, CmT, out_dtype=torch.float32) if x.dtype != torch.float32 else x @ CmT
                except (TypeError, RuntimeError):
                    dm = (x @ CmT).float()
                dm.mul_(-2.0).add_(cn[None, :])
                outs.append(dm.argmin(dim=1))
                del dm
        return outs[0] if len(outs) == 1 else (torch.cat(outs) if outs else torch.empty(0, dtype=torch.int64, device=self.X.device))

    def assign(self) -> torch.Tensor:
        """Assignment of the local points under the current centres (int64)."""
        if self.backend == "fused":
            out = torch.empty(self.n, dtype=torch.int32, device=self.X.device)
            _native.check(self.lib.dapb_kmeans_assign(self.X.data_ptr(), self.n, self.d, _DT[self.X.dtype],
                                                      self.centers.data_ptr(), self.k, out.data_ptr(), 0, 0,
                                                      _native.stream_ptr()), "kmeans_assign")
            return out.to(torch.int64)
        return self._assign_torch()
