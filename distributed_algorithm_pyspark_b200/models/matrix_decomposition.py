"""Low-rank matrix decomposition (reference: ``matrix_computation/matrix_decomposition.py``).

ALS on ``R ~ U V^T`` with ridge ``lam * (#rows of the other factor)`` (:24-33) and the rmse of :19-21.
``R`` is row-sharded and resident; per iteration

  U-step (row-local):  ``U_loc = solve(V^T V + lam n I, (R_loc V)^T)^T``                       (:52-57)
  V-step (reduced):    ``V = solve(U^T U + lam m I, (R^T U)^T)^T`` where ``U^T U`` (k x k) and
                       ``R^T U`` (n x k) are tall-skinny ``X^T Y`` products summed over the row shards --
                       kernel F5, whose epilogue performs the cross-GPU reduction                (:60-63)

The benchmark-shaped decomposition (BASELINE.json config 5) uses the same F5 kernel for the Gram
``A^T A`` of a tall-skinny ``A`` followed by an on-device Cholesky (the R factor of A's QR) and, optionally,
a symmetric eigen-decomposition of the 512 x 512 Gram.

Backends: ``oracle`` (NumPy, per-row solves as the reference), ``torch`` (matmul + all_reduce +
``torch.linalg.solve`` -- the NCCL/cuBLAS baseline), ``fused`` (csrc/gram.cu, csrc/gram_tc.cu).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch

from .. import _native
from ..config import ALSConfig
from ..oracle.algorithms import als_draw, als_oracle
from ..parallel.dist import Dist, current
from ..parallel.partition import RowSharding
from ..utils import checkpoint as ck

_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float64: 2}


@dataclass
class ALSResult:
    U: np.ndarray
    V: np.ndarray
    rmse: List[float]


class GramOp:
    """``C = X^T Y`` summed over every rank's rows (F5), result replicated."""

    def __init__(self, P: int, Q: int, dtype: torch.dtype, device, dist: Optional[Dist] = None, heap=None,
                 splits: int = 0, tag: str = ""):
        from ..parallel.symm import default_heap
        self.dist = dist or current()
        self.lib = _native.load()
        self.P, self.Q, self.dtype = P, Q, dtype
        self.acc = torch.float64 if dtype == torch.float64 else torch.float32
        eb = 8 if self.acc == torch.float64 else 4
        self.heap = heap or default_heap()
        need = 2 * _native.MAX_RANKS * ((P * Q + 3) & ~3) * eb
        self.region = self.heap.region(f"gram_{P}x{Q}_{eb}{tag}", need, _native.MAX_RANKS) if self.dist.world > 1 else None
        sms = torch.cuda.get_device_properties(device).multi_processor_count
        tiles = ((P + 31) // 32) * ((Q + 31) // 32)
        self.splits = splits or max(1, min(64, (2 * sms) // tiles))
        self.partial = torch.empty(self.splits * P * Q, dtype=self.acc, device=device)
        self.ticket = torch.zeros(1, dtype=torch.int32, device=device)
        self.out = torch.empty(P, Q, dtype=self.acc, device=device)

    def __call__(self, X: torch.Tensor, Y: torch.Tensor, allreduce: bool = True) -> torch.Tensor:
        assert X.dtype == Y.dtype == self.dtype and X.shape[0] == Y.shape[0]
        assert X.stride(1) == 1 and Y.stride(1) == 1
        reg = self.region
        _native.check(self.lib.dapb_gram(self.heap.comm_ref(), X.data_ptr(), Y.data_ptr(), X.shape[0], self.P, self.Q,
                                         X.stride(0), Y.stride(0), _DT[self.dtype], self.partial.data_ptr(), self.splits,
                                         self.out.data_ptr(), self.ticket.data_ptr(), reg.payload_off if reg else 0,
                                         reg.flag_base if reg else 0, reg.next_epoch() if (reg and allreduce) else 1,
                                         int(allreduce), _native.stream_ptr()), "gram")
        return self.out


def run_als(cfg: ALSConfig, R: Optional[np.ndarray] = None, U0: Optional[np.ndarray] = None,
            V0: Optional[np.ndarray] = None, backend: str = "auto", dist: Optional[Dist] = None,
            dtype=torch.float64, callback=None, checkpoint: Optional[str] = None, checkpoint_every: int = 0,
            resume: bool = False) -> ALSResult:
    """ALS exactly as the reference drives it; ``R``/``U0``/``V0`` default to the reference's draws (:42-44).
    ``checkpoint``: V (replicated), the rmse history and the iteration counter; U is recomputed from V each
    iteration (:52-56); it is stored only so that a finished run can be reloaded."""
    if backend == "auto":
        backend = "fused" if _native.available() else "torch"
    if R is None:
        R, U0, V0 = als_draw(cfg)
    if backend == "oracle":
        U, V, errs = als_oracle(R, U0, V0, cfg)
        if callback:
            for t, e in enumerate(errs):
                callback(t, e)
        return ALSResult(U, V, errs)
    dist = dist or current()
    dev = dist.device
    m, n = R.shape
    k = U0.shape[1]
    sh = RowSharding.even(m, dist.world, dist.world)
    r0, r1 = sh.local_row_range(dist.rank)
    Rl = torch.as_tensor(np.ascontiguousarray(R[r0:r1]), dtype=dtype, device=dev)
    V = torch.as_tensor(V0, dtype=dtype, device=dev).contiguous()
    Ul = torch.as_tensor(np.ascontiguousarray(U0[r0:r1]), dtype=dtype, device=dev)
    errs = []
    fused = backend == "fused"
    if fused:
        lib = _native.load()
        f64 = int(dtype == torch.float64)
        g_vv = GramOp(k, k, dtype, dev, dist, tag="vv")
        g_uu = GramOp(k, k, dtype, dev, dist, tag="uu")
        g_ru = GramOp(n, k, dtype, dev, dist, tag="ru")
        B = torch.empty(r1 - r0, k, dtype=dtype, device=dev)
        sq = torch.zeros(1, dtype=torch.float64, device=dev)
    t0 = 0
    meta = dict(kind="als", m=int(m), n=int(n), k=int(k))
    if checkpoint and resume and os.path.exists(checkpoint):
        stc, mt = ck.load_checkpoint(checkpoint)
        ck.check_meta(mt, **meta)
        t0, errs = int(mt["iteration"]), stc["errs"].tolist()
        V = torch.as_tensor(stc["V"], dtype=dtype, device=dev).contiguous()
        Ul = torch.as_tensor(np.ascontiguousarray(stc["U"][r0:r1]), dtype=dtype, device=dev)
    for t in range(t0, cfg.n_iterations):
        if fused:
            st = _native.stream_ptr()
            G = g_vv(V, V, allreduce=False)                                    # V is replicated
            _native.check(lib.dapb_rows_times_small(Rl.data_ptr(), V.data_ptr(), r1 - r0, n, k, Rl.stride(0), 0, f64,
                                                    B.data_ptr(), st), "rows_times_small")
            Ul = torch.empty(r1 - r0, k, dtype=dtype, device=dev)
            _native.check(lib.dapb_solve_rows(G.data_ptr(), cfg.lam * n, B.data_ptr(), r1 - r0, k, f64, Ul.data_ptr(), st),
                          "solve_rows")
            Guu = g_uu(Ul, Ul)                                                 # all-reduced k x k
            RtU = g_ru(Rl, Ul)                                                 # all-reduced n x k
            Vn = torch.empty(n, k, dtype=dtype, device=dev)
            _native.check(lib.dapb_solve_rows(Guu.data_ptr(), cfg.lam * m, RtU.data_ptr(), n, k, f64, Vn.data_ptr(), st),
                          "solve_rows")
            V = Vn
            sq.zero_()
            _native.check(lib.dapb_sq_residual(Rl.data_ptr(), Ul.data_ptr(), V.data_ptr(), r1 - r0, n, k, f64,
                                               sq.data_ptr(), st), "sq_residual")
            s = sq.clone()
        else:
            eye = torch.eye(k, dtype=dtype, device=dev)
            G = V.T @ V + cfg.lam * n * eye
            Ul = torch.linalg.solve(G, (Rl @ V).T).T.contiguous()
            Guu = Ul.T @ Ul
            RtU = Rl.T @ Ul
            dist.all_reduce_sum(Guu)
            dist.all_reduce_sum(RtU)
            V = torch.linalg.solve(Guu + cfg.lam * m * eye, RtU.T).T.contiguous()
            s = ((Rl - Ul @ V.T) ** 2).sum().double().reshape(1)
        dist.all_reduce_sum(s)
        err = float(torch.sqrt(s / (m * n)).item())                           # ref :19-21
        errs.append(err)
        if callback:
            callback(t, err)
        if checkpoint and checkpoint_every > 0 and ((t + 1) % checkpoint_every == 0 or t + 1 == cfg.n_iterations):
            U_all = np.concatenate(dist.all_gather_object(Ul.double().cpu().numpy()), axis=0)
            if dist.is_main:
                ck.save_checkpoint(checkpoint, dict(U=U_all, V=V, errs=np.asarray(errs)), dict(meta, iteration=t + 1))
    parts = dist.all_gather_object(Ul.double().cpu().numpy())
    return ALSResult(np.concatenate(parts, axis=0), V.double().cpu().numpy(), errs)


class GramDecomposition:
    """Benchmark-shaped decomposition of a row-sharded tall-skinny ``A`` [n_local, p]:
    ``G = A^T A`` (F5, fused cross-GPU reduce) -> Cholesky ``G = L L^T`` on device (``R = L^T`` is the QR
    factor of A) -> optional eigen-decomposition (singular values of A = sqrt(eig))."""

    def __init__(self, A: torch.Tensor, dist: Optional[Dist] = None, backend: str = "fused", use_tc: Optional[bool] = None):
        self.dist = dist or current()
        self.A, self.backend = A, backend
        self.n, self.p = A.shape
        self.tc = None
        if backend == "fused":
            self.lib = _native.load()
            from ..ops import gram_tc
            if use_tc is None:
                use_tc = gram_tc.supported(A, self.dist.world)
            if use_tc:
                self.tc = gram_tc.GramTC(A, self.dist)
            else:
                self.op = GramOp(self.p, self.p, A.dtype, A.device, self.dist, tag="bench")
            self.info = torch.zeros(1, dtype=torch.int32, device=A.device)

    def gram(self) -> torch.Tensor:
        if self.backend == "fused":
            return self.tc() if self.tc is not None else self.op(self.A, self.A)
        try:        # cuBLAS bf16 GEMM with fp32 output (a bf16 output would round every entry of G to 8 bits)
            G = torch.mm(self.A.T, self.A, out_dtype=torch.float32) if self.A.dtype != torch.float32 else self.A.T @ self.A
        except (TypeError, RuntimeError):
            G = (self.A.T @ self.A).float()
        self.dist.all_reduce_sum(G)
        return G

    def cholesky(self, G: torch.Tensor, jitter: float = 0.0) -> torch.Tensor:
        if self.backend == "fused":
            L = G.clone()
            if jitter:
                L.diagonal().add_(jitter)
            self.info.zero_()
            _native.check(self.lib.dapb_cholesky(L.data_ptr(), self.p, int(L.dtype == torch.float64), self.info.data_ptr(),
                                                 _native.stream_ptr()), "cholesky")
            return L
        return torch.linalg.cholesky(G + jitter * torch.eye(self.p, device=G.device, dtype=G.dtype))

    eigh_native = True

    def eigh(self, G: torch.Tensor, tol: float = 0.0, max_sweeps: int = 0):
        """(eigenvalues ascending, eigenvectors in columns) of the symmetric ``G``.  Fused backend: the block Jacobi
        solver of csrc/dense_small.cu (one cooperative launch, same dtype as ``G``); baseline: ``torch.linalg.eigh``
        (cuSOLVER) in fp64."""
        n = G.shape[0]
        if self.backend != "fused" or G.dtype not in (torch.float32, torch.float64) or n % 8 or 16 * n * 8 > 200 * 1024:
            return torch.linalg.eigh(G.double())
        # the rotations run in fp64 whatever the input: an fp32 one-sided Jacobi stalls at normalised off-diagonals of
        # ~1e-6 (rounding of an n-term dot product), which for clustered spectra leaves eigenvalue errors of ~2e-4 --
        # measured on the 512 x 512 Gram matrix; in fp64 the same sweeps reach 1e-12.  Results come back in G's dtype.
        in_dtype = G.dtype
        G = G.double().contiguous()
        work = torch.empty(2 * n * n + 8, dtype=G.dtype, device=G.device)
        evals = torch.empty(n, dtype=G.dtype, device=G.device)
        evecs = torch.empty(n, n, dtype=G.dtype, device=G.device)
        self.sweeps = torch.zeros(1, dtype=torch.int32, device=G.device)
        _native.check(self.lib.dapb_eigh(G.data_ptr(), n, 1, work.data_ptr(), evals.data_ptr(),
                                         evecs.data_ptr(), float(tol), int(max_sweeps), self.sweeps.data_ptr(),
                                         _native.stream_ptr()), "eigh")
        evals, evecs = evals.to(in_dtype), evecs.to(in_dtype)
        return evals, evecs

    def step(self):
        """One full decomposition: Gram + cross-GPU reduce + on-device Cholesky."""
        G = self.gram()
        return G, self.cholesky(G)
