"""Datasets: the reference's defaults and synthetic generators for the benchmark configs.

Reference data layer: sklearn breast-cancer with a 70/30 split seed 0
(``optimization/ssgd.py:71-76``), hard-coded toy arrays (``machine_learning/k-means.py:49-50``,
``graph_computation/pagerank.py:35-38``, ``graph_computation/transitive_closure.py:18``) and a
synthetic exact-rank matrix (``matrix_computation/matrix_decomposition.py:42-44``).
Benchmark configs (BASELINE.json) are synthetic and generated on the device that will own the shard.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

KMEANS_TOY = np.array([[1, 2], [1, 4], [1, 0], [10, 2], [10, 4], [10, 0]])      # k-means.py:49-50
PAGERANK_TOY = [(1, 2), (1, 3), (2, 3), (3, 1)]                                  # pagerank.py:35-38
TC_TOY = [(1, 2), (1, 3), (2, 3), (3, 1)]                                        # transitive_closure.py:18


@dataclass
class LogisticData:
    matrix: np.ndarray        # [n_train, D+2] = [X | 1 | y]   (ssgd.py:83-84)
    X_test_aug: np.ndarray    # [n_test, D+1] = [X_test | 1]   (ssgd.py:107-108)
    y_test: np.ndarray

    @property
    def dim(self) -> int:     # D + 1 (bias column included)
        return self.matrix.shape[1] - 1

    @property
    def X_aug(self) -> np.ndarray:
        return self.matrix[:, :-1]

    @property
    def y(self) -> np.ndarray:
        return self.matrix[:, -1]


def load_breast_cancer_split() -> LogisticData:
    """ssgd.py:71-76,83-84 -- 569x30 fp64, test_size 0.3, random_state 0 -> 398 / 171."""
    from sklearn.datasets import load_breast_cancer
    from sklearn.model_selection import train_test_split
    X, y = load_breast_cancer(return_X_y=True)
    X_train, X_test, y_train, y_test = train_test_split(
        X, y, test_size=0.3, random_state=0, shuffle=True)
    n_train, n_test = X_train.shape[0], X_test.shape[0]
    matrix = np.concatenate([X_train, np.ones((n_train, 1)), y_train.reshape(-1, 1)], axis=1)
    X_test_aug = np.concatenate([X_test, np.ones((n_test, 1))], axis=1)
    return LogisticData(matrix=matrix, X_test_aug=X_test_aug, y_test=y_test)


def synthetic_logistic_np(n: int, d: int, n_test: int = 1024, seed: int = 0,
                          dtype=np.float64) -> LogisticData:
    """Small host-side planted-model problem (tests / scaled-down CPU baseline)."""
    rs = np.random.RandomState(seed)
    w_star = rs.randn(d + 1) / np.sqrt(d)

    def make(m):
        X = rs.randn(m, d).astype(dtype)
        Xa = np.concatenate([X, np.ones((m, 1), dtype=dtype)], axis=1)
        p = 1.0 / (1.0 + np.exp(-(Xa.astype(np.float64) @ w_star)))
        y = (rs.random_sample(m) < p).astype(dtype)
        return Xa, y
    Xa, y = make(n)
    Xt, yt = make(n_test)
    return LogisticData(matrix=np.concatenate([Xa, y[:, None]], axis=1), X_test_aug=Xt,
                        y_test=yt.astype(np.int64))


def synthetic_logistic_device(n_rows: int, dim: int, device, dtype, seed: int = 0,
                              chunk_rows: int = 1 << 20):
    """Per-GPU shard of the 256M x 1024 benchmark problem (BASELINE.json config 2).

    Returns ``(X [n_rows, dim] dtype, y [n_rows] float32, w_star [dim] float32)``.  The last
    column of X is the bias column of ones (the reference appends it, ssgd.py:83).  Filled in
    chunks so the fp32 staging never exceeds ``chunk_rows * dim`` elements."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    w_star = (torch.randn(dim, device=device, generator=torch.Generator(device=device).manual_seed(1234))
              / dim ** 0.5).float()
    X = torch.empty((n_rows, dim), device=device, dtype=dtype)
    y = torch.empty((n_rows,), device=device, dtype=torch.float32)
    for r0 in range(0, n_rows, chunk_rows):
        r1 = min(n_rows, r0 + chunk_rows)
        blk = X[r0:r1]
        blk.normal_(generator=g)
        blk[:, -1] = 1
        z = blk.float() @ w_star
        y[r0:r1] = (torch.rand(r1 - r0, device=device, generator=g) < torch.sigmoid(z)).float()
    return X, y, w_star


def synthetic_clusters_device(n_rows: int, d: int, k: int, device, dtype, seed: int = 0,
                              chunk_rows: int = 1 << 22, spread: float = 4.0):
    """Per-GPU shard for the k-means benchmark (BASELINE.json config 3): k planted Gaussian blobs."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    cg = torch.Generator(device=device).manual_seed(4321)
    centres = torch.randn(k, d, device=device, generator=cg) * spread
    X = torch.empty((n_rows, d), device=device, dtype=dtype)
    for r0 in range(0, n_rows, chunk_rows):
        r1 = min(n_rows, r0 + chunk_rows)
        lab = torch.randint(0, k, (r1 - r0,), device=device, generator=g)
        X[r0:r1] = (centres[lab] + torch.randn(r1 - r0, d, device=device, generator=g)).to(dtype)
    return X, centres


def synthetic_clusters_np(n: int, d: int, k: int, seed: int = 0, spread: float = 4.0):
    rs = np.random.RandomState(seed)
    centres = rs.randn(k, d) * spread
    lab = rs.randint(0, k, n)
    return centres[lab] + rs.randn(n, d), centres


def low_rank_matrix_np(m: int, n: int, k: int, seed: Optional[int] = None):
    """matrix_decomposition.py:42: ``rand(m,k) @ rand(n,k).T``."""
    rs = np.random.RandomState(seed)
    return rs.rand(m, k) @ rs.rand(n, k).T


# ----------------------------------------------------------------------------- graphs
def rmat_edges_np(scale: int, edge_factor: int = 16, seed: int = 0,
                  a: float = 0.57, b: float = 0.19, c: float = 0.19) -> Tuple[np.ndarray, np.ndarray]:
    """Graph500-style R-MAT generator (BASELINE.json config 4), host version for tests."""
    rs = np.random.RandomState(seed)
    n_edges = edge_factor << scale
    src = np.zeros(n_edges, dtype=np.int64)
    dst = np.zeros(n_edges, dtype=np.int64)
    ab, abc = a + b, a + b + c
    for _ in range(scale):
        r = rs.random_sample(n_edges)
        sbit = r >= ab
        dbit = ((r >= a) & (r < ab)) | (r >= abc)
        src = (src << 1) | sbit
        dst = (dst << 1) | dbit
    return src, dst


def rmat_edges_device(scale: int, edge_factor: int, device, seed: int = 0,
                      a: float = 0.57, b: float = 0.19, c: float = 0.19,
                      n_edges: Optional[int] = None, edge_offset_seed: int = 0):
    """R-MAT on the device with torch ops (int64 src/dst).  ``n_edges`` lets each rank generate
    only its share; ``edge_offset_seed`` decorrelates the shares."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed * 1000003 + edge_offset_seed)
    if n_edges is None:
        n_edges = edge_factor << scale
    src = torch.zeros(n_edges, dtype=torch.int64, device=device)
    dst = torch.zeros(n_edges, dtype=torch.int64, device=device)
    ab, abc = a + b, a + b + c
    for _ in range(scale):
        r = torch.rand(n_edges, device=device, generator=g)
        sbit = (r >= ab).to(torch.int64)
        dbit = (((r >= a) & (r < ab)) | (r >= abc)).to(torch.int64)
        src = (src << 1) | sbit
        dst = (dst << 1) | dbit
    return src, dst


def random_graph_np(n_vertices: int, n_edges: int, seed: int = 0):
    """Uniform random directed graph (transitive-closure benchmark, BASELINE.json config 4)."""
    rs = np.random.RandomState(seed)
    return rs.randint(0, n_vertices, n_edges), rs.randint(0, n_vertices, n_edges)
