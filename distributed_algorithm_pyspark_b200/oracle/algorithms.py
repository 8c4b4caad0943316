"""NumPy / stdlib oracle: the ten reference scripts restated with Spark's exact
partitioning, sampling and fold-order rules.

The oracle is the numerical ground truth for every other backend and the timed
CPU baseline when PySpark is not installable.  All
randomness the reference leaves unseeded (``ssgd.py:89``, ``ma.py:86,89``,
``bmuf.py:95``, ``matrix_decomposition.py:42-44``, ``monte_carlo.py:18``) is injectable.
"""
from __future__ import annotations

import random
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from ..config import (ALSConfig, KMeansConfig, LogisticConfig, MonteCarloConfig,
                      PageRankConfig)
from . import spark_semantics as sp


# =============================================================================
# Logistic regression family
# =============================================================================
def logistic_f(x: np.ndarray, w: np.ndarray, eps: float = 0.0) -> np.ndarray:
    """``1/(exp(-x.w)+1[+1e-6])`` -- ssgd.py:23-24 (eps=0) / ma.py:25-26 (eps=1e-6).
    ``exp`` may overflow to inf for the un-normalised default data; the result is then 0."""
    with np.errstate(over="ignore"):
        return 1.0 / (np.exp(-x.dot(w)) + 1.0 + eps)


def point_gradient(point: np.ndarray, w: np.ndarray, eps: float = 0.0) -> np.ndarray:
    """ssgd.py:27-33: ``-(y - sigma(x.w)) * x`` with ``point = [x | 1 | y]``."""
    y = point[-1]
    x = point[:-1]
    return -(y - logistic_f(x, w, eps)) * x


def reg_gradient(w: np.ndarray, reg_type: str = "l2", alpha: float = 0.0):
    """ssgd.py:36-47."""
    assert reg_type in ["none", "l2", "l1", "elastic_net"]
    if reg_type == "none":
        return 0
    if reg_type == "l2":
        return w
    if reg_type == "l1":
        return np.sign(w)
    return alpha * np.sign(w) + (1 - alpha) * w


def accuracy(X_test_aug: np.ndarray, y_test: np.ndarray, w: np.ndarray, eps: float = 0.0) -> float:
    """ssgd.py:107-110: threshold 0.5 on the (variant) sigmoid, fraction correct."""
    y_pred = logistic_f(X_test_aug, w, eps)
    pred = np.where(y_pred < 0.5, 0, 1)
    return float(np.mean(pred == y_test))


def ewma_smooth(accs: Sequence[float], alpha: float = 0.9) -> np.ndarray:
    """ssgd.py:51-59."""
    s = np.zeros(len(accs))
    for i, a in enumerate(accs):
        s[i] = a if i == 0 else alpha * s[i - 1] + (1 - alpha) * a
    return s


@dataclass
class LogisticInit:
    """Everything the reference draws from the unseeded global NumPy RNG, in draw order."""
    w: np.ndarray                              # ssgd.py:89 / ma.py:89
    ws: Optional[np.ndarray] = None            # ma.py:86 (dead for ma/bmuf, live for easgd.py:88)
    delta_w: Optional[np.ndarray] = None       # bmuf.py:95

    @staticmethod
    def draw(algo: str, dim: int, n_slices: int, seed: Optional[int]) -> "LogisticInit":
        rs = np.random.RandomState(seed)
        ws = delta = None
        if algo in ("ma", "bmuf", "easgd"):
            ws = 2 * rs.random_sample((n_slices, dim)) - 1     # drawn before w (ma.py:86,89)
        w = 2 * rs.random_sample(dim) - 1
        if algo == "bmuf":
            delta = 2 * rs.random_sample(dim) - 1
        return LogisticInit(w=w, ws=ws, delta_w=delta)


@dataclass
class LogisticResult:
    w: np.ndarray
    accs: List[float]
    batch_sizes: List[int] = field(default_factory=list)
    ws: Optional[np.ndarray] = None
    delta_w: Optional[np.ndarray] = None

    @property
    def final_acc(self) -> float:
        return self.accs[-1] if self.accs else float("nan")


def logistic_oracle(matrix: np.ndarray, X_test_aug: np.ndarray, y_test: np.ndarray,
                    cfg: LogisticConfig, init: LogisticInit,
                    masks: Optional[np.ndarray] = None,
                    callback=None) -> LogisticResult:
    """Run one of ssgd / gd / ma / bmuf / easgd exactly as the reference drives it.

    ``matrix`` is ``[X_train | 1 | y]`` (ssgd.py:83-84).  ``masks`` (bool [T, N]) overrides
    the Spark sampler (used to feed Philox masks when cross-checking the device sampler).
    """
    algo = cfg.algo
    n, dim = matrix.shape[0], matrix.shape[1] - 1
    offs = sp.partition_offsets(n, cfg.n_slices)
    sizes = np.diff(offs)
    eps = cfg.sigmoid_eps
    w = np.array(init.w, dtype=np.float64, copy=True)
    ws = None if init.ws is None else np.array(init.ws, dtype=np.float64, copy=True)
    delta_w = None if init.delta_w is None else np.array(init.delta_w, dtype=np.float64, copy=True)
    accs: List[float] = []
    bsz: List[int] = []

    def mask_for(t):
        if algo == "gd":
            return np.ones(n, dtype=bool)
        if masks is not None:
            return masks[t]
        return sp.bernoulli_sample_mask(sizes, cfg.mini_batch_fraction, cfg.sample_seed + t)

    for t in range(cfg.n_iterations):
        keep = mask_for(t)
        if algo in ("ssgd", "gd"):
            # treeAggregate: per-partition left fold from the zero value, partials folded in
            # partition order on the driver (ssgd.py:99-103, logistic_regression.py:82-83).
            g_tot, cnt_tot = None, 0
            for p in range(cfg.n_slices):
                g_p, c_p = 0.0, 0
                for i in range(offs[p], offs[p + 1]):
                    if keep[i]:
                        g_p = g_p + point_gradient(matrix[i], w, eps)
                        c_p += 1
                g_tot = g_p if g_tot is None else g_tot + g_p
                cnt_tot += c_p
            bsz.append(cnt_tot)
            if algo == "gd":
                w = w - cfg.eta * g_tot                       # logistic_regression.py:85 (SUM)
            elif cnt_tot > 0:
                w = w - cfg.eta * (g_tot / cnt_tot + cfg.lam * reg_gradient(w, cfg.reg_type, cfg.reg_alpha))
        else:
            centre = w
            if algo in ("ma", "bmuf"):
                ws = np.tile(w, (cfg.n_slices, 1))            # ma.py:96 reset replicas
            n_local = cfg.n_local_iterations if algo != "easgd" else 1
            cnt_tot = int(np.count_nonzero(keep))             # mini-batch size of the round
            for _ in range(n_local):                          # same mini-batch every local step (ma.py:99)
                for p in range(cfg.n_slices):
                    rows = [i for i in range(offs[p], offs[p + 1]) if keep[i]]
                    if not rows:                              # reference raises IndexError (ma.py:41)
                        continue
                    g = np.array([point_gradient(matrix[i], ws[p], eps) for i in rows])
                    g_mean = np.mean(g, axis=0)               # ma.py:42
                    if algo == "easgd":                       # easgd.py:45
                        ws[p] = ws[p] - cfg.eta * g_mean - cfg.alpha * (ws[p] - centre)
                    else:                                     # ma.py:43
                        ws[p] = ws[p] - cfg.eta * g_mean
            bsz.append(cnt_tot)
            par_w_sum = 0.0
            for p in range(cfg.n_slices):                     # treeAggregate(0.0, add, add) (ma.py:104)
                q = 0.0 + ws[p]
                par_w_sum = q if p == 0 else par_w_sum + q
            if algo == "ma":
                w = par_w_sum / cfg.n_slices                  # ma.py:106
            elif algo == "bmuf":
                w_avg = par_w_sum / cfg.n_slices              # bmuf.py:111-114
                delta_w = cfg.mu * delta_w + cfg.zeta * (w_avg - w)
                w = w + delta_w
            else:                                             # easgd.py:106
                w = (1 - cfg.beta) * w + cfg.beta * par_w_sum / cfg.n_slices
        acc = accuracy(X_test_aug, y_test, w, eps)
        accs.append(acc)
        if callback is not None:
            callback(t, w, acc)
    return LogisticResult(w=w, accs=accs, batch_sizes=bsz, ws=ws, delta_w=delta_w)


# =============================================================================
# k-means  (machine_learning/k-means.py)
# =============================================================================
def closest_center(p: np.ndarray, centers: Sequence[np.ndarray]) -> int:
    """k-means.py:20-28: strict '<' so the lowest index wins ties."""
    closest, best = 0, float("+inf")
    for cid in range(len(centers)):
        dist = np.sqrt(np.sum((p - centers[cid]) ** 2))
        if dist < best:
            best, closest = dist, cid
    return closest


@dataclass
class KMeansResult:
    centers: np.ndarray
    assignment: np.ndarray
    init_indices: List[int]
    history: List[np.ndarray] = field(default_factory=list)


def kmeans_oracle(points: np.ndarray, cfg: KMeansConfig,
                  init_centers: Optional[np.ndarray] = None) -> KMeansResult:
    pts = np.asarray(points)
    n = pts.shape[0]
    sizes = np.diff(sp.partition_offsets(n, cfg.n_slices))
    if init_centers is None:
        idx = sp.take_sample_indices(sizes, cfg.k, cfg.seed)          # k-means.py:53
        centers = [pts[i].astype(np.float64) for i in idx]
    else:
        idx = []
        centers = [np.array(c, dtype=np.float64) for c in init_centers]
    history = []
    assign = np.zeros(n, dtype=np.int64)
    for _ in range(cfg.n_iterations):
        sums: Dict[int, np.ndarray] = {}
        cnts: Dict[int, int] = {}
        for i in range(n):                                            # k-means.py:57-63
            c = closest_center(pts[i], centers)
            assign[i] = c
            if c in sums:
                sums[c] = sums[c] + pts[i]
                cnts[c] += 1
            else:
                sums[c] = pts[i].astype(np.float64)
                cnts[c] = 1
        for c in sums:                                                # k-means.py:66-71
            centers[c] = sums[c] / cnts[c]                            # empty cluster keeps old centre
        history.append(np.array(centers))
    # k-means.py:76 re-assigns with the FINAL centres when collecting for the plot (A.7)
    final_assign = np.array([closest_center(pts[i], centers) for i in range(n)], dtype=np.int64)
    return KMeansResult(centers=np.array(centers), assignment=final_assign,
                        init_indices=list(idx), history=history)


# =============================================================================
# PageRank  (graph_computation/pagerank.py)
# =============================================================================
def pagerank_oracle(edges: Iterable[Tuple[int, int]], cfg: PageRankConfig) -> Dict[int, float]:
    """Message-passing form with Python floats, mirroring pagerank.py:41-57.

    n = number of distinct SOURCE vertices (:44); ranks exist only for vertices that
    received a contribution in the previous iteration (:52 inner join, :57)."""
    adj: Dict[int, List[int]] = {}
    for s, d in sorted(set((int(a), int(b)) for a, b in edges)):      # distinct (:41)
        adj.setdefault(s, []).append(d)
    n = len(adj)
    if n == 0:
        return {}
    ranks = {v: 1.0 / n for v in adj}
    for _ in range(cfg.n_iterations):
        contribs: Dict[int, float] = {}
        for v in sorted(adj):
            if v not in ranks:
                continue
            share = ranks[v] / len(adj[v])
            for d in adj[v]:
                contribs[d] = contribs[d] + share if d in contribs else share
        ranks = {v: cfg.q / n + (1 - cfg.q) * s for v, s in contribs.items()}
    return dict(sorted(ranks.items()))


def pagerank_vector_oracle(src: np.ndarray, dst: np.ndarray, n_vertices: int,
                           cfg: PageRankConfig) -> Tuple[np.ndarray, np.ndarray]:
    """Dense-vector restatement used at scale:
    ``s = A^T (r / outdeg)``; ``r' = where(s>0, q/n + (1-q) s, 0)``.
    Returns (ranks, present_mask) over vertex ids 0..n_vertices-1; edges must be de-duplicated."""
    src = np.asarray(src, dtype=np.int64)
    dst = np.asarray(dst, dtype=np.int64)
    outdeg = np.bincount(src, minlength=n_vertices).astype(np.float64)
    n = int(np.count_nonzero(outdeg))
    r = np.where(outdeg > 0, 1.0 / max(n, 1), 0.0)
    present = outdeg > 0
    for _ in range(cfg.n_iterations):
        c = np.where((outdeg > 0) & present, r / np.maximum(outdeg, 1.0), 0.0)
        s = np.zeros(n_vertices)
        np.add.at(s, dst, c[src])
        present = s > 0
        r = np.where(present, cfg.q / n + (1 - cfg.q) * s, 0.0)
    return r, present


# =============================================================================
# Transitive closure  (graph_computation/transitive_closure.py)
# =============================================================================
def transitive_closure_oracle(edges: Iterable[Tuple[int, int]]) -> Tuple[int, int, set]:
    """Returns (pair count, rounds, closure set).  Mirrors :29-40: each round joins
    *all* current paths with the edges, unions, de-duplicates, stops when the count is unchanged."""
    paths = set((int(a), int(b)) for a, b in edges)
    by_dst: Dict[int, List[int]] = {}
    for x, y in paths:
        by_dst.setdefault(y, []).append(x)      # edges reversed (y, x) (:27)
    rounds = 0
    while True:
        old = len(paths)
        new = set()
        for (y, z) in paths:                    # path (y,z) joined with edge (x,y) -> (x,z) (:35)
            for x in by_dst.get(y, ()):
                new.add((x, z))
        paths |= new
        rounds += 1
        if len(paths) == old:
            break
    return len(paths), rounds, paths


# =============================================================================
# ALS matrix decomposition  (matrix_computation/matrix_decomposition.py)
# =============================================================================
def rmse(R: np.ndarray, U: np.ndarray, V: np.ndarray) -> float:
    """matrix_decomposition.py:19-21 (divides by m*n)."""
    diff = R - U @ V.T
    return float(np.sqrt(np.sum(np.power(diff, 2)) / (R.shape[0] * R.shape[1])))


def als_update_row(i: int, mat: np.ndarray, ratings: np.ndarray, lam: float) -> np.ndarray:
    """matrix_decomposition.py:24-33: ridge normal equations for row i, ridge = lam * mat.shape[0]."""
    k = mat.shape[1]
    XtX = mat.T @ mat
    Xty = mat.T @ ratings[i, :].T
    for j in range(k):
        XtX[j, j] += lam * mat.shape[0]
    return np.linalg.solve(XtX, Xty)


def als_draw(cfg: ALSConfig):
    """R, U, V in the reference's draw order (matrix_decomposition.py:42-44)."""
    rs = np.random.RandomState(cfg.seed)
    R = rs.rand(cfg.m, cfg.k) @ rs.rand(cfg.n, cfg.k).T
    U = rs.rand(cfg.m, cfg.k)
    V = rs.rand(cfg.n, cfg.k)
    return R, U, V


def als_oracle(R: np.ndarray, U: np.ndarray, V: np.ndarray, cfg: ALSConfig, batched: bool = False):
    """Returns (U, V, [rmse per iteration]).  ``batched=True`` uses the algebraically
    identical ``U = R V (V^T V + lam n I)^-1``."""
    m, n = R.shape
    errs = []
    for _ in range(cfg.n_iterations):
        if batched:
            G = V.T @ V + cfg.lam * V.shape[0] * np.eye(V.shape[1])
            U = np.linalg.solve(G, (R @ V).T).T
            G = U.T @ U + cfg.lam * U.shape[0] * np.eye(U.shape[1])
            V = np.linalg.solve(G, (R.T @ U).T).T
        else:
            U = np.array([als_update_row(i, V, R, cfg.lam) for i in range(m)])      # :52-57
            V = np.array([als_update_row(j, U, R.T, cfg.lam) for j in range(n)])    # :60-63
        errs.append(rmse(R, U, V))
    return U, V, errs


def gram_eig_oracle(A: np.ndarray):
    """Benchmark-config decomposition (BASELINE.json config 5): Gram ``A^T A`` then a
    symmetric eigen-decomposition; singular values of A are sqrt of the eigenvalues."""
    G = A.T.astype(np.float64) @ A.astype(np.float64)
    evals, evecs = np.linalg.eigh(G)
    return G, evals[::-1], evecs[:, ::-1]


# =============================================================================
# Monte-Carlo pi  (randomized_algorithm/monte_carlo.py)
# =============================================================================
def monte_carlo_oracle(cfg: MonteCarloConfig) -> Tuple[float, int]:
    """monte_carlo.py:17-31 with one seeded stdlib generator per partition."""
    count = 0
    for split, (a, b) in enumerate(sp.parallelize_range_slices(cfg.n, cfg.n_slices)):
        rng = random.Random(None if cfg.seed is None else cfg.seed + split)
        c = 0
        for _ in range(a, b):
            x = rng.random() * 2 - 1
            y = rng.random() * 2 - 1
            c += 1 if x ** 2 + y ** 2 <= 1 else 0
        count += c
    return 4.0 * count / cfg.n, count
