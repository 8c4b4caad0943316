"""The in-box comparative arm: straightforward PyTorch + NCCL (+ cuBLAS) implementations of the same algorithms.

BASELINE.json's north star asks to match-or-beat "a straightforward PyTorch NCCL(+cuBLAS) implementation of the same
algorithm and config".  Those implementations live next to the fused ones as ``backend="nccl"`` (engines) /
``backend="torch"`` (``run_*`` drivers, also the CPU / gloo path) so that both arms share data layout, partitioning and
sampling; this module is the one place that names them.  ``bench.py`` builds its ``vs_baseline`` numbers from these.

What each arm does per step (kept deliberately competent):

* logistic (``ShardedLogisticEngine(backend="nccl")``, ref ``optimization/ssgd.py:93-105``): Philox mask from the library's
  one-pass mask kernel (identical mini-batches), ``nonzero`` + ``index_select``, two cuBLAS GEMVs, one ``all_reduce`` of
  ``[g | count]``, update on the device.
* k-means (``ShardedKMeans(backend="nccl")``, ref ``machine_learning/k-means.py:55-71``): 0.5 M-row chunks, cuBLAS distance GEMM
  with fp32 output, arg-min, one-hot GEMM for the per-cluster sums, one ``all_reduce``.
* PageRank (``ShardedPageRank(backend="nccl")``, ref ``graph_computation/pagerank.py:50-57``): gather + ``index_add_`` with cached
  row indices, one padded ``all_gather_into_tensor``.
* Gram (``GramDecomposition(backend="nccl")``, ref ``matrix_computation/matrix_decomposition.py:27``): cuBLAS ``A.T @ A`` with fp32
  output, ``all_reduce``, ``torch.linalg.cholesky`` / ``eigh`` (cuSOLVER).
"""
from __future__ import annotations


def logistic_engine(X, y, cfg, **kw):
    from ..models.logistic import ShardedLogisticEngine
    return ShardedLogisticEngine(X, y, cfg, backend="nccl", **kw)


def kmeans_engine(X, k, centers, **kw):
    from ..models.kmeans import ShardedKMeans
    return ShardedKMeans(X, k, centers, backend="nccl", **kw)


def pagerank_engine(graph, cfg, **kw):
    from ..models.pagerank import ShardedPageRank
    return ShardedPageRank(graph, cfg, backend="nccl", **kw)


def gram_decomposition(A, dist=None):
    from ..models.matrix_decomposition import GramDecomposition
    return GramDecomposition(A, dist, backend="nccl")


__all__ = ["logistic_engine", "kmeans_engine", "pagerank_engine", "gram_decomposition"]
