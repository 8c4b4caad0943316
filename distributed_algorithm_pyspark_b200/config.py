"""Hyper-parameter records.  Field names and defaults are the module-level
constants of the reference scripts, so that the same-named
entry scripts can expose them unchanged.
"""
from __future__ import annotations

from dataclasses import dataclass, asdict
from typing import Optional


@dataclass
class LogisticConfig:
    """Shared knobs of the five logistic-regression optimisers.

    ``algo``: ``ssgd`` (optimization/ssgd.py:17-21), ``gd``
    (machine_learning/logistic_regression.py:17-19), ``ma`` (optimization/ma.py:19-23),
    ``bmuf`` (optimization/bmuf.py:19-25), ``easgd`` (optimization/easgd.py:19-25).
    """
    algo: str = "ssgd"
    n_slices: int = 4
    n_iterations: int = 1500
    eta: float = 0.1
    mini_batch_fraction: float = 0.1
    lam: float = 0.0                 # ssgd.py:21
    reg_type: str = "l2"             # ssgd.py:105 always passes "l2"
    reg_alpha: float = 0.0           # elastic-net mixing (ssgd.py:36)
    n_local_iterations: int = 5      # ma.py:23 / bmuf.py:23
    mu: float = 0.9                  # bmuf.py:24
    zeta: float = 0.1                # bmuf.py:25
    rho: float = 0.1                 # easgd.py:23
    sample_seed: int = 42            # ssgd.py:97  (seed = 42 + t)
    sampler: str = "spark"           # "spark" (MT19937, parity) or "philox" (device)
    seed: Optional[int] = None       # init of w / replicas / delta_w (reference: unseeded)

    @property
    def sigmoid_eps(self) -> float:
        # ssgd.py:24 and logistic_regression.py:22 use the plain sigmoid;
        # ma.py:26, bmuf.py:28, easgd.py:28 add 1e-6 to the denominator.
        return 0.0 if self.algo in ("ssgd", "gd") else 1e-6

    @property
    def alpha(self) -> float:        # easgd.py:24
        return self.eta * self.rho

    @property
    def beta(self) -> float:         # easgd.py:25
        return self.n_slices * self.alpha

    @staticmethod
    def for_algo(algo: str, **kw) -> "LogisticConfig":
        defaults = {
            "ssgd": dict(n_iterations=1500),
            "gd": dict(n_iterations=1500, mini_batch_fraction=1.0),
            "ma": dict(n_iterations=300),
            "bmuf": dict(n_iterations=300),
            "easgd": dict(n_iterations=1500, n_local_iterations=1),
        }[algo]
        defaults.update(kw)
        return LogisticConfig(algo=algo, **defaults)

    def to_dict(self):
        d = asdict(self)
        d.update(sigmoid_eps=self.sigmoid_eps, alpha=self.alpha, beta=self.beta)
        return d


@dataclass
class KMeansConfig:              # machine_learning/k-means.py:15-18
    k: int = 2
    convergeDist: float = 0.1    # defined but never used by the reference (k-means.py:16)
    n_slices: int = 2
    n_iterations: int = 5
    seed: int = 42               # takeSample seed (k-means.py:53)


@dataclass
class PageRankConfig:            # graph_computation/pagerank.py:17-19
    n_slices: int = 3
    n_iterations: int = 10
    q: float = 0.15


@dataclass
class TransitiveClosureConfig:   # graph_computation/transitive_closure.py:10
    n_slices: int = 2
    max_rounds: int = 1 << 20


@dataclass
class ALSConfig:                 # matrix_computation/matrix_decomposition.py:12-17
    lam: float = 0.01
    m: int = 100
    n: int = 500
    k: int = 10
    n_iterations: int = 5
    n_slices: int = 4
    seed: Optional[int] = None


@dataclass
class MonteCarloConfig:          # randomized_algorithm/monte_carlo.py:13-15
    n_slices: int = 4
    n: int = 100000 * 4
    seed: Optional[int] = None
