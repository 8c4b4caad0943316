"""Property tests (hypothesis) of the torch backends against the NumPy oracle on random small inputs."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from distributed_algorithm_pyspark_b200.config import KMeansConfig, PageRankConfig
from distributed_algorithm_pyspark_b200.models.kmeans import run_kmeans
from distributed_algorithm_pyspark_b200.models.pagerank import run_pagerank
from distributed_algorithm_pyspark_b200.models.transitive_closure import run_transitive_closure
from distributed_algorithm_pyspark_b200.oracle.algorithms import (kmeans_oracle, pagerank_oracle,
                                                                   transitive_closure_oracle)
from distributed_algorithm_pyspark_b200.oracle import spark_semantics as sp
from distributed_algorithm_pyspark_b200.utils.philox import philox_row_mask_np

edges = st.lists(st.tuples(st.integers(0, 12), st.integers(0, 12)), min_size=1, max_size=40)
common = dict(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.too_slow])


@settings(**common)
@given(edges)
def test_transitive_closure_matches_oracle(e):
    n, rounds, pairs = transitive_closure_oracle(e)
    got = run_transitive_closure(e, backend="torch")
    assert got.n_pairs == n and got.rounds == rounds
    assert sorted(map(tuple, got.pairs.tolist())) == sorted(pairs)


@settings(**common)
@given(edges, st.integers(0, 6))
def test_pagerank_matches_oracle(e, iters):
    cfg = PageRankConfig(n_iterations=iters)
    ref = pagerank_oracle(e, cfg)
    got = run_pagerank(e, cfg, backend="torch")
    assert set(got) == set(ref)
    for v in ref:
        assert abs(got[v] - ref[v]) <= 1e-12 * max(1.0, abs(ref[v]))


@settings(**common)
@given(st.integers(6, 40), st.integers(1, 4), st.integers(2, 5), st.integers(0, 2 ** 31 - 1))
def test_kmeans_matches_oracle(n, d, k, seed):
    rs = np.random.RandomState(seed)
    pts = rs.randint(-5, 6, size=(n, d)).astype(np.float64)        # integer grid: plenty of exact ties
    cfg = KMeansConfig(k=min(k, n), n_slices=2, n_iterations=3)
    ref = kmeans_oracle(pts, cfg)
    got = run_kmeans(pts, cfg, backend="torch")
    np.testing.assert_allclose(got.centers, ref.centers, rtol=1e-12, atol=1e-12)
    assert (got.assignment == ref.assignment).all()


@settings(**common)
@given(st.integers(1, 3000), st.integers(1, 9))
def test_slices_partition_every_row_exactly_once(n, s):
    sl = sp.parallelize_slices(n, s)
    assert sl[0][0] == 0 and sl[-1][1] == n
    assert all(a <= b for a, b in sl) and all(sl[i][1] == sl[i + 1][0] for i in range(s - 1))


@settings(**common)
@given(st.integers(1, 5000), st.integers(0, 2 ** 40), st.integers(0, 1000))
def test_philox_mask_is_offset_consistent(n, seed, it):
    """A shard's mask is the corresponding slice of the global mask (what makes results GPU-count independent)."""
    full = philox_row_mask_np(n, 0.3, seed, it, 0)
    cut = n // 3
    assert (philox_row_mask_np(n - cut, 0.3, seed, it, cut) == full[cut:]).all()
