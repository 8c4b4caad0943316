"""Oracle vs the outputs the reference records in trailing comments."""
import numpy as np
import pytest

from distributed_algorithm_pyspark_b200.config import (ALSConfig, KMeansConfig, LogisticConfig, MonteCarloConfig,
                                                        PageRankConfig)
from distributed_algorithm_pyspark_b200.data.datasets import KMEANS_TOY, PAGERANK_TOY, TC_TOY
from distributed_algorithm_pyspark_b200.oracle import spark_semantics as sp
from distributed_algorithm_pyspark_b200.oracle.algorithms import (LogisticInit, als_draw, als_oracle, kmeans_oracle,
                                                                   logistic_oracle, monte_carlo_oracle, pagerank_oracle,
                                                                   pagerank_vector_oracle, reg_gradient,
                                                                   transitive_closure_oracle, ewma_smooth)


def test_parallelize_slices_match_spark():
    assert [b - a for a, b in sp.parallelize_slices(398, 4)] == [99, 99, 99, 101]
    assert sp.parallelize_slices(6, 2) == [(0, 3), (3, 6)]
    assert sp.parallelize_range_slices(400000, 4)[1] == (100000, 200000)
    for n, s in [(1, 4), (7, 3), (5000, 4), (3, 8)]:
        sl = sp.parallelize_slices(n, s)
        assert sl[0][0] == 0 and sl[-1][1] == n and all(a <= b for a, b in sl)
        assert all(sl[i][1] == sl[i + 1][0] for i in range(s - 1))


def test_sampler_is_deterministic_and_about_ten_percent():
    sizes = [99, 99, 99, 101]
    m1 = sp.bernoulli_sample_mask(sizes, 0.1, 42)
    m2 = sp.bernoulli_sample_mask(sizes, 0.1, 42)
    assert (m1 == m2).all() and m1.sum() == 40          # SSGD's first mini-batch (A.8: mean 40.56)
    counts = [sp.bernoulli_sample_mask(sizes, 0.1, 42 + t).sum() for t in range(200)]
    assert 25 <= min(counts) and max(counts) <= 62 and 36 < np.mean(counts) < 45


def test_take_sample_gives_reference_initial_centres():
    idx = sp.take_sample_indices([3, 3], 2, 42)
    assert idx == [3, 1]                                  # -> centres [10,2], [1,4] (A.5)


def test_pagerank_bit_exact_with_reference_comment():
    r = pagerank_oracle(PAGERANK_TOY, PageRankConfig())
    assert r == {1: 0.38891305880091237, 2: 0.214416470596171, 3: 0.3966704706029163}   # pagerank.py:66-68


def test_pagerank_vector_form_equals_message_form():
    rs = np.random.RandomState(1)
    e = sorted(set(zip(rs.randint(0, 40, 120).tolist(), rs.randint(0, 40, 120).tolist())))
    a = pagerank_oracle(e, PageRankConfig())
    s, d = np.array([x for x, _ in e]), np.array([y for _, y in e])
    r, present = pagerank_vector_oracle(s, d, 40, PageRankConfig())
    assert set(a) == set(np.flatnonzero(present).tolist())
    assert max(abs(a[v] - r[v]) for v in a) < 1e-15


def test_pagerank_vertex_without_in_edges_disappears():
    r = pagerank_oracle([(0, 1), (1, 2), (2, 1)], PageRankConfig(n_iterations=3))
    assert 0 not in r and set(r) == {1, 2}


def test_transitive_closure_toy_is_nine_pairs():
    n, rounds, pairs = transitive_closure_oracle(TC_TOY)
    assert n == 9 and rounds == 3 and (1, 1) in pairs


def test_kmeans_toy_centres():
    res = kmeans_oracle(KMEANS_TOY, KMeansConfig())
    assert res.init_indices == [3, 1]
    np.testing.assert_array_equal(res.centers, [[10, 2], [1, 2]])
    np.testing.assert_array_equal(res.assignment, [1, 1, 1, 0, 0, 0])


def test_kmeans_empty_cluster_keeps_centre_and_tie_goes_low():
    pts = np.array([[0.0, 0.0], [2.0, 0.0]])
    res = kmeans_oracle(pts, KMeansConfig(k=3, n_iterations=2), init_centers=np.array([[1.0, 0.0], [1.0, 0.0], [50.0, 50.0]]))
    np.testing.assert_array_equal(res.centers[2], [50.0, 50.0])      # never assigned
    np.testing.assert_array_equal(res.assignment, [0, 0])            # tie between centre 0 and 1 -> 0


def test_als_rmse_trajectory_seed0():
    cfg = ALSConfig(seed=0)
    R, U, V = als_draw(cfg)
    _, _, errs = als_oracle(R, U, V, cfg)
    np.testing.assert_allclose(errs, [0.2245, 0.0768, 0.0315, 0.0314, 0.0315], atol=1e-4)
    Ub, Vb, errs_b = als_oracle(R, U, V, cfg, batched=True)
    np.testing.assert_allclose(errs, errs_b, rtol=1e-9)


def test_monte_carlo_within_four_sigma():
    pi, count = monte_carlo_oracle(MonteCarloConfig(n=40000, seed=5))
    sigma = 4 * np.sqrt((np.pi / 4) * (1 - np.pi / 4) / 40000)
    assert abs(pi - np.pi) < 4 * sigma and count == round(pi * 40000 / 4)


def test_reg_gradient_variants():
    w = np.array([-2.0, 0.0, 3.0])
    assert reg_gradient(w, "none") == 0
    np.testing.assert_array_equal(reg_gradient(w, "l2"), w)
    np.testing.assert_array_equal(reg_gradient(w, "l1"), [-1, 0, 1])
    np.testing.assert_allclose(reg_gradient(w, "elastic_net", 0.25), 0.25 * np.sign(w) + 0.75 * w)
    with pytest.raises(AssertionError):
        reg_gradient(w, "bogus")


def test_ewma():
    np.testing.assert_allclose(ewma_smooth([1.0, 0.0, 0.0]), [1.0, 0.9, 0.81])


@pytest.mark.parametrize("algo,lead,rtol,acc_lo", [
    ("easgd", [44.10, 68.78, 259.53, 144.00], 0.03, 0.85),      # easgd.py:123
    ("bmuf", [34.15, 51.14, 204.08, 103.63], 0.25, 0.85),       # bmuf.py:131 (random block momentum: looser)
])
def test_logistic_oracle_reaches_reference_weights(breast_cancer, algo, lead, rtol, acc_lo):
    """Large-magnitude weights are insensitive to the (unseeded) init: within a few % of the comments."""
    d = breast_cancer
    cfg = LogisticConfig.for_algo(algo)
    init = LogisticInit.draw(algo, d.dim, cfg.n_slices, 0)
    res = logistic_oracle(d.matrix, d.X_test_aug, d.y_test, cfg, init)
    np.testing.assert_allclose(res.w[:4], lead, rtol=rtol)
    assert max(res.accs[-50:]) >= acc_lo
