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


@given(st.lists(st.integers(0, 40), min_size=1, max_size=60), st.integers(0, 30), st.integers(1, 17))
@settings(max_examples=60, deadline=None, suppress_health_check=list(HealthCheck))
def test_hub_split_is_a_lossless_rearrangement(lens, threshold, chunk):
    """split_hub_rows (PageRank warp layout): main CSR + hub rows/chunks together hold every non-zero exactly once, in
    row order, and the chunks tile each hub row without gaps."""
    import torch
    from distributed_algorithm_pyspark_b200.models.pagerank import split_hub_rows
    lens_t = torch.tensor(lens, dtype=torch.int64)
    row_ptr = torch.zeros(len(lens) + 1, dtype=torch.int64)
    row_ptr[1:] = torch.cumsum(lens_t, 0)
    col = torch.arange(int(row_ptr[-1]), dtype=torch.int32)            # non-zero e carries the id e
    rp, main_col, hub = split_hub_rows(row_ptr, col, threshold, chunk)
    hub_rows = set(hub.rows.tolist())
    assert hub_rows == {i for i, l in enumerate(lens) if l > threshold}
    rebuilt = {}
    for r in range(len(lens)):
        if r in hub_rows:
            assert int(rp[r + 1] - rp[r]) == 0
        else:
            rebuilt[r] = main_col[int(rp[r]):int(rp[r + 1])].tolist()
    for i, r in enumerate(hub.rows.tolist()):
        a, b = int(hub.chunk_ptr[i]), int(hub.chunk_ptr[i + 1])
        assert b > a and int(hub.chunk_start[a]) == int(hub.ptr[i]) and int(hub.chunk_end[b - 1]) == int(hub.ptr[i + 1])
        assert all(int(hub.chunk_end[c]) == int(hub.chunk_start[c + 1]) for c in range(a, b - 1))
        assert all(0 < int(hub.chunk_end[c] - hub.chunk_start[c]) <= chunk for c in range(a, b))
        rebuilt[r] = hub.col[int(hub.ptr[i]):int(hub.ptr[i + 1])].tolist()
    for r in range(len(lens)):
        assert rebuilt[r] == list(range(int(row_ptr[r]), int(row_ptr[r + 1])))


@given(st.lists(st.integers(0, 500), min_size=1, max_size=200), st.integers(1, 8), st.floats(0, 64), st.floats(0.1, 2.0))
@settings(max_examples=80, deadline=None, suppress_health_check=list(HealthCheck))
def test_vertex_bounds_cover_the_range_and_balance_cost(indeg, world, vertex_cost, hub_edge_cost):
    """balanced_vertex_bounds: monotone cuts from 0 to V, and no rank's modelled cost exceeds the ideal share by more than
    the cost of one vertex (cuts fall on vertex boundaries)."""
    import torch
    from distributed_algorithm_pyspark_b200.models.pagerank import balanced_vertex_bounds
    deg = torch.tensor(indeg, dtype=torch.int64)
    b = balanced_vertex_bounds(deg, world, vertex_cost=vertex_cost, hub_threshold=64, hub_edge_cost=hub_edge_cost)
    assert len(b) == world + 1 and b[0] == 0 and b[-1] == len(indeg)
    assert all(b[i] <= b[i + 1] for i in range(world))
    d = deg.double()
    cost = torch.where(deg > 64, d * hub_edge_cost, d) + vertex_cost * (deg > 0) + 1
    total, biggest = float(cost.sum()), float(cost.max())
    for r in range(world):
        assert float(cost[b[r]:b[r + 1]].sum()) <= total / world + biggest + 1e-6


# ------------------------------------------------------------------------------------------ PageRank host-side layout
@settings(**common)
@given(st.integers(0, 2 ** 31 - 1), st.integers(2, 8), st.integers(1, 40), st.integers(4, 64))
def test_hub_split_and_partition_invariants(seed, world, threshold, chunk):
    """split_hub_rows: every non-zero lands exactly once in the main CSR or in a hub chunk, chunks tile their rows in order
    and know their row; balanced_vertex_bounds: contiguous cover of [0, V), modelled cost of every rank within one vertex
    of the mean."""
    import torch
    from distributed_algorithm_pyspark_b200.models.pagerank import balanced_vertex_bounds, split_hub_rows
    g = torch.Generator().manual_seed(seed)
    V = 200
    lens = (torch.rand(V, generator=g) ** 4 * 120).long()
    lens[torch.rand(V, generator=g) < 0.3] = 0
    row_ptr = torch.zeros(V + 1, dtype=torch.int64)
    row_ptr[1:] = torch.cumsum(lens, 0)
    col = torch.randint(0, V, (int(row_ptr[-1]),), generator=g).to(torch.int32)
    rp, main_col, hub = split_hub_rows(row_ptr, col, threshold, chunk)
    main_lens = rp[1:] - rp[:-1]
    is_hub = lens > threshold
    assert torch.equal(main_lens[~is_hub], lens[~is_hub]) and int(main_lens[is_hub].sum()) == 0
    assert torch.equal(hub.rows.long(), torch.nonzero(is_hub).flatten())
    assert main_col.numel() + hub.col.numel() == col.numel()
    if hub.rows.numel():
        assert torch.equal(hub.ptr[1:] - hub.ptr[:-1], lens[is_hub])
        # chunks: consecutive, at most `chunk` long, each inside its own row, chunk_row consistent with chunk_ptr
        for i in range(hub.rows.numel()):
            c0, c1 = int(hub.chunk_ptr[i]), int(hub.chunk_ptr[i + 1])
            assert c1 > c0 and int(hub.chunk_start[c0]) == int(hub.ptr[i]) and int(hub.chunk_end[c1 - 1]) == int(hub.ptr[i + 1])
            for ch in range(c0, c1):
                assert int(hub.chunk_row[ch]) == i and 0 < int(hub.chunk_end[ch] - hub.chunk_start[ch]) <= chunk
                if ch > c0:
                    assert int(hub.chunk_start[ch]) == int(hub.chunk_end[ch - 1])
        # the hub non-zeros are the original ones, row after row
        want = torch.cat([col[int(row_ptr[r]):int(row_ptr[r + 1])] for r in hub.rows.tolist()])
        assert torch.equal(hub.col, want)
    outdeg = torch.bincount(col.long(), minlength=V)
    b = balanced_vertex_bounds(lens, world, vertex_cost=3.0, hub_threshold=threshold, hub_edge_cost=0.65,
                               outdeg_all=outdeg, push_cost=2.0)
    assert b[0] == 0 and b[-1] == V and all(b[i] <= b[i + 1] for i in range(world))
    deg = lens.double()
    per_v = torch.where(lens > threshold, deg * 0.65, deg) + 3.0 * (lens > 0) + 1 + 2.0 * ((lens > 0) & (outdeg > 0))
    total, vmax = float(per_v.sum()), float(per_v.max())
    for r in range(world):
        share = float(per_v[b[r]:b[r + 1]].sum())
        assert share <= total / world + vmax + 1e-9
