"""Host-side distributed logic on CPU (gloo): results must not depend on the number of processes
(logical n_slices is fixed) and must match the oracle."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from distributed_algorithm_pyspark_b200.config import (ALSConfig, KMeansConfig, LogisticConfig, PageRankConfig)
from distributed_algorithm_pyspark_b200.data.datasets import synthetic_clusters_np
from distributed_algorithm_pyspark_b200.oracle.algorithms import (LogisticInit, als_draw, als_oracle, kmeans_oracle,
                                                                   logistic_oracle, pagerank_oracle,
                                                                   transitive_closure_oracle)

HERE = os.path.dirname(os.path.abspath(__file__))


def _run_worker(tmp_path, nproc, port):
    out = tmp_path / f"w{nproc}.json"
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "_dist_worker.py"), str(out), "torch"]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-3000:]
    return json.loads(out.read_text())


@pytest.fixture(scope="module")
def runs(tmp_path_factory):
    tmp = tmp_path_factory.mktemp("dist")
    return {1: _run_worker(tmp, 1, 29631), 2: _run_worker(tmp, 2, 29632)}


def test_world_size_independence(runs):
    a, b = runs[1], runs[2]
    assert b["world"] == 2
    for algo in ["ssgd", "gd", "ma", "bmuf", "easgd"]:
        np.testing.assert_allclose(a[algo]["w"], b[algo]["w"], rtol=1e-9, atol=1e-12)
        assert a[algo]["bsz"] == b[algo]["bsz"]
    assert a["kmeans_toy"] == b["kmeans_toy"] == [[10.0, 2.0], [1.0, 2.0]]
    np.testing.assert_allclose(a["kmeans_syn"]["centers"], b["kmeans_syn"]["centers"], rtol=1e-12)
    assert a["kmeans_syn"]["assign"] == b["kmeans_syn"]["assign"]
    for k in a["pagerank"]:
        assert abs(a["pagerank"][k] - b["pagerank"][k]) < 1e-15
    assert a["tc"] == b["tc"]
    np.testing.assert_allclose(a["als"]["rmse"], b["als"]["rmse"], rtol=1e-9)


def test_resume_under_torchrun(runs):
    """Checkpoints written and resumed by 1 and by 2 processes reproduce the uninterrupted runs exactly."""
    for r in runs.values():
        assert r["pagerank_resumed"].keys() == r["pagerank"].keys()
        assert max(abs(r["pagerank_resumed"][k] - r["pagerank"][k]) for k in r["pagerank"]) == 0.0
        assert r["easgd_resumed"]["w"] == r["easgd"]["w"] and r["easgd_resumed"]["accs"] == r["easgd"]["accs"]
        assert r["easgd_resumed"]["bsz"] == r["easgd"]["bsz"]
        assert r["kmeans_resumed"] == r["kmeans_syn"]["centers"]


def test_matches_oracle(runs, breast_cancer):
    b = runs[2]
    d = breast_cancer
    for algo in ["ssgd", "gd", "ma", "bmuf", "easgd"]:
        cfg = LogisticConfig.for_algo(algo, n_iterations=12, seed=0)
        ref = logistic_oracle(d.matrix, d.X_test_aug, d.y_test, cfg, LogisticInit.draw(algo, d.dim, cfg.n_slices, 0))
        np.testing.assert_allclose(b[algo]["w"], ref.w, rtol=1e-8, atol=1e-10)
        assert b[algo]["accs"] == pytest.approx(ref.accs)
        assert b[algo]["bsz"] == ref.batch_sizes
    pts, _ = synthetic_clusters_np(500, 5, 4, seed=1)
    km = kmeans_oracle(pts, KMeansConfig(k=4, n_slices=4, n_iterations=4))
    np.testing.assert_allclose(b["kmeans_syn"]["centers"], km.centers, rtol=1e-10)
    assert b["kmeans_syn"]["assign"] == km.assignment.tolist()
    rs = np.random.RandomState(0)
    e = list(zip(rs.randint(0, 60, 150).tolist(), rs.randint(0, 60, 150).tolist()))
    pr = pagerank_oracle(e, PageRankConfig())
    assert set(map(int, b["pagerank"])) == set(pr)
    assert max(abs(b["pagerank"][str(k)] - v) for k, v in pr.items()) < 1e-14
    n, rounds, pairs = transitive_closure_oracle(e)
    assert b["tc"]["n"] == n and b["tc"]["rounds"] == rounds
    assert sorted(map(tuple, b["tc"]["pairs"])) == sorted(pairs)
    cfg = ALSConfig(m=40, n=60, k=5, seed=0, n_iterations=3)
    R, U, V = als_draw(cfg)
    _, _, errs = als_oracle(R, U, V, cfg)
    np.testing.assert_allclose(b["als"]["rmse"], errs, rtol=1e-8)
