"""Checkpoint / resume (utils/checkpoint.py): a resumed run must follow the uninterrupted trajectory exactly.
The reference has no checkpointing; this is an addition for long GPU runs."""
import os
import subprocess
import sys

import numpy as np
import pytest

from distributed_algorithm_pyspark_b200.config import ALSConfig, KMeansConfig, LogisticConfig, PageRankConfig
from distributed_algorithm_pyspark_b200.utils import checkpoint as ck

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_atomic_roundtrip_and_validation(tmp_path):
    p = str(tmp_path / "a" / "c.npz")
    ck.save_checkpoint(p, {"w": np.arange(5.0), "none": None, "t": np.int64(7)}, {"kind": "x", "iteration": 7})
    st, meta = ck.load_checkpoint(p)
    assert set(st) == {"w", "t"} and meta["iteration"] == 7 and int(st["t"]) == 7
    np.testing.assert_array_equal(st["w"], np.arange(5.0))
    assert [f for f in os.listdir(tmp_path / "a") if f.endswith(".tmp")] == []      # no temp file left behind
    with pytest.raises(ValueError):
        ck.check_meta(meta, kind="y")
    np.savez(str(tmp_path / "foreign.npz"), w=np.zeros(2))
    with pytest.raises(ValueError):
        ck.load_checkpoint(str(tmp_path / "foreign.npz"))


@pytest.mark.parametrize("backend", ["oracle", "torch"])
@pytest.mark.parametrize("algo", ["ssgd", "gd", "ma", "bmuf", "easgd"])
def test_logistic_resume_is_bit_identical(tmp_path, algo, backend):
    from distributed_algorithm_pyspark_b200.data.datasets import load_breast_cancer_split
    from distributed_algorithm_pyspark_b200.models.logistic import run_logistic
    from distributed_algorithm_pyspark_b200.oracle.algorithms import LogisticInit
    data = load_breast_cancer_split()
    cfg = LogisticConfig.for_algo(algo, n_iterations=17, seed=0)
    init = LogisticInit.draw(algo, data.dim, cfg.n_slices, 0)
    full = run_logistic(data, cfg, init, backend=backend)
    p = str(tmp_path / "lr.npz")
    run_logistic(data, LogisticConfig.for_algo(algo, n_iterations=8, seed=0), init, backend=backend, checkpoint=p,
                 checkpoint_every=3)                                            # "killed" after 8 iterations
    assert ck.load_checkpoint(p)[1]["iteration"] == 8
    res = run_logistic(data, cfg, None, backend=backend, checkpoint=p, checkpoint_every=5, resume=True)
    np.testing.assert_array_equal(res.w, full.w)
    assert res.accs == full.accs and res.batch_sizes == full.batch_sizes
    with pytest.raises(ValueError):                                             # a different algorithm must not resume it
        other = "ma" if algo != "ma" else "ssgd"
        run_logistic(data, LogisticConfig.for_algo(other, n_iterations=9, seed=0), None, backend=backend, checkpoint=p,
                     resume=True)


def test_kmeans_pagerank_als_resume(tmp_path):
    from distributed_algorithm_pyspark_b200.data.datasets import synthetic_clusters_np
    from distributed_algorithm_pyspark_b200.models.kmeans import run_kmeans
    from distributed_algorithm_pyspark_b200.models.matrix_decomposition import run_als
    from distributed_algorithm_pyspark_b200.models.pagerank import run_pagerank
    pts, _ = synthetic_clusters_np(400, 4, 5, seed=2)
    p = str(tmp_path / "km.npz")
    full = run_kmeans(pts, KMeansConfig(k=5, n_iterations=7), backend="torch")
    run_kmeans(pts, KMeansConfig(k=5, n_iterations=3), backend="torch", checkpoint=p, checkpoint_every=1)
    res = run_kmeans(pts, KMeansConfig(k=5, n_iterations=7), backend="torch", checkpoint=p, checkpoint_every=2, resume=True)
    np.testing.assert_array_equal(res.centers, full.centers)
    assert len(res.history) == 7

    rs = np.random.RandomState(0)
    e = list(zip(rs.randint(0, 50, 200).tolist(), rs.randint(0, 50, 200).tolist()))
    p = str(tmp_path / "pr.npz")
    full = run_pagerank(e, PageRankConfig(n_iterations=9), backend="torch")
    run_pagerank(e, PageRankConfig(n_iterations=4), backend="torch", checkpoint=p, checkpoint_every=2)
    res = run_pagerank(e, PageRankConfig(n_iterations=9), backend="torch", checkpoint=p, checkpoint_every=3, resume=True)
    assert res.keys() == full.keys() and max(abs(res[k] - full[k]) for k in full) < 1e-15

    cfg = ALSConfig(m=30, n=40, k=4, seed=0, n_iterations=6)
    p = str(tmp_path / "als.npz")
    full = run_als(cfg, backend="torch")
    run_als(ALSConfig(m=30, n=40, k=4, seed=0, n_iterations=2), backend="torch", checkpoint=p, checkpoint_every=1)
    res = run_als(cfg, backend="torch", checkpoint=p, checkpoint_every=2, resume=True)
    np.testing.assert_array_equal(res.V, full.V)
    np.testing.assert_array_equal(res.U, full.U)
    assert res.rmse == full.rmse


def test_script_flags_resume(tmp_path):
    """`optimization/ssgd.py --checkpoint F --checkpoint-every K`, then `--resume` to more iterations."""
    p = str(tmp_path / "s.npz")
    base = [sys.executable, os.path.join(ROOT, "optimization", "ssgd.py"), "--backend", "oracle", "--seed", "0", "--quiet",
            "--no-save", "--checkpoint", p]
    a = subprocess.run(base + ["--n-iterations", "6", "--checkpoint-every", "2"], capture_output=True, text=True, timeout=300)
    assert a.returncode == 0, a.stderr
    b = subprocess.run(base + ["--n-iterations", "12", "--resume"], capture_output=True, text=True, timeout=300)
    assert b.returncode == 0, b.stderr
    c = subprocess.run(base[:-2] + ["--n-iterations", "12"], capture_output=True, text=True, timeout=300)
    final = lambda out: [l for l in out.splitlines() if l.startswith("Final")]
    assert final(b.stdout) == final(c.stdout) and len(final(b.stdout)) == 2
    bad = subprocess.run(base[:-2] + ["--resume"], capture_output=True, text=True, timeout=300)
    assert bad.returncode != 0
