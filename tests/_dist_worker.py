"""Worker run under torch.distributed.run by the multi-process tests (gloo on CPU, NCCL on GPUs).
Writes rank 0's results as JSON to argv[1]; argv[2] = backend."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from distributed_algorithm_pyspark_b200.config import (ALSConfig, KMeansConfig, LogisticConfig, MonteCarloConfig,  # noqa: E402
                                                        PageRankConfig)
from distributed_algorithm_pyspark_b200.data.datasets import (KMEANS_TOY, load_breast_cancer_split,  # noqa: E402
                                                               synthetic_clusters_np)
from distributed_algorithm_pyspark_b200.models.kmeans import run_kmeans  # noqa: E402
from distributed_algorithm_pyspark_b200.models.logistic import run_logistic  # noqa: E402
from distributed_algorithm_pyspark_b200.models.matrix_decomposition import run_als  # noqa: E402
from distributed_algorithm_pyspark_b200.models.monte_carlo import run_monte_carlo  # noqa: E402
from distributed_algorithm_pyspark_b200.models.pagerank import run_pagerank  # noqa: E402
from distributed_algorithm_pyspark_b200.models.transitive_closure import run_transitive_closure  # noqa: E402
from distributed_algorithm_pyspark_b200.oracle.algorithms import LogisticInit  # noqa: E402
from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed, shutdown  # noqa: E402


def main():
    out_path, backend = sys.argv[1], sys.argv[2]
    dist = init_distributed()
    out = {"world": dist.world}
    data = load_breast_cancer_split()
    for algo in ["ssgd", "gd", "ma", "bmuf", "easgd"]:
        cfg = LogisticConfig.for_algo(algo, n_iterations=12, seed=0)
        init = LogisticInit.draw(algo, data.dim, cfg.n_slices, 0)
        r = run_logistic(data, cfg, init, backend=backend, dist=dist)
        out[algo] = {"w": r.w.tolist(), "accs": r.accs, "bsz": r.batch_sizes}
    out["kmeans_toy"] = run_kmeans(KMEANS_TOY, KMeansConfig(), backend=backend, dist=dist).centers.tolist()
    pts, _ = synthetic_clusters_np(500, 5, 4, seed=1)
    km = run_kmeans(pts, KMeansConfig(k=4, n_slices=4, n_iterations=4), backend=backend, dist=dist)
    out["kmeans_syn"] = {"centers": km.centers.tolist(), "assign": km.assignment.tolist()}
    rs = np.random.RandomState(0)
    e = list(zip(rs.randint(0, 60, 150).tolist(), rs.randint(0, 60, 150).tolist()))
    pr = run_pagerank(e, PageRankConfig(), backend=backend, dist=dist)
    out["pagerank"] = {str(k): v for k, v in pr.items()}
    tc = run_transitive_closure(e, backend=backend, dist=dist)
    out["tc"] = {"n": tc.n_pairs, "rounds": tc.rounds, "pairs": tc.pairs.tolist()}
    als = run_als(ALSConfig(m=40, n=60, k=5, seed=0, n_iterations=3), backend=backend, dist=dist)
    out["als"] = {"rmse": als.rmse, "U": als.U.tolist()}
    # ---- checkpoint / resume under torchrun: stop after a few iterations, resume to the end, compare with the
    #      uninterrupted results above (PageRank writes one file per rank, the others write on rank 0)
    ck_dir = out_path + ".ckpt"
    os.makedirs(ck_dir, exist_ok=True)
    run_pagerank(e, PageRankConfig(n_iterations=4), backend=backend, dist=dist, checkpoint=ck_dir + "/pr.npz", checkpoint_every=2)
    dist.barrier()
    pr2 = run_pagerank(e, PageRankConfig(), backend=backend, dist=dist, checkpoint=ck_dir + "/pr.npz", checkpoint_every=5,
                       resume=True)
    out["pagerank_resumed"] = {str(k): v for k, v in pr2.items()}
    cfg_full = LogisticConfig.for_algo("easgd", n_iterations=12, seed=0)
    init = LogisticInit.draw("easgd", data.dim, cfg_full.n_slices, 0)
    run_logistic(data, LogisticConfig.for_algo("easgd", n_iterations=5, seed=0), init, backend=backend, dist=dist,
                 checkpoint=ck_dir + "/easgd.npz", checkpoint_every=5)
    dist.barrier()
    r2 = run_logistic(data, cfg_full, None, backend=backend, dist=dist, checkpoint=ck_dir + "/easgd.npz", resume=True)
    out["easgd_resumed"] = {"w": r2.w.tolist(), "accs": r2.accs, "bsz": r2.batch_sizes}
    run_kmeans(pts, KMeansConfig(k=4, n_slices=4, n_iterations=2), backend=backend, dist=dist, checkpoint=ck_dir + "/km.npz",
               checkpoint_every=1)
    dist.barrier()
    km2 = run_kmeans(pts, KMeansConfig(k=4, n_slices=4, n_iterations=4), backend=backend, dist=dist,
                     checkpoint=ck_dir + "/km.npz", resume=True)
    out["kmeans_resumed"] = km2.centers.tolist()
    out["mc"] = run_monte_carlo(MonteCarloConfig(n=20001, seed=11), backend="fused" if backend == "fused" else "torch", dist=dist)[1]
    if dist.is_main:
        with open(out_path, "w") as f:
            json.dump(out, f)
    if backend == "fused":
        from distributed_algorithm_pyspark_b200.parallel.symm import reset_default_heap
        reset_default_heap()
    shutdown()


if __name__ == "__main__":
    main()
