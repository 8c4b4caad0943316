"""Lloyd's k-means on the reference's six 2-D points (or synthetic data with --synthetic).

Same entry point, constants, printed line and plot file as the reference script
``machine_learning/k-means.py``: centres from takeSample(False, k, 42), ties to the lowest index, empty
clusters keep their centre, fixed iteration count (convergeDist is defined but unused there as well).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

k = 2
convergeDist = 0.1
n_slices = 2
n_iterations = 5


def main(argv=None):
    from distributed_algorithm_pyspark_b200.config import KMeansConfig
    from distributed_algorithm_pyspark_b200.data.datasets import KMEANS_TOY, synthetic_clusters_np
    from distributed_algorithm_pyspark_b200.models.kmeans import run_kmeans
    from distributed_algorithm_pyspark_b200.utils import cli, report

    ap = cli.base_parser(__doc__.splitlines()[0])
    ap.add_argument("--k", type=int, default=k)
    ap.add_argument("--n-iterations", type=int, default=n_iterations)
    ap.add_argument("--n-slices", type=int, default=n_slices)
    ap.add_argument("--synthetic", type=int, nargs=2, metavar=("N", "D"), default=None,
                    help="cluster N synthetic D-dimensional points instead of the toy matrix")
    args = ap.parse_args(argv)
    dist = cli.setup(args) if args.backend != "oracle" else None

    if args.synthetic:
        matrix, _ = synthetic_clusters_np(args.synthetic[0], args.synthetic[1], args.k, seed=args.seed or 0)
    else:
        matrix = KMEANS_TOY
    cfg = KMeansConfig(k=args.k, convergeDist=convergeDist, n_slices=args.n_slices, n_iterations=args.n_iterations)
    res = run_kmeans(matrix, cfg, backend=args.backend, dist=dist, **cli.checkpoint_kwargs(args))
    k_centers = [c for c in res.centers]
    cli.main_print(dist, "Final centers: " + str(k_centers))
    if (dist is None or dist.is_main) and not args.no_save:
        report.save_result("kmeans_result", args.out_dir, centers=res.centers, assignment=res.assignment)
        if matrix.shape[1] == 2:
            report.display_clusters(matrix, res.assignment, cfg.k, "kmeans_clusters_display.png", args.out_dir)
    return res


if __name__ == "__main__":
    main()
