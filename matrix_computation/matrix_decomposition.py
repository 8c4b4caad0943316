"""Low-rank matrix decomposition by alternating least squares.

Same entry point, constants and printed lines as ``matrix_computation/matrix_decomposition.py``
(R = rand(m,k) rand(n,k)^T, ridge lam * rows, rmse / (m n) per iteration); the factors are also saved.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

lam = 0.01   # regularization coefficient
m = 100  # number of users
n = 500  # number of items
k = 10  # dim of the latent vectors of users and items
n_iterations = 5  # number of iterations
n_slices = 4  # number of slices


def main(argv=None):
    from distributed_algorithm_pyspark_b200.config import ALSConfig
    from distributed_algorithm_pyspark_b200.models.matrix_decomposition import run_als
    from distributed_algorithm_pyspark_b200.utils import cli, report

    ap = cli.base_parser(__doc__.splitlines()[0])
    ap.add_argument("--m", type=int, default=m)
    ap.add_argument("--n", type=int, default=n)
    ap.add_argument("--k", type=int, default=k)
    ap.add_argument("--n-iterations", type=int, default=n_iterations)
    args = ap.parse_args(argv)
    dist = cli.setup(args) if args.backend != "oracle" else None

    cfg = ALSConfig(lam=lam, m=args.m, n=args.n, k=args.k, n_iterations=args.n_iterations, n_slices=n_slices, seed=args.seed)
    res = run_als(cfg, backend=args.backend, dist=dist, **cli.checkpoint_kwargs(args),
                  callback=lambda t, error: cli.main_print(dist, "iterations: %d, rmse: %f" % (t, error)))
    if (dist is None or dist.is_main) and not args.no_save:
        report.save_result("matrix_decomposition_result", args.out_dir, U=res.U, V=res.V, rmse=res.rmse)
    return res


if __name__ == "__main__":
    main()
