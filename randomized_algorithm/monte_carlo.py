"""Monte-Carlo estimate of pi.

Same entry point, constants and printed line as ``randomized_algorithm/monte_carlo.py``: n points uniform
in the square, accepted inside the unit circle, pi ~ 4 * count / n.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

n_slices = 4
# times of sampling
n = 100000 * n_slices


def main(argv=None):
    from distributed_algorithm_pyspark_b200.config import MonteCarloConfig
    from distributed_algorithm_pyspark_b200.models.monte_carlo import run_monte_carlo
    from distributed_algorithm_pyspark_b200.utils import cli, report

    ap = cli.base_parser(__doc__.splitlines()[0])
    ap.add_argument("--n", type=int, default=n)
    args = ap.parse_args(argv)
    dist = cli.setup(args) if args.backend != "oracle" else None

    pi, count = run_monte_carlo(MonteCarloConfig(n_slices=n_slices, n=args.n, seed=args.seed), backend=args.backend, dist=dist)
    cli.main_print(dist, "Pi is roughly %f" % pi)
    if (dist is None or dist.is_main) and not args.no_save:
        report.save_result("monte_carlo_result", args.out_dir, pi=pi, count=count, n=args.n)
    return pi


if __name__ == "__main__":
    main()
