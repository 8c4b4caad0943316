"""Transitive closure of the reference's four-edge graph (or a random graph with --random V E).

Same entry point, constant and printed line as ``graph_computation/transitive_closure.py``: paths grow
by one edge per round through a join with the edge relation until a round adds nothing.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

n_slices = 2  # Number of Slices


def main(argv=None):
    from distributed_algorithm_pyspark_b200.config import TransitiveClosureConfig
    from distributed_algorithm_pyspark_b200.data.datasets import TC_TOY, random_graph_np
    from distributed_algorithm_pyspark_b200.models.transitive_closure import run_transitive_closure
    from distributed_algorithm_pyspark_b200.utils import cli, report

    ap = cli.base_parser(__doc__.splitlines()[0])
    ap.add_argument("--random", type=int, nargs=2, metavar=("V", "E"), default=None)
    args = ap.parse_args(argv)
    dist = cli.setup(args) if args.backend != "oracle" else None

    if args.random:
        s, d = random_graph_np(args.random[0], args.random[1], seed=args.seed or 0)
        edges = list(zip(s.tolist(), d.tolist()))
    else:
        edges = TC_TOY
    res = run_transitive_closure(edges, TransitiveClosureConfig(n_slices=n_slices), backend=args.backend, dist=dist)
    cli.main_print(dist, "The original graph has %i paths" % res.n_pairs)
    if (dist is None or dist.is_main) and not args.no_save:
        report.save_result("transitive_closure_result", args.out_dir, n_pairs=res.n_pairs, rounds=res.rounds,
                           pairs=res.pairs if res.pairs is not None else [])
    return res


if __name__ == "__main__":
    main()
