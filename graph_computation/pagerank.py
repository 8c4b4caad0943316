"""PageRank on the reference's four-edge graph (or an R-MAT graph with --rmat SCALE).

Same entry point, constants and printed lines as ``graph_computation/pagerank.py``: duplicate links
dropped, n = number of distinct source vertices, rank = q/n + (1-q) * (sum of contributions), vertices
that receive nothing drop out.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

n_slices = 3  # Number of Slices
n_iterations = 10  # Number of iterations
q = 0.15  # the default value of q is 0.15


def main(argv=None):
    from distributed_algorithm_pyspark_b200.config import PageRankConfig
    from distributed_algorithm_pyspark_b200.data.datasets import PAGERANK_TOY, rmat_edges_np
    from distributed_algorithm_pyspark_b200.models.pagerank import run_pagerank
    from distributed_algorithm_pyspark_b200.utils import cli, report

    ap = cli.base_parser(__doc__.splitlines()[0])
    ap.add_argument("--n-iterations", type=int, default=n_iterations)
    ap.add_argument("--q", type=float, default=q)
    ap.add_argument("--rmat", type=int, default=None, metavar="SCALE", help="use an R-MAT graph of 2^SCALE vertices")
    ap.add_argument("--edge-factor", type=int, default=16)
    args = ap.parse_args(argv)
    dist = cli.setup(args) if args.backend != "oracle" else None

    if args.rmat is not None:
        s, d = rmat_edges_np(args.rmat, args.edge_factor, seed=args.seed or 0)
        links = list(zip(s.tolist(), d.tolist()))
    else:
        links = PAGERANK_TOY
    cfg = PageRankConfig(n_slices=n_slices, n_iterations=args.n_iterations, q=args.q)
    ranks = run_pagerank(links, cfg, backend=args.backend, dist=dist, **cli.checkpoint_kwargs(args))
    if not args.quiet or len(ranks) <= 64:
        for vertex, rank in ranks.items():
            cli.main_print(dist, "%s has rank: %s." % (vertex, rank))
    if (dist is None or dist.is_main) and not args.no_save:
        report.save_result("pagerank_result", args.out_dir, vertices=list(ranks.keys()), ranks=list(ranks.values()))
    return ranks


if __name__ == "__main__":
    main()
