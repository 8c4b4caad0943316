"""Elastic Averaging SGD (synchronous EASGD) for logistic regression on the breast-cancer data.

Same entry point, hyper-parameters, printed lines and plot file as the reference script
``optimization/easgd.py``; the Spark RDD pipeline is replaced by GPU-resident shards and fused
sm_100a kernels (``--backend fused``), with a PyTorch+NCCL baseline (``--backend torch``) and the
NumPy Spark-semantics oracle (``--backend oracle``).  Run under ``torchrun`` for several GPUs.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

n_slices = 4  # Number of Slices
n_iterations = 1500  # Number of iterations
eta = 0.1
mini_batch_fraction = 0.1  # the fraction of mini batch sample
rho = 0.1  # penalty constraint coefficient
alpha = eta * rho  # iterative constraint coefficient
beta = n_slices * alpha  # the parameter of history information


def main(argv=None):
    from distributed_algorithm_pyspark_b200.config import LogisticConfig
    from distributed_algorithm_pyspark_b200.data.datasets import load_breast_cancer_split
    from distributed_algorithm_pyspark_b200.models.logistic import run_logistic
    from distributed_algorithm_pyspark_b200.oracle.algorithms import LogisticInit
    from distributed_algorithm_pyspark_b200.utils import cli, report

    ap = cli.base_parser(__doc__.splitlines()[0])
    ap.add_argument("--n-iterations", type=int, default=n_iterations)
    ap.add_argument("--n-slices", type=int, default=n_slices)
    ap.add_argument("--sampler", default="spark", choices=["spark", "philox"])
    args = ap.parse_args(argv)
    dist = cli.setup(args) if args.backend != "oracle" else None

    data = load_breast_cancer_split()
    cfg = LogisticConfig.for_algo("easgd", n_slices=n_slices, n_iterations=n_iterations, eta=eta, mini_batch_fraction=mini_batch_fraction, rho=rho, seed=args.seed, sampler=args.sampler)
    cfg.n_iterations, cfg.n_slices = args.n_iterations, args.n_slices
    init = LogisticInit.draw("easgd", data.dim, cfg.n_slices, args.seed)
    cli.main_print(dist, "Initial w: " + str(init.w))

    res = run_logistic(data, cfg, init, backend=args.backend, dist=dist, **cli.checkpoint_kwargs(args))
    if not args.quiet:
        for t, acc in enumerate(res.accs):
            cli.main_print(dist, "On iteration %d" % (t + 1))
            cli.main_print(dist, "iterations: %d, accuracy: %f" % (t, acc))

    cli.main_print(dist, "Final w: %s " % res.w)
    cli.main_print(dist, "Final acc: %f" % res.final_acc)
    if (dist is None or dist.is_main) and not args.no_save:
        report.draw_acc_plot(res.accs, len(res.accs), "easgd_acc_plot.png", args.out_dir)
        report.save_result("easgd_result", args.out_dir, w=res.w, accs=res.accs, batch_sizes=res.batch_sizes)
    return res


if __name__ == "__main__":
    main()
