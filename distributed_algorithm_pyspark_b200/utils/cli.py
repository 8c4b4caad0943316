"""Shared command-line plumbing of the ten entry scripts.

The reference configures itself with module-level constants only; the scripts
keep those names and defaults and add optional flags on top.
"""
from __future__ import annotations

import argparse
import sys


def base_parser(description: str) -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description=description)
    ap.add_argument("--backend", default="auto", choices=["auto", "fused", "torch", "oracle"],
                    help="fused = sm_100a kernels, torch = PyTorch+NCCL/gloo baseline, oracle = NumPy Spark-semantics restatement")
    ap.add_argument("--seed", type=int, default=None, help="seed for everything the reference leaves unseeded")
    ap.add_argument("--out-dir", default=".", help="where plots / saved results go")
    ap.add_argument("--device", default=None, choices=[None, "cpu", "cuda"])
    ap.add_argument("--quiet", action="store_true", help="suppress per-iteration prints")
    ap.add_argument("--no-save", action="store_true", help="do not write result files")
    ap.add_argument("--checkpoint", default=None, metavar="FILE.npz",
                    help="iterative drivers: file that receives the optimiser state (utils/checkpoint.py)")
    ap.add_argument("--checkpoint-every", type=int, default=0, metavar="K", help="write --checkpoint every K iterations")
    ap.add_argument("--resume", action="store_true", help="continue from --checkpoint if it exists")
    return ap


def checkpoint_kwargs(args) -> dict:
    """The three checkpoint flags as keyword arguments of the ``run_*`` drivers."""
    if not args.checkpoint:
        if args.resume or args.checkpoint_every:
            raise SystemExit("--resume / --checkpoint-every need --checkpoint FILE")
        return {}
    return dict(checkpoint=args.checkpoint, checkpoint_every=args.checkpoint_every, resume=args.resume)


def setup(args):
    """Join the process group (if launched by torchrun) and return the Dist handle."""
    from ..parallel.dist import init_distributed
    return init_distributed(args.device)


def main_print(dist, *a, **kw):
    """Print on rank 0 only (the reference's driver prints)."""
    if dist is None or dist.is_main:
        print(*a, **kw)
        sys.stdout.flush()
