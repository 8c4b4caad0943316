"""Printing / plotting / saving helpers shared by the entry scripts.

The reference plots with matplotlib (``optimization/ssgd.py:50-66``, ``machine_learning/k-means.py:30-40``)
and only prints its results.  matplotlib is optional here (it is absent on the build image): the raw
curves / assignments are always written next to where the PNG would go, and final results are saved as
``.npy`` / ``.json``.
"""
from __future__ import annotations

import json
import os
import random
from typing import Dict, Sequence

import numpy as np

from ..oracle.algorithms import ewma_smooth


def _pyplot():
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        return plt
    except Exception:
        return None


def draw_acc_plot(accs: Sequence[float], n_iterations: int, png_name: str, out_dir: str = ".") -> Dict[str, str]:
    """ref ssgd.py:50-66: raw accuracy (alpha 0.3) + EWMA(0.9) curve, title / axis labels, savefig."""
    accs = np.asarray(list(accs), dtype=np.float64)
    s_accs = ewma_smooth(accs, alpha=0.9)
    stem = os.path.splitext(png_name)[0]
    written = {}
    os.makedirs(out_dir, exist_ok=True)
    raw_path = os.path.join(out_dir, stem + ".npy")
    np.save(raw_path, np.stack([accs, s_accs]))
    written["curve"] = raw_path
    plt = _pyplot()
    if plt is not None:
        plt.figure()
        x = np.arange(1, n_iterations + 1)
        plt.plot(x, accs, color="C0", alpha=0.3)
        plt.plot(x, s_accs, color="C0")
        plt.title(label="Accuracy on test dataset")
        plt.xlabel("Round")
        plt.ylabel("Accuracy")
        path = os.path.join(out_dir, png_name)
        plt.savefig(path)
        plt.close()
        written["png"] = path
    return written


def display_clusters(points: np.ndarray, assignment: np.ndarray, k: int,
                     png_name: str = "kmeans_clusters_display.png", out_dir: str = ".") -> Dict[str, str]:
    """ref k-means.py:30-40: one scatter per cluster with a random hex colour (only called when d == 2)."""
    written = {}
    os.makedirs(out_dir, exist_ok=True)
    raw = os.path.join(out_dir, os.path.splitext(png_name)[0] + ".npy")
    np.save(raw, np.concatenate([np.asarray(points, dtype=np.float64), np.asarray(assignment)[:, None]], axis=1))
    written["assignment"] = raw
    plt = _pyplot()
    if plt is not None:
        plt.figure()
        for c_id in range(k):
            pts = np.asarray(points)[np.asarray(assignment) == c_id]
            if len(pts) == 0:
                continue
            color = "#" + "".join(random.choice("0123456789ABCDEF") for _ in range(6))
            plt.scatter(pts[:, 0], pts[:, 1], c=color)
        path = os.path.join(out_dir, png_name)
        plt.savefig(path)
        plt.close()
        written["png"] = path
    return written


def save_result(name: str, out_dir: str = ".", **arrays) -> str:
    """Save final outputs (weights, centroids, ranks, factors, counts) as ``<name>.npz`` + ``<name>.json`` summary."""
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, name + ".npz")
    np.savez(path, **{k: np.asarray(v) for k, v in arrays.items()})
    summary = {k: (np.asarray(v).tolist() if np.asarray(v).size <= 64 else list(np.asarray(v).shape)) for k, v in arrays.items()}
    with open(os.path.join(out_dir, name + ".json"), "w") as f:
        json.dump(summary, f)
    return path
