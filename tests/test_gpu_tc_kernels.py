"""tcgen05 / TMEM / TMA kernels vs plain PyTorch fp32 math.  Each case runs in its own process with a timeout:
a mis-programmed barrier traps after a bounded wait instead of hanging the box, and a faulted context cannot
poison the other tests."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def run_case(*args, timeout=180):
    res = subprocess.run([sys.executable, os.path.join(HERE, "_tc_worker.py"), *map(str, args)], capture_output=True,
                         text=True, timeout=timeout)
    assert res.returncode == 0, (res.stdout[-1500:], res.stderr[-3000:])
    line = [l for l in res.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


@pytest.mark.parametrize("n,d,k,grid", [
    (128, 128, 128, 1),           # one tile, one chunk
    (512, 128, 256, 1),           # one super tile, two chunks
    (128 * 4 * 7 + 77, 128, 256, 2),     # several super tiles per CTA (slot / stage / TMEM ring wrap-around) + ragged tail
    (128 * 37 + 5, 64, 384, 3),          # d = 64, k not a power of two
    (200000, 128, 1024, 0),              # benchmark-shaped k, full grid
    (3000, 128, 100, 2),                 # k below one chunk: 28 padding centroids that must never win
    (5000, 64, 1000, 0),                 # k not a multiple of 128
    (20000, 128, 2048, 0),               # 16 chunks
])
def test_kmeans_tc_assignment_and_sums(n, d, k, grid):
    r = run_case("kmeans", n, d, k, grid)
    assert r["bad_range"] == 0
    assert r["wrong"] == 0, r                 # every disagreement with fp32 argmin must be a genuine near-tie
    assert r["mismatch_vs_fp32"] <= max(2, n // 500), r
    assert r["sum_rel_err"] < 1e-4 and r["counts_ok"], r


def test_kmeans_tc_exact_ties_lowest_index_wins():
    r = run_case("kmeans_ties", 20000, 128, 512)
    assert r["to_duplicates"] == 0 and r["to_originals"] == r["n"], r


@pytest.mark.parametrize("n,P,splits", [
    (64, 128, 1),              # one stage, one strip
    (64 * 7 + 13, 256, 2),     # ring wrap-around + ragged last tile, two strips
    (5000, 512, 3),            # full-width strip (all 512 TMEM columns), uneven splits
    (300000, 512, 0),          # benchmark-shaped width, full grid
])
def test_gram_tc_vs_fp64(n, P, splits):
    r = run_case("gram", n, P, splits)
    tol = 2e-5 if n < 100000 else 1e-4        # fp32 accumulation over n rows
    assert r["err1"] < tol and r["err2"] < tol, r
    assert r["sym"] < tol, r


@pytest.mark.parametrize("n,frac_pct,grid", [
    (300, 100, 1),           # 3 tiles (one ragged) in one CTA, every row
    (5000, 30, 2),           # several tiles per CTA, Bernoulli gather, software pipeline F/B interleave
    (200000, 10, 0),         # benchmark-like sampling, full grid
    (1000, 1, 4),            # most CTAs see fewer than one tile of selected rows (some none at all)
])
def test_lr_tc_step_vs_fp32_and_cuda_core_kernel(n, frac_pct, grid):
    r = run_case("lr", n, frac_pct, grid)
    assert r["cnt_ok_0"] and r["cnt_ok_2"], r
    assert r["grad_err_0"] < 2e-3 and r["grad_err_2"] < 2e-3, r        # same tolerance as the CUDA-core kernel test
    assert r["step_err"] < 2e-3, r
