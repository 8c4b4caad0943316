"""Secondary benchmark configurations (BASELINE.json configs 3-5), reported under "extra" by bench.py and
runnable on their own:  python -m benchmarks.extras [--only kmeans,pagerank,...] [--impl ours|nccl]

Every number is CUDA-event time on the launching stream after warm-up, max over ranks; inputs are far larger
than L2.  Shapes are the per-GPU shard of the 8-GPU configuration (weak scaling) unless stated.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "fallback": True}


def _timed(dist, fn, iters, warmup):
    import torch
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    dist.barrier()
    return dist.max_over_ranks(a.elapsed_time(b)) / iters


def bench_kmeans(dist, backend, n=125_000_000, d=128, k=1024, iters=3, warmup=1):
    """BASELINE.json config 3: 1B x 128 points, k = 1024 over 8 GPUs -> 125M points per GPU (bf16 storage)."""
    import torch
    from distributed_algorithm_pyspark_b200.data.datasets import synthetic_clusters_device
    from distributed_algorithm_pyspark_b200.models.kmeans import ShardedKMeans
    dev = dist.device
    free, _ = torch.cuda.mem_get_info(dev)
    while n * d * 2 > 0.5 * free:
        n //= 2
    X, centres = synthetic_clusters_device(n, d, k, dev, torch.bfloat16, seed=7 + dist.rank)
    c0 = (centres + 0.5 * torch.randn_like(centres)).float()
    if dist.world > 1:
        dist.broadcast(c0, 0)
    eng = ShardedKMeans(X, k, c0, dist=dist, backend="fused" if backend == "fused" else "nccl")
    ms = _timed(dist, eng.step, iters, warmup)
    pts = n * dist.world / (ms / 1e3)
    flops = 2.0 * n * k * d / (ms / 1e3)
    pk = _peaks()
    out = {"kmeans_pts_per_s_iter": pts, "kmeans_ms_per_iter": ms, "kmeans_n_per_gpu": n, "kmeans_k": k, "kmeans_d": d,
           "kmeans_tflops_per_gpu": flops / 1e12, "kmeans_frac_of_measured_bf16": flops / 1e12 / pk["bf16_tflops_sustained"],
           "kmeans_tensor_core_path": bool(getattr(eng, "tc", None) is not None)}
    del X, eng
    torch.cuda.empty_cache()
    return out


def bench_pagerank(dist, backend, scale=26, edge_factor=16, iters=30, warmup=1, stats=True, tag="pagerank"):
    """BASELINE.json config 4: R-MAT scale-26 (edge factor 16), 30 iterations; GTEPS = de-duplicated edges * iterations /
    time.  The fused backend runs the 30 iterations as ONE persistent launch (``ShardedPageRank.run``); a second,
    untimed run collects the in-kernel %globaltimer stamps for the per-rank breakdown."""
    import torch
    from distributed_algorithm_pyspark_b200.config import PageRankConfig
    from distributed_algorithm_pyspark_b200.data.datasets import rmat_edges_device
    from distributed_algorithm_pyspark_b200.models.pagerank import ShardedPageRank, build_sharded_graph
    dev = dist.device
    fused = backend == "fused"
    copies = None
    if fused and dist.world > 1:
        from distributed_algorithm_pyspark_b200.parallel.symm import default_heap
        heap = default_heap(max(1 << 30, 2 * (1 << scale) * 4 + (64 << 20)))
        copies = 1 if heap.multicast else dist.world - 1
    src, dst = rmat_edges_device(scale, edge_factor, dev, seed=1)
    g = build_sharded_graph(src, dst, 1 << scale, dist, push_copies=copies)
    del src, dst
    torch.cuda.empty_cache()
    pr = ShardedPageRank(g, PageRankConfig(n_iterations=iters), dist, backend="fused" if fused else "nccl")
    ms = _timed(dist, lambda: pr.run(iters), 1, warmup) / iters
    pr.finish()
    out = {f"{tag}_gteps": g.n_edges / (ms / 1e3) / 1e9, f"{tag}_ms_per_iter": ms, f"{tag}_scale": scale,
           f"{tag}_iterations": iters, f"{tag}_edges_dedup": g.n_edges}
    if fused:
        out[f"{tag}_launches_per_run"] = 1
        out[f"{tag}_multicast"] = bool(dist.world > 1 and pr.heap.multicast)
    if fused and stats:
        pr._collect_stats = True
        pr.reset()
        pr.run(iters)
        pr.finish()
        st = pr.stats.double().cpu()                       # [iters, 4] ns: enter, wait passed, last ticket, published
        mine = {"rank": dist.rank, "local_vertices": int(g.v1 - g.v0), "local_nnz": int(g.col_idx.numel() + g.hub.col.numel()),
                "push_bytes_per_iter": pr.push_bytes_per_iter(),
                "flag_wait_us_mean": float((st[1:, 1] - st[1:, 0]).mean() / 1e3),
                "compute_us_mean": float((st[1:, 2] - st[1:, 1]).mean() / 1e3),
                "publish_us_mean": float((st[1:, 3] - st[1:, 2]).mean() / 1e3),
                "iter_us_mean": float((st[1:, 0] - st[:-1, 0]).mean() / 1e3)}
        out[f"{tag}_per_rank"] = dist.all_gather_object(mine)
    del pr, g
    torch.cuda.empty_cache()
    return out


def bench_gram(dist, backend, n=1 << 24, p=512, iters=3, warmup=1):
    """BASELINE.json config 5a: 2^27 x 512 over 8 GPUs -> 2^24 rows per GPU (bf16), Gram + cross-GPU reduce + Cholesky."""
    import torch
    from distributed_algorithm_pyspark_b200.models.matrix_decomposition import GramDecomposition
    dev = dist.device
    free, _ = torch.cuda.mem_get_info(dev)
    while n * p * 2 > 0.5 * free:
        n //= 2
    A = torch.empty(n, p, dtype=torch.bfloat16, device=dev)
    for r0 in range(0, n, 1 << 22):
        A[r0:r0 + (1 << 22)].normal_()
    dec = GramDecomposition(A, dist, backend="fused" if backend == "fused" else "nccl")
    if backend == "fused" and dec.tc is None:          # CUDA-core fallback is only meant for small shapes
        n_small = 1 << 18
        dec = GramDecomposition(A[:n_small], dist, backend="fused")
        n = n_small
    ms = _timed(dist, dec.gram, iters, warmup)
    G = dec.gram()
    t_chol = _timed(dist, lambda: dec.cholesky(G, jitter=1e-3 * float(G.diagonal().mean().item())), 1, 1)
    flops = 2.0 * n * p * p / (ms / 1e3)
    pk = _peaks()
    out = {"gram_ms": ms, "gram_rows_per_gpu": n, "gram_tflops_per_gpu": flops / 1e12,
           "gram_frac_of_measured_bf16": flops / 1e12 / pk["bf16_tflops_sustained"],
           "gram_hbm_gbs": n * p * 2 / (ms / 1e3) / 1e9, "cholesky_ms": t_chol,
           "gram_tensor_core_path": bool(dec.tc is not None)}
    del A, dec
    torch.cuda.empty_cache()
    return out


def bench_monte_carlo(dist, backend, n_total=None):
    """BASELINE.json config 5b: 10^12 samples over 8 GPUs -> 1.25e11 per GPU."""
    import torch
    from distributed_algorithm_pyspark_b200 import _native
    n = 125_000_000_000 if backend == "fused" else 2_000_000_000
    dev = dist.device
    if backend == "fused":
        lib = _native.load()
        cnt = torch.zeros(1, dtype=torch.int64, device=dev)
        first = dist.rank * n

        def run():
            _native.check(lib.dapb_mc_pi(first, n, 12345, cnt.data_ptr(), _native.stream_ptr()))
        ms = _timed(dist, run, 1, 0)
        dist.all_reduce_sum(cnt)
        pi = 4.0 * float(cnt.item()) / (n * dist.world)
    else:
        from distributed_algorithm_pyspark_b200.config import MonteCarloConfig
        from distributed_algorithm_pyspark_b200.models.monte_carlo import run_monte_carlo
        t0 = time.perf_counter()
        pi, _ = run_monte_carlo(MonteCarloConfig(n=n * dist.world, seed=1), backend="torch", dist=dist)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
    return {"mc_samples_per_s": n * dist.world / (ms / 1e3), "mc_ms": ms, "mc_samples": n * dist.world, "mc_pi": pi}


def bench_tc(dist, backend, V=2_000_000, E=1_000_000):
    """BASELINE.json config 4b: transitive closure of a 1M-edge random graph (V = 2M, sub-critical)."""
    import torch
    from distributed_algorithm_pyspark_b200.models.transitive_closure import closure_device
    dev = dist.device
    g = torch.Generator(device=dev).manual_seed(3)
    src = torch.randint(0, V, (E,), device=dev, generator=g)
    dst = torch.randint(0, V, (E,), device=dev, generator=g)
    be = "fused" if backend == "fused" else "torch"
    closure_device(src, dst, V, be, dist)          # warm-up at full size (allocator pools, heap regions, module load)
    best = None
    for _ in range(3):
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        total, rounds, _ = closure_device(src, dst, V, be, dist)   # includes de-duplication and the CSR build
        torch.cuda.synchronize()
        ms = dist.max_over_ranks((time.perf_counter() - t0) * 1e3)
        best = ms if best is None else min(best, ms)
    return {"tc_pairs": total, "tc_rounds": rounds, "tc_ms": best, "tc_pairs_per_s": total / (best / 1e3)}


ALL = {"kmeans": bench_kmeans, "pagerank": bench_pagerank, "gram": bench_gram, "mc": bench_monte_carlo, "tc": bench_tc}


def run_all(dist, backend, only=None, **kw):
    out = {}
    for name, fn in ALL.items():
        if only and name not in only:
            continue
        try:
            out.update(fn(dist, backend, **kw.get(name, {})))
        except Exception as e:                          # one failing extra must not hide the others
            out[name + "_error"] = repr(e)[:300]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--impl", default="ours", choices=["ours", "nccl"])
    ap.add_argument("--pagerank-scale", type=int, default=26)
    ap.add_argument("--pagerank-iters", type=int, default=30)
    a = ap.parse_args()
    from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed, shutdown
    dist = init_distributed()
    only = [s for s in a.only.split(",") if s] or None
    out = run_all(dist, "fused" if a.impl == "ours" else "nccl", only, pagerank={"scale": a.pagerank_scale, "iters": a.pagerank_iters})
    out["impl"], out["n_gpus"] = a.impl, dist.world
    if dist.is_main:
        print(json.dumps(out))
    if a.impl == "ours":
        from distributed_algorithm_pyspark_b200.parallel.symm import reset_default_heap
        reset_default_heap()
    shutdown()


if __name__ == "__main__":
    main()
