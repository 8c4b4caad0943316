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


def bench_kmeans(dist, backend, n=125_000_000, d=128, k=1024, iters=20, warmup=1, baseline=False, checks=None):
    """BASELINE.json config 3: 1B x 128 points, k = 1024 over 8 GPUs -> 125M points per GPU (bf16 storage), 20 iterations."""
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
           "kmeans_iterations": iters, "kmeans_launches_per_iter": getattr(eng, "launches_per_iter", None),
           "kmeans_tflops_per_gpu": flops / 1e12, "kmeans_frac_of_measured_bf16": flops / 1e12 / pk["bf16_tflops_sustained"],
           "kmeans_tensor_core_path": bool(getattr(eng, "tc", None) is not None)}
    if checks is not None:
        checks["kmeans_counts_sum_to_n"] = bool(int(eng.counts_out.long().sum().item()) == n * dist.world)
    if baseline and backend == "fused":
        del eng
        torch.cuda.empty_cache()
        base = ShardedKMeans(X, k, c0, dist=dist, backend="nccl")
        bms = _timed(dist, base.step, 2, 1)
        out["kmeans_baseline_ms_per_iter"] = bms
        out["kmeans_vs_baseline"] = bms / ms
        # lower bound for ANY cuBLAS-based formulation: just the two GEMMs an iteration needs (distances, one-hot sums),
        # bf16 in / bf16 out, no elementwise work, no arg-min, no collective
        CmT = c0.to(torch.bfloat16).T.contiguous()
        chunk = 1 << 20
        oh = torch.zeros(chunk, k, dtype=torch.bfloat16, device=dev)

        def gemms_only():
            for i in range(0, n, chunk):
                x = X[i:i + chunk]
                torch.mm(x, CmT)
                torch.mm(oh[:x.shape[0]].T, x)
        gms = _timed(dist, gemms_only, 1, 1)
        out["kmeans_cublas_gemms_only_ms_per_iter"] = gms
        out["kmeans_vs_cublas_gemms_only"] = gms / ms
        del oh, CmT
        eng = base
    del X, eng
    torch.cuda.empty_cache()
    return out


def bench_kmeans_fp32(dist, n=125_000_000, d=128, k=1024, iters=5, warmup=1):
    """Same shape with fp32 storage (64 GB per GPU): distances on the tensor cores from a bf16 shadow copy (+32 GB), sums
    from the fp32 rows.  Roofline: max(tensor 2 n k d / sustained bf16, HBM (64 + 32 GB) / measured copy bandwidth)."""
    import torch
    from distributed_algorithm_pyspark_b200.data.datasets import synthetic_clusters_device
    from distributed_algorithm_pyspark_b200.models.kmeans import ShardedKMeans
    dev = dist.device
    free, _ = torch.cuda.mem_get_info(dev)
    while n * d * 6 > 0.7 * free:
        n //= 2
    X, centres = synthetic_clusters_device(n, d, k, dev, torch.float32, seed=7 + dist.rank)
    c0 = (centres + 0.5 * torch.randn_like(centres)).float()
    if dist.world > 1:
        dist.broadcast(c0, 0)
    eng = ShardedKMeans(X, k, c0, dist=dist, backend="fused")
    ms = _timed(dist, eng.step, iters, warmup)
    pk = _peaks()
    t_tensor = 2.0 * n * k * d / (pk["bf16_tflops_sustained"] * 1e12) * 1e3
    t_hbm = n * d * 6 / (pk["hbm_gbs"] * 1e9) * 1e3
    out = {"kmeans_fp32_ms_per_iter": ms, "kmeans_fp32_n_per_gpu": n, "kmeans_fp32_pts_per_s_iter": n * dist.world / (ms / 1e3),
           "kmeans_fp32_tensor_core_path": bool(eng.tc is not None and eng.tc.X32 is not None),
           "kmeans_fp32_roofline_ms": max(t_tensor, t_hbm), "kmeans_fp32_frac_of_roofline": max(t_tensor, t_hbm) / ms}
    del X, eng
    torch.cuda.empty_cache()
    return out


def bench_pagerank(dist, backend, scale=26, edge_factor=16, iters=30, warmup=1, stats=True, tag="pagerank",
                   baseline=False, checks=None):
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
        pr._collect_stats = False
    if baseline and fused:
        # same graph, same partition through the PyTorch + NCCL arm (index_add_ SpMV + all_gather); also a result check:
        # after `iters` iterations from the same start both arms must agree
        pr.reset()
        pr.run(iters)
        ours = pr.ranks_local.double().clone()
        base = ShardedPageRank(g, PageRankConfig(n_iterations=iters), dist, backend="nccl")
        bms = _timed(dist, lambda: base.run(3), 1, 0) / 3
        base.run(iters - 3)
        out[f"{tag}_baseline_ms_per_iter"] = bms
        out[f"{tag}_vs_baseline"] = bms / ms
        ref = base.ranks_local.double()
        rel = float(dist.max_over_ranks(float(((ours - ref).abs() / ref.clamp(min=1e-30))[ref > 0].max().item()) if ref.numel() else 0.0))
        same = bool(torch.equal(ours > 0, ref > 0))
        if checks is not None:
            checks[f"{tag}_max_rel_diff_vs_nccl_arm"] = rel
            checks[f"{tag}_agrees_with_nccl_arm"] = bool(rel < 1e-3 and all(dist.all_gather_object(same)))
        del base
    del pr, g
    torch.cuda.empty_cache()
    return out


def check_pagerank_vs_single_gpu(dist, checks, scale=18):
    """N-GPU fused ranks against a 1-GPU recomputation (fp64 torch on this rank alone) of the same scale-18 R-MAT graph."""
    import torch
    from distributed_algorithm_pyspark_b200.config import PageRankConfig
    from distributed_algorithm_pyspark_b200.data.datasets import rmat_edges_device
    from distributed_algorithm_pyspark_b200.models.pagerank import ShardedPageRank, build_sharded_graph
    from distributed_algorithm_pyspark_b200.parallel.dist import Dist
    dev = dist.device
    src, dst = rmat_edges_device(scale, 16, dev, seed=2)
    cfg = PageRankConfig(n_iterations=10)
    g = build_sharded_graph(src, dst, 1 << scale, dist)
    pr = ShardedPageRank(g, cfg, dist, backend="fused")
    pr.run(10)
    got = pr.ranks().double()
    solo = Dist(rank=0, world=1, local_rank=dist.local_rank, device=dev)
    g1 = build_sharded_graph(src, dst, 1 << scale, solo, layout="cta")
    ref = ShardedPageRank(g1, cfg, solo, backend="nccl", dtype=torch.float64)
    ref.run(10)
    want = ref.ranks_local
    rel = float(((got - want).abs() / want.clamp(min=1e-30))[want > 0].max().item())
    checks["pagerank18_ngpu_vs_1gpu_max_rel_diff"] = dist.max_over_ranks(rel)
    checks["pagerank18_ngpu_vs_1gpu_ok"] = bool(dist.max_over_ranks(rel) < 1e-5 * 10 and torch.equal(got > 0, want > 0))
    del pr, ref, g, g1
    torch.cuda.empty_cache()


def bench_gram(dist, backend, n=1 << 24, p=512, iters=3, warmup=1, baseline=False, checks=None):
    """BASELINE.json config 5a: 2^27 x 512 over 8 GPUs -> 2^24 rows per GPU (bf16), Gram + cross-GPU reduce + the small
    on-device factorisations (Cholesky R factor, symmetric eigen-decomposition)."""
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
    jit = 1e-3 * float(G.diagonal().mean().item())
    t_chol = _timed(dist, lambda: dec.cholesky(G, jitter=jit), 2, 1)
    flops = 2.0 * n * p * p / (ms / 1e3)
    pk = _peaks()
    out = {"gram_ms": ms, "gram_rows_per_gpu": n, "gram_tflops_per_gpu": flops / 1e12,
           "gram_frac_of_measured_bf16": flops / 1e12 / pk["bf16_tflops_sustained"],
           "gram_hbm_gbs": n * p * 2 / (ms / 1e3) / 1e9, "cholesky_ms": t_chol,
           "gram_tensor_core_path": bool(dec.tc is not None)}
    if hasattr(dec, "eigh_native") and backend == "fused":
        try:
            out["eigh_ms"] = _timed(dist, lambda: dec.eigh(G), 1, 1)
            w, _ = dec.eigh(G)
            wr = torch.linalg.eigvalsh(G.double())
            if checks is not None:
                checks["gram_eigenvalues_max_rel_err"] = float(((w.double().sort().values - wr).abs() / wr.abs().max()).max().item())
        except Exception as e:
            out["eigh_error"] = repr(e)[:200]
    if baseline and backend == "fused":
        base = GramDecomposition(A, dist, backend="nccl")
        bms = _timed(dist, base.gram, 2, 1)
        Gb = base.gram()
        out["gram_baseline_ms"] = bms
        out["gram_vs_baseline"] = bms / ms
        if checks is not None:
            checks["gram_max_rel_diff_vs_cublas_nccl"] = float((G - Gb).abs().max().item() / Gb.abs().max().item())
        del base, Gb
    del A, dec
    torch.cuda.empty_cache()
    return out


def bench_monte_carlo(dist, backend, n_total=None, baseline=False, checks=None):
    """BASELINE.json config 5b: 10^12 samples over 8 GPUs -> 1.25e11 per GPU; sampler, reduction and the cross-GPU total
    are one launch per rank."""
    import math
    import torch
    n = 125_000_000_000 if backend == "fused" else 2_000_000_000
    if backend == "fused":
        from distributed_algorithm_pyspark_b200.models.monte_carlo import mc_count_fused
        first = dist.rank * n
        mc_count_fused(first, 1 << 20, 12345, dist)                 # warm-up: module load, heap region
        res = {}

        def run():
            res["t"] = mc_count_fused(first, n, 12345, dist, return_tensor=True)
        ms = _timed(dist, run, 1, 0)
        pi = 4.0 * float(res["t"].item()) / (n * dist.world)
    else:
        from distributed_algorithm_pyspark_b200.config import MonteCarloConfig
        from distributed_algorithm_pyspark_b200.models.monte_carlo import run_monte_carlo
        t0 = time.perf_counter()
        pi, _ = run_monte_carlo(MonteCarloConfig(n=n * dist.world, seed=1), backend="torch", dist=dist)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
    out = {"mc_samples_per_s": n * dist.world / (ms / 1e3), "mc_ms": ms, "mc_samples": n * dist.world, "mc_pi": pi}
    if checks is not None:
        sigma = 4.0 * math.sqrt((math.pi / 4) * (1 - math.pi / 4) / (n * dist.world))
        checks["mc_pi_within_4_sigma"] = bool(abs(pi - math.pi) < 4 * sigma)
    if baseline and backend == "fused":
        from distributed_algorithm_pyspark_b200.config import MonteCarloConfig
        from distributed_algorithm_pyspark_b200.models.monte_carlo import run_monte_carlo
        nb = 2_000_000_000
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        run_monte_carlo(MonteCarloConfig(n=nb * dist.world, seed=1), backend="torch", dist=dist)
        torch.cuda.synchronize()
        bms = dist.max_over_ranks((time.perf_counter() - t0) * 1e3)
        out["mc_baseline_samples_per_s"] = nb * dist.world / (bms / 1e3)
        out["mc_vs_baseline"] = out["mc_samples_per_s"] / out["mc_baseline_samples_per_s"]
    return out


def _tc_case(dist, be, V, E, seed, reps=3):
    import torch
    from distributed_algorithm_pyspark_b200.models.transitive_closure import closure_device
    dev = dist.device
    g = torch.Generator(device=dev).manual_seed(seed)
    src = torch.randint(0, V, (E,), device=dev, generator=g)
    dst = torch.randint(0, V, (E,), device=dev, generator=g)
    closure_device(src, dst, V, be, dist)          # warm-up at full size (allocator pools, heap regions, module load)
    best = None
    for _ in range(reps):
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        total, rounds, _ = closure_device(src, dst, V, be, dist)   # includes de-duplication and the CSR build
        torch.cuda.synchronize()
        ms = dist.max_over_ranks((time.perf_counter() - t0) * 1e3)
        best = ms if best is None else min(best, ms)
    return total, rounds, best


def bench_tc(dist, backend, V=2_000_000, E=1_000_000, baseline=False, checks=None):
    """BASELINE.json config 4b: transitive closure of a 1M-edge random graph.  Two regimes: V = 2M (sub-critical: 2 M pairs,
    latency-bound, 20 rounds) and V = 2^18 with the same 1M edges... (mean degree 4 is super-critical and would need V^2
    pairs), so the throughput case is V = 1.25 M, E = 1 M: mean degree 0.8, ~1e8-pair closure."""
    be = "fused" if backend == "fused" else "torch"
    total, rounds, best = _tc_case(dist, be, V, E, 3)
    out = {"tc_pairs": total, "tc_rounds": rounds, "tc_ms": best, "tc_pairs_per_s": total / (best / 1e3)}
    try:
        Vb = 1_020_000
        tb, rb, bb = _tc_case(dist, be, Vb, E, 3, reps=2)
        out.update({"tc_big_V": Vb, "tc_big_pairs": tb, "tc_big_rounds": rb, "tc_big_ms": bb, "tc_big_pairs_per_s": tb / (bb / 1e3)})
        if baseline and backend == "fused" and tb < 3e8:
            t2, r2, b2 = _tc_case(dist, "torch", Vb, E, 3, reps=1)
            out["tc_big_baseline_ms"] = b2
            out["tc_big_vs_baseline"] = b2 / bb
            if checks is not None:
                checks["tc_big_pairs_agree_with_torch_arm"] = bool(t2 == tb and r2 == rb)
    except Exception as e:
        out["tc_big_error"] = repr(e)[:300]
    return out


ALL = {"kmeans": bench_kmeans, "pagerank": bench_pagerank, "gram": bench_gram, "mc": bench_monte_carlo, "tc": bench_tc}


def run_all(dist, backend, only=None, baseline=False, checks=None, **kw):
    out = {}
    for name, fn in ALL.items():
        if only and name not in only:
            continue
        try:
            out.update(fn(dist, backend, baseline=baseline, checks=checks, **kw.get(name, {})))
        except Exception as e:                          # one failing extra must not hide the others
            out[name + "_error"] = repr(e)[:300]
    if (not only or "kmeans" in only) and backend == "fused":
        try:    # fp32 points (SURVEY 7.4 sized k-means for fp32 storage): tensor-core path via a resident bf16 shadow copy
            out.update(bench_kmeans_fp32(dist))
        except Exception as e:
            out["kmeans_fp32_error"] = repr(e)[:300]
    if (not only or "pagerank" in only) and backend == "fused" and not os.environ.get("DAPB_BENCH_SKIP_PR24"):
        try:    # continuity with round 1 (strong scaling of the smaller graph) + the N-GPU vs 1-GPU result check
            out.update(bench_pagerank(dist, backend, scale=24, iters=30, stats=False, tag="pagerank24"))
            if checks is not None:
                check_pagerank_vs_single_gpu(dist, checks)
        except Exception as e:
            out["pagerank24_error"] = repr(e)[:300]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--impl", default="ours", choices=["ours", "nccl"])
    ap.add_argument("--pagerank-scale", type=int, default=26)
    ap.add_argument("--pagerank-iters", type=int, default=30)
    ap.add_argument("--baseline", action="store_true", help="also run the PyTorch+NCCL arm of each config")
    a = ap.parse_args()
    from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed, shutdown
    dist = init_distributed()
    only = [s for s in a.only.split(",") if s] or None
    checks = {}
    out = run_all(dist, "fused" if a.impl == "ours" else "nccl", only, baseline=a.baseline, checks=checks,
                  pagerank={"scale": a.pagerank_scale, "iters": a.pagerank_iters})
    out["impl"], out["n_gpus"], out["checks"] = a.impl, dist.world, checks
    if dist.is_main:
        print(json.dumps(out))
    if a.impl == "ours":
        from distributed_algorithm_pyspark_b200.parallel.symm import reset_default_heap
        reset_default_heap()
    shutdown()


if __name__ == "__main__":
    main()
