"""Multi-GPU worker: fused engines vs the NCCL baseline at world > 1 (run under torch.distributed.run)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from distributed_algorithm_pyspark_b200 import _native  # noqa: E402
from distributed_algorithm_pyspark_b200.config import LogisticConfig, PageRankConfig  # noqa: E402
from distributed_algorithm_pyspark_b200.data.datasets import rmat_edges_np  # noqa: E402
from distributed_algorithm_pyspark_b200.models.kmeans import ShardedKMeans  # noqa: E402
from distributed_algorithm_pyspark_b200.models.logistic import ShardedLogisticEngine  # noqa: E402
from distributed_algorithm_pyspark_b200.models.matrix_decomposition import GramOp  # noqa: E402
from distributed_algorithm_pyspark_b200.models.pagerank import ShardedPageRank, build_sharded_graph  # noqa: E402
from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed, shutdown  # noqa: E402
from distributed_algorithm_pyspark_b200.parallel.symm import default_heap, reset_default_heap  # noqa: E402


def main():
    out_path = sys.argv[1]
    dist = init_distributed()
    dev = dist.device
    heap = default_heap(256 << 20)
    lib = _native.load()
    out = {"world": dist.world}
    # ---- stand-alone all-reduce, many epochs
    reg = heap.region("ar_test", 2 * 8 * 4096 * 4, 8)
    ok = True
    for it in range(50):
        src = torch.full((4000,), float(dist.rank + 1 + it), device=dev)
        dst = torch.zeros_like(src)
        _native.check(lib.dapb_allreduce_f32(heap.comm_ref(), reg.payload_off, reg.flag_base, reg.next_epoch(),
                                             src.data_ptr(), dst.data_ptr(), 4000, _native.stream_ptr()))
        want = sum(r + 1 + it for r in range(dist.world))
        ok = ok and bool((dst == want).all().item())
    out["allreduce_ok"] = ok
    out["heap_backend"], out["multicast"] = heap.backend, bool(heap.multicast)
    # ---- flag-protocol stress on real peers: 100k epochs in one persistent launch per rank, random stalls, bitwise check
    fail = torch.zeros(1, dtype=torch.int32, device=dev)
    iters = 100_000
    _native.check(lib.dapb_allreduce_stress(heap.comm_ref(), reg.payload_off, reg.flag_base, reg.epoch + 1, 1024, iters,
                                            17, fail.data_ptr(), _native.stream_ptr()))
    reg.epoch += iters
    torch.cuda.synchronize()
    out["stress_fail"] = int(fail.item())
    # ---- grid-wide collectives (multimem.* when the heap is multicast-mapped) against torch.distributed
    import torch.distributed as tdist
    W = dist.world
    big = heap.region("coll_test", 8 << 20, 8)
    coll_ok = {}
    for rep in range(3):
        n = 4096 * W * (rep + 1)
        x = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(100 * rep + dist.rank))
        want = x.clone()
        tdist.all_reduce(want)
        got = torch.empty_like(x)
        _native.check(lib.dapb_allreduce_twoshot_f32(heap.comm_ref(), big.payload_off, big.flag_base, big.next_epoch(),
                                                     x.data_ptr(), got.data_ptr(), n, _native.stream_ptr()))
        gs = dist.all_gather(got)
        coll_ok[f"twoshot{rep}"] = bool(torch.allclose(got, want, rtol=1e-5, atol=1e-5)) and all(torch.equal(gs[0], t) for t in gs)
        rs_got = torch.empty(n // W, device=dev)
        _native.check(lib.dapb_reduce_scatter_f32(heap.comm_ref(), big.payload_off, big.flag_base, big.next_epoch(),
                                                  x.data_ptr(), rs_got.data_ptr(), n, _native.stream_ptr()))
        coll_ok[f"rs{rep}"] = bool(torch.allclose(rs_got, want[dist.rank * (n // W):(dist.rank + 1) * (n // W)], rtol=1e-5, atol=1e-5))
        for dt, eb in ((torch.float32, 4), (torch.float64, 8)):
            m = 1000 * (rep + 1) + 3
            xs = torch.randn(m, device=dev, dtype=dt) + dist.rank
            ag = torch.empty(m * W, device=dev, dtype=dt)
            _native.check(lib.dapb_allgather(heap.comm_ref(), big.payload_off, big.flag_base, big.next_epoch(),
                                             xs.data_ptr(), ag.data_ptr(), m, eb, _native.stream_ptr()))
            coll_ok[f"ag{rep}_{eb}"] = bool(torch.equal(ag, torch.cat(dist.all_gather(xs))))
        m = 500 * (rep + 1)
        snd = torch.arange(W * m, device=dev, dtype=torch.int64) + 1_000_000 * dist.rank
        rcv = torch.empty_like(snd)
        _native.check(lib.dapb_alltoall_u64(heap.comm_ref(), big.payload_off, big.flag_base, big.next_epoch(),
                                            snd.data_ptr(), rcv.data_ptr(), m, _native.stream_ptr()))
        want_rcv = torch.cat([torch.arange(dist.rank * m, (dist.rank + 1) * m, device=dev, dtype=torch.int64) + 1_000_000 * r
                              for r in range(W)])
        coll_ok[f"a2a{rep}"] = bool(torch.equal(rcv, want_rcv))
    torch.cuda.synchronize()
    out["collectives"] = coll_ok
    # ---- logistic engines: fused vs nccl, identical Philox masks
    torch.manual_seed(100 + dist.rank)
    n, D = 20000, 256
    X = (torch.randn(n, D, device=dev) * 0.3).to(torch.bfloat16).contiguous()
    X[:, -1] = 1
    wstar = torch.randn(D, device=dev, generator=torch.Generator(device=dev).manual_seed(5)) * 0.2
    y = (torch.rand(n, device=dev) < torch.sigmoid(X.float() @ wstar)).float()
    out["lr"] = {}
    for algo in ["ssgd", "gd", "ma", "bmuf", "easgd"]:
        cfg = LogisticConfig.for_algo(algo, eta=0.05 if algo != "gd" else 1e-5, mini_batch_fraction=0.1, seed=3,
                                      n_slices=dist.world)
        a = ShardedLogisticEngine(X, y, cfg, row_base=dist.rank * n, dist=dist, backend="fused")
        b = ShardedLogisticEngine(X, y, cfg, row_base=dist.rank * n, dist=dist, backend="nccl")
        for _ in range(6):
            a.step()
            b.step()
        torch.cuda.synchronize()
        diff = float((a.w - b.w).abs().max().item() / (b.w.abs().max().item() + 1e-12))
        ws = dist.all_gather(a.w)
        same = all(bool(torch.equal(ws[0], t)) for t in ws)
        out["lr"][algo] = {"rel_diff": diff, "bit_identical_across_ranks": same}
    # ---- k-means engine
    Xk = torch.randn(8000, 16, device=dev) + (dist.rank * 3)
    c0 = torch.randn(24, 16, device=torch.device("cpu"), generator=torch.Generator().manual_seed(1)).to(dev) * 2
    ka = ShardedKMeans(Xk, 24, c0, dist=dist, backend="fused")
    kb = ShardedKMeans(Xk, 24, c0, dist=dist, backend="nccl")
    for _ in range(4):
        ka.step()
        kb.step()
    torch.cuda.synchronize()
    out["kmeans_rel_diff"] = float((ka.centers - kb.centers).abs().max().item())
    out["kmeans_counts_equal"] = bool(torch.equal(ka.counts_out, kb.counts_out))
    # counts beyond 2^24 per cluster must survive the cross-rank exchange exactly (they travel as integer bit patterns):
    # feed the finalize kernel hand-made partial sums / counts
    kk, dd = 24, 16
    big = torch.full((kk,), 20_000_001 + dist.rank, dtype=torch.int32, device=dev)
    psum = torch.ones(kk, dd, device=dev) * big[:, None].float()
    cen = torch.zeros(kk, dd, device=dev)
    cout = torch.zeros(kk, dtype=torch.int32, device=dev)
    _native.check(lib.dapb_kmeans_finalize(heap.comm_ref(), psum.data_ptr(), big.data_ptr(), cen.data_ptr(), kk, dd, 0,
                                           ka.region.payload_off, ka.region.flag_base, ka.region.next_epoch(),
                                           cout.data_ptr(), 0, _native.stream_ptr()))
    torch.cuda.synchronize()
    want = sum(20_000_001 + r for r in range(dist.world))
    out["kmeans_big_counts_exact"] = bool((cout == want).all().item())
    # ---- PageRank engine
    s, d = rmat_edges_np(12, 16, seed=3)
    src, dst = torch.as_tensor(s, device=dev), torch.as_tensor(d, device=dev)
    g = build_sharded_graph(src, dst, 1 << 12, dist, block_nnz=512)
    cfg = PageRankConfig(n_iterations=10)
    pa = ShardedPageRank(g, cfg, dist, backend="fused", heap=heap)
    pb = ShardedPageRank(g, cfg, dist, backend="nccl")
    for _ in range(10):
        pa.step()
        pb.step()
    ra, rb = pa.ranks(), pb.ranks()
    out["pagerank_rel_diff"] = float(((ra - rb).abs() / (rb.abs() + 1e-12)).max().item())
    # persistent launch (all 10 iterations in one kernel), hub rows, fp32 and fp64; must equal the stepped run bit for bit
    g2 = build_sharded_graph(src, dst, 1 << 12, dist, hub_threshold=32, hub_chunk=128)
    out["pagerank_run"] = {}
    for dt in (torch.float32, torch.float64):
        p1 = ShardedPageRank(g2, cfg, dist, backend="fused", heap=heap, dtype=dt, collect_stats=True)
        p1.run(10)
        r1 = p1.ranks().clone()
        p2 = ShardedPageRank(g2, cfg, dist, backend="fused", heap=heap, dtype=dt)
        for _ in range(10):
            p2.step()
        r2 = p2.ranks()
        p3 = ShardedPageRank(g2, cfg, dist, backend="nccl", dtype=torch.float64)
        p3.run(10)
        r3 = p3.ranks()
        out["pagerank_run"][str(dt)] = {
            "run_equals_steps": bool(torch.equal(r1, r2)),
            "rel_diff_vs_fp64": float(((r1.double() - r3).abs() / r3.clamp(min=1e-300))[r3 > 0].max().item()),
            "same_support": bool(torch.equal(r1 > 0, r3 > 0)),
            "push_bytes_per_iter": p1.push_bytes_per_iter()}
    # ---- transitive closure of a supercritical graph: buffers must grow mid-run on every rank (persistent kernel:
    #      status GROW decided from the in-kernel message matrix, no host collective), three drivers must agree
    from distributed_algorithm_pyspark_b200.models.transitive_closure import closure_device
    rs = np.random.RandomState(11)
    ts, td = torch.as_tensor(rs.randint(0, 700, 1500), device=dev), torch.as_tensor(rs.randint(0, 700, 1500), device=dev)
    out["tc"] = [list(closure_device(ts, td, 700, "fused", dist, persistent=True)[:2]),
                 list(closure_device(ts, td, 700, "fused", dist, persistent=False)[:2]),
                 list(closure_device(ts, td, 700, "torch", dist)[:2])]
    # ---- Gram with fused all-reduce
    A = torch.randn(5000, 96, device=dev)
    op = GramOp(96, 96, torch.float32, dev, dist, heap=heap, tag="mg")
    G = op(A, A).clone()
    Gr = A.T @ A
    dist.all_reduce_sum(Gr)
    out["gram_rel_diff"] = float((G - Gr).abs().max().item() / Gr.abs().max().item())
    # ---- tensor-core Gram (SYRK roles) with the grid-wide in-kernel two-shot all-reduce, twice (epoch / buffer reuse)
    from distributed_algorithm_pyspark_b200.ops.gram_tc import GramTC
    out["gram_tc_rel_diff"] = []
    for Pn in (256, 512):
        At = (torch.randn(6000 + 100 * dist.rank, Pn, device=dev)).to(torch.bfloat16).contiguous()
        tc = GramTC(At, dist, heap=heap)
        ref_g = At.float().T @ At.float()
        dist.all_reduce_sum(ref_g)
        for _ in range(2):
            Gt = tc().clone()
            gs = dist.all_gather(Gt)
            out["gram_tc_rel_diff"].append(float((Gt - ref_g).abs().max().item() / ref_g.abs().max().item()))
            out.setdefault("gram_tc_identical_across_ranks", True)
            out["gram_tc_identical_across_ranks"] = out["gram_tc_identical_across_ranks"] and all(bool(torch.equal(gs[0], t)) for t in gs)
    if dist.is_main:
        with open(out_path, "w") as f:
            json.dump(out, f)
    reset_default_heap()
    shutdown()


if __name__ == "__main__":
    main()
