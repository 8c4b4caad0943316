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
    if dist.is_main:
        with open(out_path, "w") as f:
            json.dump(out, f)
    reset_default_heap()
    shutdown()


if __name__ == "__main__":
    main()
