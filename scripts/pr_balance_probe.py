"""Compute-side load balance of the 8-way PageRank partition, measured on ONE GPU: shard r of 8 is built with the real
partitioner and its local step is timed with a world-1 communicator (pushes stay local), for several cost models."""
import os, sys, types
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_algorithm_pyspark_b200.config import PageRankConfig
from distributed_algorithm_pyspark_b200.data.datasets import rmat_edges_device
from distributed_algorithm_pyspark_b200.models.pagerank import ShardedPageRank, build_sharded_graph
from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 26
W = int(sys.argv[2]) if len(sys.argv) > 2 else 8
settings = [s.split(":") for s in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["8:1.0"])]
dist = init_distributed()
src, dst = rmat_edges_device(scale, 16, dist.device, seed=1)
for vc, hc in settings:
    os.environ["DAPB_PR_VERTEX_COST"], os.environ["DAPB_PR_HUB_EDGE_COST"] = vc, hc
    times = []
    for r in range(W):
        fake = types.SimpleNamespace(world=W, rank=r, device=dist.device)
        g = build_sharded_graph(src, dst, 1 << scale, fake)
        pr = ShardedPageRank(g, PageRankConfig(), dist, backend="fused")
        for _ in range(3):
            pr.step()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            pr.step()
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b) / 10)
        nv, nnz, hub = g.v1 - g.v0, int(g.col_idx.numel()), int(g.hub.col.numel())
        print(f"  vc={vc} hc={hc} rank {r}: {times[-1]:.3f} ms  vertices {nv:>9} main nnz {nnz:>10} hub nnz {hub:>10}", flush=True)
        del pr, g
        torch.cuda.empty_cache()
    print(f"vertex_cost={vc} hub_edge_cost={hc}: max {max(times):.3f} ms, mean {sum(times)/len(times):.3f} ms, "
          f"imbalance {max(times)/(sum(times)/len(times)):.2f}", flush=True)
