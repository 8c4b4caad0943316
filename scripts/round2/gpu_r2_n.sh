#!/bin/bash
# round-2 call N (8 GPUs): PageRank after the fence fix -- hub edge cost re-tuned
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
export DAPB_BENCH_SKIP_PR24=1
run() {  # scale hub_edge_cost vertex_cost tag
  DAPB_PR_HUB_EDGE_COST=$2 DAPB_PR_VERTEX_COST=$3 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29681 \
    -m benchmarks.extras --only pagerank --pagerank-scale $1 --pagerank-iters 30 > gpurun_out/n_pr$1_$4.json 2> gpurun_out/n_pr$1_$4.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/n_pr$1_$4.json").read().strip().splitlines()[-1])
    pr=d.get("pagerank_per_rank") or []
    print("scale $1 hub_cost=$2 vc=$3:", round(d["pagerank_ms_per_iter"],4), "ms", round(d["pagerank_gteps"],1), "GTEPS",
          "| compute us", [round(r["compute_us_mean"]) for r in pr], "| wait us", [round(r["flag_wait_us_mean"]) for r in pr])
except Exception as e: print("parse fail $1 $4", e)
PY
}
run 26 0.65 3 a
run 26 0.8 3 b
run 24 0.65 3 a
run 24 0.8 3 b
run 26 0.65 5 c
