#!/bin/bash
# round-2 call G (8 GPUs): multi-GPU tests at 4 ranks, PageRank partition sweep at scale 24 / 26
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -x -q --timeout 800 > gpurun_out/g_multigpu.log 2>&1; echo "multigpu rc=$?"; tail -3 gpurun_out/g_multigpu.log
export DAPB_BENCH_SKIP_PR24=1
run() {  # scale vertex_cost push_cost tag
  DAPB_PR_VERTEX_COST=$2 DAPB_PR_PUSH_COST=$3 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29660 \
    -m benchmarks.extras --only pagerank --pagerank-scale $1 --pagerank-iters 30 > gpurun_out/g_pr$1_$4.json 2> gpurun_out/g_pr$1_$4.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/g_pr$1_$4.json").read().strip().splitlines()[-1])
    pr=d.get("pagerank_per_rank") or []
    print("scale $1 vc=$2 pc=$3:", round(d["pagerank_ms_per_iter"],4), "ms", round(d["pagerank_gteps"],1), "GTEPS", "| s24:", d.get("pagerank24_ms_per_iter"),
          "| compute us", [round(r["compute_us_mean"]) for r in pr], "| wait us", [round(r["flag_wait_us_mean"]) for r in pr])
except Exception as e: print("parse fail $1 $4", e)
PY
}
run 24 8 2 a
run 24 3 2 b
run 24 1.5 1 c
run 24 1 0.5 d
run 26 3 2 b
run 26 1.5 1 c
