#!/bin/bash
# round-2 call D (2 GPUs): SYRK Gram + cooperative Cholesky + Jacobi eigen-solver; PageRank at 6 CTAs/SM
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/d_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/d_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/d_smoke.log
for n in 1 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29654 \
    -m benchmarks.extras --only gram,pagerank --pagerank-scale 24 --baseline > gpurun_out/d_gram_pr_n$n.json 2> gpurun_out/d_gram_pr_n$n.err
  echo "gram/pr n=$n rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/d_gram_pr_n$n.json").read().strip().splitlines()[-1])
    print({k:v for k,v in d.items() if "per_rank" not in k})
except Exception as e: print("parse fail", e)
PY
done
