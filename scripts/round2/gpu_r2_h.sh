#!/bin/bash
# round-2 call H (2 GPUs): PageRank tail experiments (grab size, block size) on small per-rank work + fp32 k-means path
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
export DAPB_BENCH_SKIP_PR24=1
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q --timeout 600 -k "kmeans or pagerank" > gpurun_out/h_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/h_pytest.log
run() {  # scale grab block
  DAPB_PR_GRAB=$2 DAPB_PR_BLOCK_NNZ=$3 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29661 \
    -m benchmarks.extras --only pagerank --pagerank-scale $1 --pagerank-iters 30 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); pr=d.get('pagerank_per_rank') or []
print('scale $1 grab $2 block $3:', round(d['pagerank_ms_per_iter'],4), 'ms', round(d['pagerank_gteps'],1), 'GTEPS | compute', [round(r['compute_us_mean']) for r in pr], '| wait', [round(r['flag_wait_us_mean']) for r in pr])"
}
run 22 4 512
run 22 1 512
run 22 1 256
run 22 2 256
run 24 4 512
run 24 1 512
run 24 1 256
