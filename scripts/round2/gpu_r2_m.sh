#!/bin/bash
# round-2 call M (2 GPUs): PageRank after the fence fix (tests, 1- and 2-GPU timing), k-means ncu capture
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
export DAPB_BENCH_SKIP_PR24=1
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_multigpu.py -x -q --timeout 800 -k "pagerank or multi_gpu or resume or entry" > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/m_pytest.log
for n in 1 2; do for sc in 24 26; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29690 \
    -m benchmarks.extras --only pagerank --pagerank-scale $sc --pagerank-iters 30 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); pr=d.get('pagerank_per_rank') or []
print('n=$n scale $sc:', round(d['pagerank_ms_per_iter'],4), 'ms', round(d['pagerank_gteps'],1), 'GTEPS | compute', [round(r['compute_us_mean']) for r in pr], '| wait', [round(r['flag_wait_us_mean']) for r in pr])"
done; done | tee gpurun_out/m_pr.txt
python tests/_tc_worker.py kmeans 16000000 128 1024 0 > gpurun_out/ncu_plain_kmeans_tc_r2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:kmeans_tc_kernel -s 0 -c 1 -f -o gpurun_out/prof_kmeans_tc_r2 python tests/_tc_worker.py kmeans 16000000 128 1024 0 > gpurun_out/ncu_kmeans_tc_r2.log 2>&1; echo "ncu kmeans rc=$?"
