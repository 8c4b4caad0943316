#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q --timeout 300 -k "pagerank or entry" > gpurun_out/pytest_pr.log 2>&1; echo "pytest rc=$?"
tail -n 6 gpurun_out/pytest_pr.log
timeout 600 python -m benchmarks.extras --only pagerank > gpurun_out/extras_pagerank.json 2> gpurun_out/extras_pagerank.err; echo "pagerank rc=$?"
cat gpurun_out/extras_pagerank.json; tail -n 3 gpurun_out/extras_pagerank.err
timeout 900 python -m benchmarks.extras --only pagerank --pagerank-scale 26 > gpurun_out/extras_pagerank26_n1.json 2> gpurun_out/extras_pagerank26_n1.err; echo "pagerank26 rc=$?"
cat gpurun_out/extras_pagerank26_n1.json; tail -n 3 gpurun_out/extras_pagerank26_n1.err
