#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "pagerank" --timeout 600 > gpurun_out/pr4_tests.log 2>&1; echo "tests rc=$?"
tail -n 4 gpurun_out/pr4_tests.log
for thr in 32 64 256 1024; do
DAPB_PR_HUB_THRESHOLD=$thr timeout 600 python benchmarks/extras.py --only pagerank --pagerank-scale 24 > gpurun_out/pr4_s24_$thr.json 2> gpurun_out/pr4_s24_$thr.err; echo "s24 thr=$thr rc=$? $(cut -c1-90 gpurun_out/pr4_s24_$thr.json)"
done
for thr in 64 256; do
DAPB_PR_HUB_THRESHOLD=$thr timeout 900 python benchmarks/extras.py --only pagerank --pagerank-scale 26 > gpurun_out/pr4_s26_$thr.json 2> gpurun_out/pr4_s26_$thr.err; echo "s26 thr=$thr rc=$? $(cut -c1-90 gpurun_out/pr4_s26_$thr.json)"
done
