#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "pagerank" --timeout 600 > gpurun_out/pr5_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/pr5_tests.log
for sc in 24 26; do
timeout 900 python benchmarks/extras.py --only pagerank --pagerank-scale $sc > gpurun_out/pr5_s$sc.json 2> gpurun_out/pr5_s$sc.err; echo "s$sc rc=$? $(cut -c1-90 gpurun_out/pr5_s$sc.json)"
done
