#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "pagerank or state_dict or entry_scripts" --timeout 600 > gpurun_out/pr2_tests.log 2>&1; echo "tests rc=$?"
tail -n 8 gpurun_out/pr2_tests.log
for lay in warp cta; do
DAPB_PR_LAYOUT=$lay timeout 600 python benchmarks/extras.py --only pagerank --pagerank-scale 24 > gpurun_out/pr2_s24_$lay.json 2> gpurun_out/pr2_s24_$lay.err; echo "$lay rc=$?"; cat gpurun_out/pr2_s24_$lay.json
done
