#!/bin/bash
mkdir -p gpurun_out
DAPB_PR_HUB_MODE=window DAPB_PR_HUB_CHUNK=64 timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "pagerank" --timeout 600 > gpurun_out/pr4_tests.log 2>&1; echo "tests(window) rc=$?"
tail -n 4 gpurun_out/pr4_tests.log
for mode in chunk window; do for sc in 24 26; do
DAPB_PR_HUB_MODE=$mode DAPB_PR_HUB_CHUNK=$([ $mode = window ] && echo 1024 || echo 2048) timeout 900 python benchmarks/extras.py --only pagerank --pagerank-scale $sc > gpurun_out/pr4_${mode}_s$sc.json 2> gpurun_out/pr4_${mode}_s$sc.err; echo "$mode s$sc rc=$? $(cut -c1-90 gpurun_out/pr4_${mode}_s$sc.json)"
done; done
DAPB_PR_HUB_MODE=window DAPB_PR_HUB_CHUNK=512 timeout 900 python benchmarks/extras.py --only pagerank --pagerank-scale 26 > gpurun_out/pr4_window512_s26.json 2> gpurun_out/pr4_window512_s26.err; echo "window512 s26 rc=$? $(cut -c1-90 gpurun_out/pr4_window512_s26.json)"
