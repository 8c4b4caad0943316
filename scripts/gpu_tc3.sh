#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py -m gpu -q --timeout 300 > gpurun_out/pytest_tc.log 2>&1; echo "tc pytest rc=$?"
tail -n 15 gpurun_out/pytest_tc.log
timeout 300 python scripts/kmeans_tc_probe.py 32000000
