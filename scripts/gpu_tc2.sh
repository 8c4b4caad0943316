#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py -m gpu -q --timeout 300 > gpurun_out/pytest_tc.log 2>&1; echo "tc pytest rc=$?"
tail -n 15 gpurun_out/pytest_tc.log
timeout 600 python -m benchmarks.extras --only kmeans > gpurun_out/extras_kmeans.json 2> gpurun_out/extras_kmeans.err; echo "kmeans rc=$?"
cat gpurun_out/extras_kmeans.json; tail -n 3 gpurun_out/extras_kmeans.err
