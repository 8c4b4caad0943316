#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py tests/test_gpu_kernels.py -m gpu -q --timeout 300 -k "gram or cholesky or als" > gpurun_out/pytest_gram.log 2>&1; echo "pytest rc=$?"
tail -n 8 gpurun_out/pytest_gram.log
timeout 600 python -m benchmarks.extras --only gram > gpurun_out/extras_gram.json 2> gpurun_out/extras_gram.err; echo "gram bench rc=$?"
cat gpurun_out/extras_gram.json; tail -n 3 gpurun_out/extras_gram.err
timeout 600 python -m benchmarks.extras --only gram --impl nccl > gpurun_out/extras_nccl_gram.json 2> gpurun_out/extras_nccl_gram.err; echo "gram nccl rc=$?"
cat gpurun_out/extras_nccl_gram.json
