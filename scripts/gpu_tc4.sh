#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py -m gpu -q --timeout 300 -k gram > gpurun_out/pytest_tc_gram.log 2>&1; echo "gram pytest rc=$?"
tail -n 25 gpurun_out/pytest_tc_gram.log
timeout 600 python -m benchmarks.extras --only gram > gpurun_out/extras_gram.json 2> gpurun_out/extras_gram.err; echo "gram bench rc=$?"
cat gpurun_out/extras_gram.json; tail -n 3 gpurun_out/extras_gram.err
