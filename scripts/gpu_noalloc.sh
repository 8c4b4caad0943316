#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"
tail -n 5 gpurun_out/noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "all rc=$?"
tail -n 5 gpurun_out/pytest_gpu.log
