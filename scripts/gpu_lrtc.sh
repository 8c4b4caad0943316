#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py -m gpu -q --timeout 300 -k lr_tc > gpurun_out/pytest_lrtc.log 2>&1; echo "lr_tc pytest rc=$?"
tail -n 30 gpurun_out/pytest_lrtc.log | cut -c1-400
bash scripts/gpu_pr.sh
