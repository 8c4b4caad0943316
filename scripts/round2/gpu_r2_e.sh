#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_tc_kernels.py -x -q --timeout 600 -k "eigh or cholesky or gram or kmeans or pagerank" > gpurun_out/e_pytest.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/e_pytest.log
for g in 0 16 48; do DAPB_CHOL_GRID=$g timeout 120 python scripts/chol_eigh_probe.py 2>/dev/null | tail -1; done | tee gpurun_out/e_chol_eigh.jsonl
timeout 600 python -m benchmarks.extras --only gram --baseline 2>gpurun_out/e_gram.err | tail -1 | tee gpurun_out/e_gram_n1.json
