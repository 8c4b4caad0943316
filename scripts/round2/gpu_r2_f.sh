#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py tests/test_gpu_kernels.py -x -q --timeout 600 -k "kmeans or eigh or cholesky or gram" > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/f_pytest.log
for b in 8 4; do DAPB_EIGH_B=$b timeout 120 python scripts/chol_eigh_probe.py 2>/dev/null | tail -1; done | tee gpurun_out/f_chol_eigh.jsonl
for pe in 0 32 128; do
  DAPB_GRAM_PACE_EVERY=$pe DAPB_GRAM_PACE_WINDOW=$((pe*2)) timeout 300 python -m benchmarks.extras --only gram 2>gpurun_out/f_gram.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('pace', $pe, d.get('gram_ms'), d.get('gram_hbm_gbs'))"
done | tee gpurun_out/f_gram_pace.txt
timeout 600 python -m benchmarks.extras --only kmeans 2>gpurun_out/f_kmeans.err | tail -1 | tee gpurun_out/f_kmeans_n1.json
DAPB_GRAM_PACE_EVERY=0 timeout 300 python -m benchmarks.extras --only gram > gpurun_out/f_plain.log 2>&1 && \
DAPB_GRAM_PACE_EVERY=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:gram_tc_kernel -s 1 -c 1 -o gpurun_out/f_gram_prof python -m benchmarks.extras --only gram > gpurun_out/f_ncu.log 2>&1; echo "ncu rc=$?"
