#!/bin/bash
# First contact: smoke, GPU tests, headline bench (ours vs NCCL baseline).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
python -c "import pyspark" > gpurun_out/pyspark_probe.txt 2>&1; which java >> gpurun_out/pyspark_probe.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary.txt
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/summary.txt
tail -30 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/bench_ours.json
timeout 600 python bench.py --steps 5 --warmup 3 --no-extras --no-e2e --impl nccl > gpurun_out/bench_nccl.json 2> gpurun_out/bench_nccl.err; echo "bench nccl rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/bench_nccl.json
tail -n 5 gpurun_out/smoke.log gpurun_out/bench_ours.err gpurun_out/bench_nccl.err
