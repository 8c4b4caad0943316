#!/bin/bash
# round-2 call L (1 GPU): ncu --set full of the round-2 kernels at benchmark shape (each after the plain command exited 0)
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
export DAPB_BENCH_SKIP_PR24=1
run() {  # name regex skip cmd...
  local name=$1 regex=$2 skip=$3; shift 3
  "$@" > gpurun_out/ncu_plain_$name.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 -f -o gpurun_out/prof_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
}
run lr_step_r2 lr_step_kernel 4 python bench.py --steps 3 --warmup 3 --no-extras --no-e2e --no-baseline
run kmeans_tc_r2 kmeans_tc_kernel 1 python tests/_tc_worker.py kmeans 16000000 128 1024 0
run pagerank_run_r2 pagerank_run_kernel 1 python -m benchmarks.extras --only pagerank --pagerank-scale 24
run gram_tc_r2 gram_tc_kernel 1 python -m benchmarks.extras --only gram
ls -la gpurun_out/prof_*_r2.ncu-rep
