#!/bin/bash
mkdir -p gpurun_out
for what in kmeans pagerank mc tc gram; do
  timeout 600 python -m benchmarks.extras --only $what > gpurun_out/extras_$what.json 2> gpurun_out/extras_$what.err; echo "$what rc=$?"
  cat gpurun_out/extras_$what.json; tail -n 3 gpurun_out/extras_$what.err
done
for what in kmeans pagerank; do
  timeout 900 python -m benchmarks.extras --only $what --impl nccl > gpurun_out/extras_nccl_$what.json 2> gpurun_out/extras_nccl_$what.err; echo "nccl $what rc=$?"
  cat gpurun_out/extras_nccl_$what.json; tail -n 3 gpurun_out/extras_nccl_$what.err
done
