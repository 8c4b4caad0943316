#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}; SC=${2:-26}
for np in 0 1; do
DAPB_PR_NOPUSH=$np timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2975$np benchmarks/extras.py --only pagerank --pagerank-scale $SC > gpurun_out/prpush_${np}_n$N.json 2> gpurun_out/prpush_${np}_n$N.err; echo "nopush=$np rc=$? $(grep '^{' gpurun_out/prpush_${np}_n$N.json | cut -c1-90)"
done
