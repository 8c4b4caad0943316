#!/bin/bash
mkdir -p gpurun_out
for cfg in "32 2048" "64 2048" "128 2048" "256 2048" "512 2048"; do
set -- $cfg
DAPB_PR_HUB_THRESHOLD=$1 DAPB_PR_HUB_CHUNK=$2 timeout 600 python benchmarks/extras.py --only pagerank --pagerank-scale 24 > gpurun_out/pr3_$1_$2.json 2> gpurun_out/pr3_$1_$2.err; echo "thr=$1 chunk=$2 rc=$? $(cut -c1-90 gpurun_out/pr3_$1_$2.json)"
done
for thr in 64 256 512; do
DAPB_PR_HUB_THRESHOLD=$thr timeout 900 python benchmarks/extras.py --only pagerank --pagerank-scale 26 > gpurun_out/pr3_s26_$thr.json 2> gpurun_out/pr3_s26_$thr.err; echo "s26 thr=$thr rc=$? $(cut -c1-90 gpurun_out/pr3_s26_$thr.json)"
done
