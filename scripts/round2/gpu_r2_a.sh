#!/bin/bash
# round-2 call A (2 GPUs): NVLS probe + new collectives + persistent PageRank
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/a_gpus.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/a_multigpu.log 2>&1; echo "multigpu rc=$?" 
tail -5 gpurun_out/a_multigpu.log
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -k "pagerank or entry_scripts or resume" > gpurun_out/a_pr_tests.log 2>&1; echo "pr tests rc=$?"
tail -5 gpurun_out/a_pr_tests.log
for n in 1 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29650 \
    -m benchmarks.extras --only pagerank --pagerank-scale 24 --pagerank-iters 30 > gpurun_out/a_pr24_n$n.json 2> gpurun_out/a_pr24_n$n.err
  echo "pr24 n=$n rc=$?"; tail -c 1500 gpurun_out/a_pr24_n$n.json
done
