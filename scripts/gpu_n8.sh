#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q -x --timeout 600 > gpurun_out/pytest_multigpu_n$N.log 2>&1; echo "multigpu pytest rc=$?"
tail -n 5 gpurun_out/pytest_multigpu_n$N.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29760 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_full_n$N.json 2> gpurun_out/bench_full_n$N.err ) 2> gpurun_out/bench_full_n$N.time; echo "bench rc=$?"
cat gpurun_out/bench_full_n$N.json; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/bench_full_n$N.err | tail -n 5; tail -n 3 gpurun_out/bench_full_n$N.time
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29761 bench.py --gpus $N --steps 10 --warmup 3 --impl nccl --no-extras --no-e2e > gpurun_out/bench_nccl_n$N.json 2> gpurun_out/bench_nccl_n$N.err; echo "bench nccl rc=$?"
cat gpurun_out/bench_nccl_n$N.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29762 -m benchmarks.extras --only pagerank --pagerank-scale 26 > gpurun_out/extras_pagerank26_n$N.json 2> gpurun_out/extras_pagerank26_n$N.err; echo "pagerank26 rc=$?"
cat gpurun_out/extras_pagerank26_n$N.json; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/extras_pagerank26_n$N.err | tail -n 3
