#!/bin/bash
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 1500 python -m pytest tests/test_multigpu.py -m gpu -q -x --timeout 900 > gpurun_out/pytest_multigpu.log 2>&1; echo "multigpu pytest rc=$?"
tail -n 40 gpurun_out/pytest_multigpu.log
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29750 bench.py --gpus $N --steps 10 --warmup 3 --no-extras > gpurun_out/bench_ours_n$N.json 2> gpurun_out/bench_ours_n$N.err; echo "bench N=$N rc=$?"
cat gpurun_out/bench_ours_n$N.json; tail -n 5 gpurun_out/bench_ours_n$N.err
