#!/bin/bash
mkdir -p gpurun_out
N=${1:-1}
if [ "$N" = "1" ]; then
  ( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_full_n1.json 2> gpurun_out/bench_full_n1.err ) 2> gpurun_out/bench_full_n1.time; echo "bench rc=$?"
else
  ( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29760 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_full_n$N.json 2> gpurun_out/bench_full_n$N.err ) 2> gpurun_out/bench_full_n$N.time; echo "bench rc=$?"
fi
cat gpurun_out/bench_full_n$N.json; tail -n 5 gpurun_out/bench_full_n$N.err; cat gpurun_out/bench_full_n$N.time
