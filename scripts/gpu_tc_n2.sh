#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 600 > gpurun_out/mg_tests_n$N.log 2>&1; echo "multigpu tests rc=$?"
tail -n 6 gpurun_out/mg_tests_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29731 benchmarks/extras.py --only tc > gpurun_out/tc_persist_n$N.json 2> gpurun_out/tc_persist_n$N.err; echo "persistent rc=$?"; tail -n 1 gpurun_out/tc_persist_n$N.json
DAPB_TC_ROUNDS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29732 benchmarks/extras.py --only tc > gpurun_out/tc_rounds_n$N.json 2> gpurun_out/tc_rounds_n$N.err; echo "rounds rc=$?"; tail -n 1 gpurun_out/tc_rounds_n$N.json
