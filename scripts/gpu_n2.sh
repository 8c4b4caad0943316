#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 600 > gpurun_out/mg2_tests_n$N.log 2>&1; echo "multigpu tests rc=$?"
tail -n 6 gpurun_out/mg2_tests_n$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 benchmarks/extras.py --only pagerank,tc --pagerank-scale 24 > gpurun_out/extras_pr_tc_n$N.json 2> gpurun_out/extras_pr_tc_n$N.err; echo "extras rc=$?"; tail -n 1 gpurun_out/extras_pr_tc_n$N.json
