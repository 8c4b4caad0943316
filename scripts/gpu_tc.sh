#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "transitive or entry_scripts" --timeout 300 > gpurun_out/tc_tests.log 2>&1; echo "tc tests rc=$?"
tail -n 8 gpurun_out/tc_tests.log
timeout 300 python benchmarks/extras.py --only tc > gpurun_out/tc_persist.json 2> gpurun_out/tc_persist.err; echo "persistent rc=$?"; cat gpurun_out/tc_persist.json
DAPB_TC_ROUNDS=1 timeout 300 python benchmarks/extras.py --only tc > gpurun_out/tc_rounds.json 2> gpurun_out/tc_rounds.err; echo "rounds rc=$?"; cat gpurun_out/tc_rounds.json
