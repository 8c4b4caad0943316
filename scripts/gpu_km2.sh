#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py tests/test_gpu_kernels.py -m gpu -q -k "kmeans or state_dict or transitive" --timeout 600 > gpurun_out/km2_tests.log 2>&1; echo "tests rc=$?"
tail -n 8 gpurun_out/km2_tests.log
timeout 300 python scripts/kmeans_tc_probe.py 32000000 > gpurun_out/km2_probe.log 2>&1; echo "probe rc=$?"; cat gpurun_out/km2_probe.log
timeout 600 python benchmarks/extras.py --only kmeans > gpurun_out/km2_extras.json 2> gpurun_out/km2_extras.err; echo "extras rc=$?"; cat gpurun_out/km2_extras.json
