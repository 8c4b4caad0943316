#!/bin/bash
# round-2 call J (1 GPU): full GPU test-suite, smoke, full bench at N = 1, launch list under ncu
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/j_pytest.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/j_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/j_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/j_smoke.log
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/j_bench_n1.json 2> gpurun_out/j_bench_n1.err ) 2> gpurun_out/j_bench_n1.time; echo "bench n=1 rc=$?"; tail -3 gpurun_out/j_bench_n1.time
python - <<'PY'
import json
d = json.loads(open("gpurun_out/j_bench_n1.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["sustained_ms_per_step"], d["vs_baseline"], d["checks"])
print({k: v for k, v in d["extra"].items() if "per_rank" not in k})
PY
timeout 300 python __graft_entry__.py smoke > gpurun_out/j_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/j_smoke_launches.csv python __graft_entry__.py smoke > gpurun_out/j_ncu.log 2>&1; echo "ncu rc=$?"
