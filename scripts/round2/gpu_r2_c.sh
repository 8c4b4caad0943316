#!/bin/bash
# round-2 call C (2 GPUs): tests + smoke + reworked bench (baselines, checks, full configs)
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/c_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/c_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/c_smoke.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29652 \
  -m benchmarks.extras --only tc,mc --baseline > gpurun_out/c_tc_mc_n2.json 2> gpurun_out/c_tc_mc_n2.err; echo "tc/mc n=2 rc=$?"; tail -c 1200 gpurun_out/c_tc_mc_n2.json
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/c_bench_n1.json 2> gpurun_out/c_bench_n1.err ) 2> gpurun_out/c_bench_n1.time; echo "bench n=1 rc=$?"; tail -3 gpurun_out/c_bench_n1.time
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29653 \
  bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/c_bench_n2.json 2> gpurun_out/c_bench_n2.err ) 2> gpurun_out/c_bench_n2.time; echo "bench n=2 rc=$?"; tail -3 gpurun_out/c_bench_n2.time
python - <<'PY'
import json
for n in (1, 2):
    try:
        d = json.loads(open(f"gpurun_out/c_bench_n{n}.json").read().strip().splitlines()[-1])
        print(n, d["ms_per_step"], d["sustained_ms_per_step"], d["vs_baseline"], d["checks"])
        print({k: v for k, v in d["extra"].items() if "per_rank" not in k})
    except Exception as e:
        print(n, "parse fail", e)
PY
