#!/bin/bash
# round-2 call I (8 GPUs): full bench at N = 8 (+ headline-only N = 1 on the same box)
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29670 \
  bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/i_bench_n8.json 2> gpurun_out/i_bench_n8.err ) 2> gpurun_out/i_bench_n8.time; echo "bench n=8 rc=$?"; tail -3 gpurun_out/i_bench_n8.time
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-extras --no-e2e > gpurun_out/i_bench_n1_head.json 2> gpurun_out/i_bench_n1_head.err; echo "bench n=1 rc=$?"
python - <<'PY'
import json
for f in ("i_bench_n8", "i_bench_n1_head"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["sustained_ms_per_step"], d["vs_baseline"], d["checks"])
        print({k: v for k, v in d["extra"].items() if "per_rank" not in k})
        print(d.get("finish_breakdown_last_step"))
        print(d.get("e2e"))
    except Exception as e:
        print(f, "parse fail", e)
PY
