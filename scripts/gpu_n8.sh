#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q -x --timeout 600 > gpurun_out/pytest_multigpu_n$N.log 2>&1; echo "multigpu pytest rc=$?"
tail -n 3 gpurun_out/pytest_multigpu_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29762 -m benchmarks.extras --only pagerank --pagerank-scale 26 > gpurun_out/extras_pagerank26_n$N.json 2> gpurun_out/extras_pagerank26_n$N.err; echo "pagerank26 rc=$?"
grep "^{" gpurun_out/extras_pagerank26_n$N.json; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/extras_pagerank26_n$N.err | tail -n 3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29763 -m benchmarks.extras --only pagerank --pagerank-scale 24 > gpurun_out/extras_pagerank24_n$N.json 2> gpurun_out/extras_pagerank24_n$N.err; echo "pagerank24 rc=$?"
grep "^{" gpurun_out/extras_pagerank24_n$N.json
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29760 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_full_n$N.json 2> gpurun_out/bench_full_n$N.err ) 2> gpurun_out/bench_full_n$N.time; echo "bench rc=$?"
grep "^{" gpurun_out/bench_full_n$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['extra']
print('ssgd', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
print({k:(round(v,3) if isinstance(v,float) else v) for k,v in e.items() if 'per_s' in k or 'gteps' in k or '_ms' in k or 'error' in k})"
