#!/bin/bash
mkdir -p gpurun_out
for k in cuda tc cuda tc; do
  timeout 600 python bench.py --steps 20 --warmup 3 --no-extras --no-e2e --lr-kernel $k > gpurun_out/bench_lr_$k.json 2> gpurun_out/bench_lr_$k.err; echo "bench $k rc=$?"
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_lr_$k.json') if l.startswith('{')][-1])
print('$k', round(d['ms_per_step'],4), round(d['hbm_gbs_per_gpu'],1), round(d['roofline_frac_of_measured_hbm'],3), d['clocks'].get('sm_mhz'))"
  tail -n 2 gpurun_out/bench_lr_$k.err
done
for v in skip noskip skip noskip; do
  if [ $v = noskip ]; then export DAPB_PR_NOSKIP=1; else unset DAPB_PR_NOSKIP; fi
  timeout 600 python -m benchmarks.extras --only pagerank 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', round(d['pagerank_gteps'],1), round(d['pagerank_ms_per_iter'],3))"
done
