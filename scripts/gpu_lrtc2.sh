#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py -m gpu -q --timeout 300 -k lr_tc > gpurun_out/pytest_lrtc.log 2>&1; echo "lr_tc pytest rc=$?"
tail -n 3 gpurun_out/pytest_lrtc.log | cut -c1-300
for k in tc cuda; do
  timeout 600 python bench.py --steps 20 --warmup 3 --no-extras --no-e2e --lr-kernel $k > gpurun_out/bench_lr_$k.json 2> gpurun_out/bench_lr_$k.err; echo "bench $k rc=$?"
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_lr_$k.json') if l.startswith('{')][-1])
print('$k', round(d['ms_per_step'],4), round(d['hbm_gbs_per_gpu'],1), round(d['roofline_frac_of_measured_hbm'],3), d['clocks'].get('sm_mhz'))"
done
