#!/bin/bash
# round-2 call B (2 GPUs): full GPU test-suite + SSGD headline with the slice-parallel finish (PDL on / off)
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/b_pytest.log
for n in 1 2; do
  for pdl in 1 0; do
    DAPB_LR_PDL=$pdl timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29651 \
      bench.py --gpus $n --steps 40 --warmup 5 --no-extras --no-e2e > gpurun_out/b_bench_n${n}_pdl$pdl.json 2> gpurun_out/b_bench_n${n}_pdl$pdl.err
    echo "bench n=$n pdl=$pdl rc=$?"; python -c "
import json,sys
try:
    d=json.loads(open('gpurun_out/b_bench_n${n}_pdl$pdl.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d.get('roofline_frac_of_measured_hbm'))
except Exception as e: print('parse fail', e)
"
  done
done
