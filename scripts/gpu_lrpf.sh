#!/bin/bash
mkdir -p gpurun_out
for pf in 1 2 default; do
[ $pf = default ] && unset DAPB_LR_PREFETCH || export DAPB_LR_PREFETCH=$pf
timeout 600 python bench.py --steps 20 --warmup 5 --no-extras --no-e2e > gpurun_out/lrpf_${pf}_bf16.json 2> gpurun_out/lrpf_${pf}_bf16.err; echo "pf=$pf bf16 rc=$? $(python -c "
import json;d=json.loads([l for l in open('gpurun_out/lrpf_${pf}_bf16.json') if l.startswith('{')][-1]);print(d['value'], d['ms_per_step'])")"
done
unset DAPB_LR_PREFETCH
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -m gpu -q -k "lr or logistic or ssgd or step or alloc" --timeout 600 > gpurun_out/lrpf_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/lrpf_tests.log
