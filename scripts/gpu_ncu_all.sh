#!/bin/bash
# One ncu capture per hot kernel (each only after the identical plain command exited 0).
mkdir -p gpurun_out
run() {  # name regex skip cmd...
  local name=$1 regex=$2 skip=$3; shift 3
  "$@" > gpurun_out/ncu_plain_$name.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 -f -o gpurun_out/prof_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
}
run kmeans_tc kmeans_tc_kernel 0 python tests/_tc_worker.py kmeans 8000000 128 1024 0
run gram_tc gram_tc_kernel 0 python tests/_tc_worker.py gram 2000000 512 0
run lr_step lr_step_kernel 2 python bench.py --steps 3 --warmup 1 --no-extras --no-e2e --rows-per-gpu 8388608
run lr_tc lr_tc_kernel 2 python bench.py --steps 3 --warmup 1 --no-extras --no-e2e --rows-per-gpu 8388608 --lr-kernel tc
run pagerank_warp pagerank_warp_kernel 3 python -m benchmarks.extras --only pagerank --pagerank-scale 24
run pagerank_hub pagerank_hub_partial_kernel 3 python -m benchmarks.extras --only pagerank --pagerank-scale 24
run pagerank pagerank_step_kernel 3 env DAPB_PR_LAYOUT=cta python -m benchmarks.extras --only pagerank --pagerank-scale 22
run tc_closure tc_closure_kernel 1 python -m benchmarks.extras --only tc
run mc_pi mc_pi_kernel 0 python -m benchmarks.extras --only mc
ls -la gpurun_out/*.ncu-rep
