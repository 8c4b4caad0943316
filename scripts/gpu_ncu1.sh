#!/bin/bash
mkdir -p gpurun_out
CMD1="python tests/_tc_worker.py kmeans 8000000 128 1024 0"
$CMD1 > gpurun_out/ncu_plain_kmeans.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:kmeans_tc_kernel -c 1 -f -o gpurun_out/prof_kmeans_tc $CMD1 > gpurun_out/ncu_kmeans.log 2>&1
echo "ncu kmeans rc=$?"; tail -n 3 gpurun_out/ncu_kmeans.log
CMD2="python bench.py --steps 3 --warmup 1 --no-extras --no-e2e --rows-per-gpu 4194304"
$CMD2 > gpurun_out/ncu_plain_lr.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lr_step_kernel -s 2 -c 1 -f -o gpurun_out/prof_lr_step $CMD2 > gpurun_out/ncu_lr.log 2>&1
echo "ncu lr rc=$?"; tail -n 3 gpurun_out/ncu_lr.log
ls -la gpurun_out/*.ncu-rep
