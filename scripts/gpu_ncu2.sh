#!/bin/bash
mkdir -p gpurun_out
CMD1="python tests/_tc_worker.py kmeans 8000000 128 1024 0"
$CMD1 > gpurun_out/ncu_plain_kmeans.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:kmeans_tc_kernel -c 1 -f -o gpurun_out/prof_kmeans_tc $CMD1 > gpurun_out/ncu_kmeans.log 2>&1
echo "ncu kmeans rc=$?"; tail -n 2 gpurun_out/ncu_kmeans.log
