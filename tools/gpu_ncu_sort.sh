#!/bin/bash
# ncu --set full of the sort pass kernels (first pass + one middle pass), after a plain run
mkdir -p gpurun_out
timeout 300 python tools/sort_bench.py --n 26 --dist uniform --iters 1 > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sort_pass_kernel -s 4 -c 2 \
    -o gpurun_out/prof_lab5_v5 -f python tools/sort_bench.py --n 26 --dist uniform --iters 1 > gpurun_out/ncu.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/ncu.log
