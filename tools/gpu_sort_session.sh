#!/bin/bash
# One GPU call: sort correctness tests, A/B timing of the pass-kernel variants, per-kernel times.
set -u
mkdir -p gpurun_out
OUT=gpurun_out/sort_session.log
: > $OUT
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/sort_clocks.csv &
SMI=$!
echo "== tests" >> $OUT
timeout 900 python -m pytest tests/test_gpu_kernels.py -k sort tests/test_zz_no_alloc.py tests/test_gpu_multirank_single_gpu.py -x -q >> $OUT 2>&1
echo "pytest exit $?" >> $OUT
echo "== A/B" >> $OUT
B200LABS_SORT_IMPL=4 timeout 300 python tools/sort_bench.py --n 26,28 --dist uniform >> $OUT 2>&1
B200LABS_SORT_IMPL=5 B200LABS_SORT_BLOCKS=3 timeout 300 python tools/sort_bench.py --n 20,24,26,28 --dist uniform >> $OUT 2>&1
B200LABS_SORT_IMPL=5 B200LABS_SORT_BLOCKS=2 timeout 300 python tools/sort_bench.py --n 26,28 --dist uniform >> $OUT 2>&1
echo "== distributions (impl 5)" >> $OUT
timeout 600 python tools/sort_bench.py --n 26 --dist sorted,reverse,constant,small,bits16 >> $OUT 2>&1
timeout 300 python tools/sort_bench.py --n 26 --dist uniform,small --dtype float32 >> $OUT 2>&1
B200LABS_SORT_IMPL=4 timeout 600 python tools/sort_bench.py --n 26 --dist sorted,constant,small >> $OUT 2>&1
echo "== per-kernel times" >> $OUT
timeout 300 python tools/sort_bench.py --n 26 --dist uniform --iters 2 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/sort_launches.csv python tools/sort_bench.py --n 26 --dist uniform --iters 2 > gpurun_out/ncu.log 2>&1
kill $SMI
tail -5 gpurun_out/sort_launches.csv >> $OUT 2>/dev/null
cat $OUT | tail -60
