#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cached.py -x -q 2>&1 | tail -3
GRLCU_DEBUG=1 python bench.py --no-cpu --steps 3 --warmup 3 > gpurun_out/bench_v5.json 2> gpurun_out/bench_v5.err; tail -1 gpurun_out/bench_v5.json | cut -c1-150; tail -2 gpurun_out/bench_v5.err
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_run -s 3 -c 1 -o gpurun_out/prof_k_run_v5 \
    python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full_v5.log 2>&1
