#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cached.py -x -q 2>&1 | tail -15
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --no-cpu --steps 5 --warmup 3 > gpurun_out/bench_v5.json 2> gpurun_out/bench_v5.err; tail -1 gpurun_out/bench_v5.json
GRLCU_MODE=1 python bench.py --no-cpu --steps 3 --warmup 2 > gpurun_out/bench_uncached.json 2>&1; tail -1 gpurun_out/bench_uncached.json | cut -c1-200
