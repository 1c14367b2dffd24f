#!/bin/bash
set -x
mkdir -p gpurun_out
for w in cart_pole cart_double_pole_ac sweep; do
  python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "exit $?"; tail -1 gpurun_out/bench_$w.json | cut -c1-120; tail -3 gpurun_out/bench_$w.err
done
