#!/bin/bash
set -x
python bench.py --workload cart_pole --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_cp.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_run -s 3 -c 1 -o gpurun_out/prof_k_run_cartpole \
    python bench.py --workload cart_pole --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full_cp.log 2>&1
tail -2 gpurun_out/ncu_full_cp.log
