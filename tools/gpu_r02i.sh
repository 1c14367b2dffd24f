#!/bin/bash
# round 2: trace value cache in shared memory (GRLCU_SMEM=1): parity, then bench
GRLCU_SMEM=1 python -m pytest tests/test_gpu_cached.py tests/test_gpu_free_running.py -m gpu -x -q 2>&1 | tail -3
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "kernel ms/pass %.3f"%d["roofline"]["kernel_ms_per_pass"])'
for rep in 1 2; do
python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "pendulum(global cache)"
GRLCU_SMEM=1 python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "pendulum(smem cache)"
GRLCU_SMEM=1 GRLCU_BLOCK=32 python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "pendulum(smem cache, block 32)"
done
GRLCU_SMEM=1 python bench.py --workload cart_pole --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "cart_pole(smem cache)"
python bench.py --workload cart_pole --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "cart_pole(global cache)"
