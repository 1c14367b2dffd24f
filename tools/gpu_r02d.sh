#!/bin/bash
# round 2: cohort-blocked trace value cache + persisting-L2 window on the cohort's block (GRLCU_PERSIST)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cached.py -m gpu -x -q 2>&1 | tail -3
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "frac %.3f"%d["roofline"]["frac"])'
for rep in 1 2; do
python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "pendulum"
GRLCU_PERSIST=1 python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "pendulum(persist)"
done
python bench.py --workload sweep --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "sweep"
GRLCU_PERSIST=1 python bench.py --workload cart_double_pole_ac --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "ac(persist)"
python bench.py --workload cart_double_pole_ac --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "ac"
