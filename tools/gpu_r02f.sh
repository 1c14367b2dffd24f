#!/bin/bash
# round 2: Q(p) read at the end of the iteration that chose p ("registration"): parity, then A/B on one box
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cached.py tests/test_gpu_free_running.py tests/test_gpu_full_size.py -m gpu -x -q 2>&1 | tail -4
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "frac %.3f"%d["roofline"]["frac"])'
for rep in 1 2; do
for w in pendulum cart_pole sweep; do
GRLCU_REGISTER=0 python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(register 0)"
GRLCU_REGISTER=1 python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(register 1)"
done
done
