#!/bin/bash
# parity + the four workloads, and A/B of the 128-register instantiation on the 32,768-agent workloads
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "frac %.3f"%d["roofline"]["frac"], "ra %.3f"%d["roofline"]["random_access"]["frac"])'
for w in pendulum cart_pole cart_double_pole_ac sweep; do
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | tee gpurun_out/bench_$w.json | python -c "$fmt" $w
done
for w in cart_pole cart_double_pole_ac; do
GRLCU_MINB7=1 python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(minb7)"
done
