#!/bin/bash
# round 2: lazy per-entry trace accumulators (default, GRLCU_MODE=0) against the exact cache (GRLCU_MODE=3)
python -m pytest tests/test_gpu_cached.py tests/test_gpu_free_running.py tests/test_gpu_full_size.py -m gpu -q -s 2>&1 | grep -E "lazy vs exact|agents equal|passed|failed|Error|assert" | tail -60
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "frac %.3f"%d["roofline"]["frac"], "kernel ms/pass %.3f"%d["roofline"]["kernel_ms_per_pass"])'
for w in pendulum cart_pole sweep cart_double_pole_ac; do
GRLCU_MODE=3 python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(exact cache)"
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(lazy cache)"
done
