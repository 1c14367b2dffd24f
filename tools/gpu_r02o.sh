#!/bin/bash
# round 2: sincos in the cart-pole / double-pole dynamics (tree) and two-entry flushes in trace clears (alt.bin) against HEAD (prev.bin)
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"])'
python -m pytest tests/test_gpu_parity.py tests/test_gpu_ac.py -m gpu -x -q 2>&1 | tail -2
GRLCU_LIB=$PWD/tools/alt.bin python -m pytest tests/test_gpu_cached.py tests/test_gpu_free_running.py -m gpu -x -q 2>&1 | tail -2
for w in cart_double_pole_ac cart_pole cart_double_pole cart_pole_ac; do
GRLCU_LIB=$PWD/tools/prev.bin python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(prev)"
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(sincos)"
GRLCU_LIB=$PWD/tools/alt.bin python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(sincos + 2-entry flush)"
done
