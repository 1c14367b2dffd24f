#!/bin/bash
# round 2: phase-1 (slot computation) loop unrolled by 2 / 4 so that the slotmap / perm lookups of several tilings are in flight together
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"])'
for w in pendulum cart_pole sweep; do
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(unroll 1)"
GRLCU_LIB=$PWD/tools/unroll2.bin python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(unroll 2)"
GRLCU_LIB=$PWD/tools/unroll4.bin python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(unroll 4)"
done
