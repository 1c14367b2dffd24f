#!/bin/bash
# A/B of two builds of libgrlcu.so on the same box: tools/prev.bin (GRLCU_LIB) vs the in-tree build
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"])'
for rep in $(seq ${REPS:-2}); do
for w in ${WORKLOADS:-pendulum}; do
GRLCU_LIB=$PWD/tools/prev.bin python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(prev)"
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(new)"
done
done
