#!/bin/bash
# same-box A/B of the in-tree library against an alternative build ($ALT, default tools/alt2.bin)
ALT=${ALT:-tools/alt2.bin}
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "kernel ms/pass %.3f"%d["roofline"]["kernel_ms_per_pass"])'
for rep in 1 2; do
for w in ${WORKLOADS:-pendulum}; do
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(tree)"
GRLCU_LIB=$PWD/$ALT python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w($ALT)"
done
done
