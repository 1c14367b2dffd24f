#!/bin/bash
# same-box comparison of several builds of libgrlcu.so: "tree" (in-tree) and every file named in $LIBS
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "kernel ms/pass %.3f"%d["roofline"]["kernel_ms_per_pass"])'
for rep in $(seq ${REPS:-1}); do
for w in ${WORKLOADS:-pendulum}; do
for lib in ${LIBS:-tools/prev.bin}; do
GRLCU_LIB=$PWD/$lib python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w($lib)"
done
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w(tree)"
done
done
