#!/bin/bash
# one thread per agent (k_run) in cohorts: does L2 reuse pay before latency bites?
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"])'
for cc in 0 32768 16384 12288 8192 6144 4096; do
python bench.py --no-cpu --steps 3 --warmup 3 --cohort $cc | python -c "$fmt" "pendulum k_run cohort=$cc"
done
