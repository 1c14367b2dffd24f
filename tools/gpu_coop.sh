#!/bin/bash
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"])'
for w in pendulum; do
for L in 2 4 8; do
for cc in 8192 12288 16384 24576 32768; do
GRLCU_MODE=2 GRLCU_COOP_LANES=$L GRLCU_COOP_COHORT=$cc timeout 300 python bench.py --workload $w --no-cpu --steps 3 --warmup 3 | python -c "$fmt" "$w(coop L=$L C=$cc regs128)"
done
done
done
GRLCU_MODE=2 GRLCU_COOP_LANES=4 GRLCU_COOP_COHORT=16384 timeout 300 python bench.py --workload cart_pole --no-cpu --steps 3 --warmup 3 | python -c "$fmt" "cart_pole(coop L=4 C=16384 regs128)"
GRLCU_MODE=2 GRLCU_COOP_LANES=2 GRLCU_COOP_COHORT=32768 timeout 300 python bench.py --workload cart_pole --no-cpu --steps 3 --warmup 3 | python -c "$fmt" "cart_pole(coop L=2 C=32768 regs128)"
