#!/bin/bash
# round 2: upper bound of "early registration" (table side of the next update at act time; no staleness handling, results
# not valid): tools/alt.bin is the estimate build, GRLCU_FORCE_SLOW=2 switches it on
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "kernel ms/pass %.3f"%d["roofline"]["kernel_ms_per_pass"], "W %.1f"%d["roofline"]["random_access"]["writes_per_agent_step"], "ret %.0f"%d["mean_last_return"])'
for rep in 1 2; do
python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "tree"
GRLCU_LIB=$PWD/tools/alt.bin python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "alt(off)"
GRLCU_LIB=$PWD/tools/alt.bin GRLCU_FORCE_SLOW=2 python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "alt(early registration estimate)"
done
