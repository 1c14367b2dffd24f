#!/bin/bash
# round 2: per-agent-lambda contexts launched in lambda-sorted 255-register cohorts (chunk map) against one 128-register launch
python -m pytest tests/test_gpu_cached.py tests/test_gpu_full_size.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "frac %.3f"%d["roofline"]["frac"])'
for rep in 1 2; do
GRLCU_NO_CHUNK_MAP=1 python bench.py --workload sweep --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "sweep(one launch, 128 regs)"
python bench.py --workload sweep --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "sweep(lambda-sorted cohorts, 255 regs)"
done
