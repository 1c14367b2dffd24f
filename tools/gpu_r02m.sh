#!/bin/bash
# round 2: visit-ordered table layout (pilot run) for the tasks whose tiles do not fit the table: parity, then A/B
python -m pytest tests/test_gpu_cached.py tests/test_gpu_parity.py tests/test_gpu_free_running.py tests/test_gpu_full_size.py -m gpu -x -q 2>&1 | tail -4
WORKLOADS="cart_pole cart_double_pole pendulum sweep" bash tools/gpu_ab.sh
python bench.py --workload cart_pole --no-cpu --steps 3 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['roofline']['table_layout'])"
