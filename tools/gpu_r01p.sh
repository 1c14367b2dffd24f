#!/bin/bash
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for w in pendulum cart_pole cart_double_pole_ac sweep; do
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w', '%.4e'%d['value'], '%.4e'%d['e2e']['value'], 'frac %.3f'%d['roofline']['frac'], 'ra %.3f'%d['roofline']['random_access']['frac'])"
done
