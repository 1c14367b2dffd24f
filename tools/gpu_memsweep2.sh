#!/bin/bash
# pendulum: tiny tables (L2-resident) and agent-count sweep, to see what bounds the kernel when the table gathers are cheap
for cfg in "65536 32" "65536 64" "65536 128" "65536 256" "32768 64" "32768 65536" "16384 64" "16384 65536" "131072 64" "131072 32768"; do
set -- $cfg
python bench.py --steps 3 --warmup 3 --no-cpu --agents $1 --memory $2 2>&1 | python -c "
import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; ra=r['random_access']
print('agents %6d memory %6d  tables %8.1f MiB  value %.3e  e2e %.3e  V %.1f  W %.1f' % ($1, $2, $1*$2*8/2**20, d['value'], d['e2e']['value'], r['V_mean'], ra['writes_per_agent_step']))"
done
