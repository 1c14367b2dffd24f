#!/bin/bash
# pendulum, 65,536 agents: throughput against the per-agent table size (L2-resident tables at the small end)
for mem in 512 1024 2048 4096 8192 16384 32768 65536; do
python bench.py --steps 3 --warmup 3 --no-cpu --agents 65536 --memory $mem 2>&1 | python -c "
import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; ra=r['random_access']
print('memory %6d  table bytes %5.1f MiB/agent-set  value %.3e  e2e %.3e  V %.1f  W %.1f  byte-frac %.3f' % ($mem, 65536*$mem*8/2**20, d['value'], d['e2e']['value'], r['V_mean'], ra['writes_per_agent_step'], r['frac']))"
done
