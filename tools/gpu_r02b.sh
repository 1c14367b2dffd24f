#!/bin/bash
# round 2: how the tile-major kernel responds to resident agents / register budget (one box, back to back)
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "kernel ms/launch %.3f"%d["roofline"]["kernel_ms_per_launch"])'
run() { tag=$1; shift; "$@" python bench.py --no-cpu --steps 5 --warmup 3 ${ARGS} | python -c "$fmt" "$tag"; }
ARGS="" run "default(65536 agents, 128 regs)" env
ARGS="--cohort 32768" run "cohort 32768 (255 regs, eager close)" env
ARGS="--agents 32768" run "32768 agents (255 regs)" env
ARGS="--agents 131072" run "131072 agents (128 regs)" env
ARGS="--agents 98304" run "98304 agents (128 regs)" env
ARGS="" run "block 32" env GRLCU_BLOCK=32
ARGS="--agents 32768" run "32768 agents forced 128 regs" env GRLCU_MINB7=1
