#!/bin/bash
# round 2: latency- or throughput-bound?  throughput against resident agents at 255 registers; whole-line gathers
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "kernel ms/pass %.3f"%d["roofline"]["kernel_ms_per_pass"])'
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "run_host or replay_pendulum" 2>&1 | tail -2
for n in 8192 16384 24576 32768 37888 65536; do
python bench.py --agents $n --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "agents $n"
done
for rep in 1 2; do
python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "64B gathers"
GRLCU_LIB=$PWD/tools/alt.bin python bench.py --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "128B gathers"
done
