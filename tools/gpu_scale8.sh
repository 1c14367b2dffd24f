#!/bin/bash
# BASELINE configs[2] and [4] at their full 8-GPU size (and the pendulum again), launched like the driver does
mkdir -p gpurun_out
port=29540
for w in cart_pole sweep pendulum cart_double_pole_ac; do
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus 8 --workload $w --steps 5 --warmup 3 > gpurun_out/scale8_$w.json 2> gpurun_out/scale8_$w.err; echo "$w exit $?"
  python -c "
import json
d=json.loads(open('gpurun_out/scale8_$w.json').read().strip().splitlines()[-1]); print('$w', d['n_gpus'], '%.4e'%d['value'], '%.4e'%d['e2e']['value'], d['config']['agents_total'])"
done
