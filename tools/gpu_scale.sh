#!/bin/bash
# weak-scaling run on one box: N = 1, 2, 4, 8 (as many as are visible), driver-style launch
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo "n1 exit $?"
port=29520
for n in 2 4 8; do
  [ $n -le $NG ] || continue
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "n$n exit $?"
done
for f in gpurun_out/scale_n*.json; do python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['n_gpus'], '%.4e'%d['value'], '%.4e'%d['e2e']['value'], d['config']['agents_total'])"; done
