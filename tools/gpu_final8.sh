#!/bin/bash
# final multi-GPU pass on one 8 x B200 box, launched like the driver does: weak scaling of the headline config
# (N = 1, 2, 4, 8), then the other three BASELINE workloads at their full 8-GPU size
TAG=${TAG:-v9}
mkdir -p gpurun_out
fmt="import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d['n_gpus'], '%.4e'%d['value'], '%.4e'%d['e2e']['value'], d['config']['agents_total'])"
python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_n1_$TAG.json 2> gpurun_out/scale_n1_$TAG.err; echo "n1 exit $?"
python -c "$fmt" gpurun_out/scale_n1_$TAG.json
port=29560
for n in 2 4 8; do
  port=$((port+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/scale_n${n}_$TAG.json 2> gpurun_out/scale_n${n}_$TAG.err; echo "n$n exit $?"
  python -c "$fmt" gpurun_out/scale_n${n}_$TAG.json
done
for w in cart_pole sweep cart_double_pole_ac; do
  port=$((port+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus 8 --workload $w --steps 5 --warmup 3 > gpurun_out/scale8_${w}_$TAG.json 2> gpurun_out/scale8_${w}_$TAG.err; echo "$w exit $?"
  python -c "$fmt" gpurun_out/scale8_${w}_$TAG.json
done
