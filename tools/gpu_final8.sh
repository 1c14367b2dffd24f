#!/bin/bash
# final multi-GPU pass on one 8 x B200 box, launched like the driver does: the headline config at N = 1 and N = 8, the other
# three BASELINE workloads at their full 8-GPU size, and the plugin path (grld_b200, one process, 8 contexts, NCCL
# all-gather) on BASELINE configs[4]: 1,048,576 sweep agents
TAG=${TAG:-r02}
mkdir -p gpurun_out
fmt="import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d['n_gpus'], '%.4e'%d['value'], '%.4e'%d['e2e']['value'], d['config']['agents_total'], 'collective_ms', d.get('collective_ms'))"
python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_n1_$TAG.json 2> gpurun_out/scale_n1_$TAG.err; echo "n1 exit $?"
python -c "$fmt" gpurun_out/scale_n1_$TAG.json
port=29560
for n in ${NS:-8}; do
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
# the plugin path on all 8 GPUs
sed -e 's/agents: 131072/agents: 1048576/' -e 's/steps: 10000/steps: 1000/' cfg/pendulum/sarsa_tc_sweep_cuda.yaml > gpurun_out/sweep8.yaml
echo "  output: gpurun_out/grld_sweep8_$TAG" >> gpurun_out/sweep8.yaml
( time timeout 900 grl_b200/host/grld_b200 -v gpurun_out/sweep8.yaml ) > gpurun_out/grld_sweep8_$TAG.log 2>&1; echo "grld exit $?"
tail -6 gpurun_out/grld_sweep8_$TAG.log
head -c 2000 gpurun_out/grld_sweep8_$TAG-0.txt | head -12
rm -f gpurun_out/grld_sweep8_$TAG-0.txt gpurun_out/grld_sweep8_$TAG-0-cells.txt.bak
python -m pytest tests/test_host_shims.py -m gpu -q -k two_gpus 2>&1 | tail -2
