#!/bin/bash
# r01 pass after the actor-critic siblings (pendulum / cart-pole AC, end_stop_penalty): parity, the two new bench
# workloads, and a same-box A/B of the four existing workloads against the previous build (tools/prev.bin)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs -x > gpurun_out/pytest_gpu_v10.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_v10.log
tail -3 gpurun_out/pytest_gpu_v10.log
for w in pendulum_ac cart_pole_ac; do
  python bench.py --workload $w --no-cpu > gpurun_out/bench_${w}_v10.json 2> gpurun_out/bench_${w}_v10.err; echo "$w exit $?"
done
WORKLOADS="pendulum cart_pole cart_double_pole_ac sweep" bash tools/gpu_ab.sh > gpurun_out/ab_v10.txt 2>&1
cat gpurun_out/ab_v10.txt
