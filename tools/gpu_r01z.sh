#!/bin/bash
# r01 pass after the sibling configurations (pendulum / cart-pole AC, cart-double-pole SARSA, end_stop_penalty):
# parity and the bench lines of the new workloads
TAG=${TAG:-v10}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs -x > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_${TAG}.log
tail -3 gpurun_out/pytest_gpu_${TAG}.log
for w in ${WORKLOADS:-cart_double_pole pendulum_ac cart_pole_ac}; do
  python bench.py --workload $w --no-cpu > gpurun_out/bench_${w}_${TAG}.json 2> gpurun_out/bench_${w}_${TAG}.err; echo "$w exit $?"
done
