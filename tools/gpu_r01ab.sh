#!/bin/bash
# parity, then same-box A/B of tools/prev.bin against the in-tree build on the workloads in $WORKLOADS
TAG=${TAG:-ab}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs -x > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_${TAG}.log
tail -3 gpurun_out/pytest_gpu_${TAG}.log
bash tools/gpu_ab.sh > gpurun_out/ab_${TAG}.txt 2>&1
cat gpurun_out/ab_${TAG}.txt
