#!/bin/bash
# round 2, first layout check: parity of the tile-major layout, then same-box A/B against the previous build
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_cached.py tests/test_gpu_free_running.py -m gpu -x -q -s 2>&1 | tail -15 > gpurun_out/r02a_pytest.log
cat gpurun_out/r02a_pytest.log
WORKLOADS="pendulum sweep" bash tools/gpu_ab.sh 2>&1 | tee gpurun_out/r02a_ab.txt
