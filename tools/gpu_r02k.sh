#!/bin/bash
# round 2: actor-critic tables interleaved as {critic, actor} pairs: parity, then same-box A/B against the previous build
python -m pytest tests/test_gpu_ac.py tests/test_gpu_cached.py tests/test_gpu_free_running.py tests/test_gpu_full_size.py tests/test_host_shims.py -m gpu -x -q 2>&1 | tail -5
WORKLOADS="cart_double_pole_ac pendulum_ac cart_pole_ac" bash tools/gpu_ab.sh
