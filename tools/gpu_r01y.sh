#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
WORKLOADS="pendulum cart_pole cart_double_pole_ac sweep" bash tools/gpu_ab.sh
