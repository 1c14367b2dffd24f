#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
WORKLOADS="sweep pendulum" bash tools/gpu_ab.sh
