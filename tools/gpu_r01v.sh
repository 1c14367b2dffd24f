#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
grl_b200/host/grld_b200 cfg/cart_double_pole/ac_tc_cuda.yaml 2>&1 | tail -14
