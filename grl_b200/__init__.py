"""grl_b200 -- B200-native batched inner learning loop of wcaarls/grl (one hot path).

The product is libgrlcu.so (grl_b200/csrc, C ABI in include/grlcu.h) plus the C++ Configurable shims in
grl_b200/host.  This Python package is only the ctypes harness used by tests/ and bench.py.
"""
from .capi import Batch, Config, GrlcuError, load, pendulum_config, cart_pole_config, cart_double_pole_config, cart_double_pole_ac_config, pendulum_ac_config, cart_pole_ac_config  # noqa: F401
