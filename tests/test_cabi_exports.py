"""CPU checks of the drop-in boundary: libgrlcu.so builds for sm_100a, loads, and exports every symbol
include/grlcu.h declares; without a GPU every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from grl_b200 import build
    return C.CDLL(build.build())


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "grlcu.h")).read()
    return sorted(set(re.findall(r"\b(grlcu_[a-z_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(lib, s), s
    from grl_b200.capi import EXPORTS
    assert sorted(EXPORTS) == syms


def test_config_struct_matches_header_layout():
    from grl_b200.capi import Config
    import _oracle
    assert C.sizeof(Config) == C.sizeof(_oracle.Config)
    # 4 int32 + 3 double ... : spot-check offsets the C compiler would give
    assert Config.control_step.offset == 8 and Config.resolution.offset == 48
    assert Config.seed.offset == Config.init_max.offset + 8


def test_no_cpu_fallback_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from grl_b200 import Batch, GrlcuError, pendulum_config
    with pytest.raises(GrlcuError) as e:
        Batch(pendulum_config(memory=1024), 4)
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)


def test_unsupported_configs_are_rejected():
    from grl_b200 import Batch, GrlcuError, pendulum_config
    with pytest.raises(GrlcuError):
        Batch(pendulum_config(memory=1024, tilings=7), 4)
    with pytest.raises(GrlcuError):
        Batch(pendulum_config(memory=0), 4)
