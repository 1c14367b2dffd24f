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


def test_config_struct_matches_the_c_compiler(tmp_path):
    """sizeof / offsetof of every grlcu_config and orc_config field as gcc lays them out, against the ctypes mirrors
    (a field added to one header and not to a mirror shifts everything after it)."""
    import subprocess
    from grl_b200.capi import Config
    import _oracle
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    names = [f[0] for f in Config._fields_]
    cname = lambda n: "lambda" if n == "lambda_" else n
    for header, struct, mirror in (("include/grlcu.h", "grlcu_config", Config), ("oracle/grl_oracle.h", "orc_config", _oracle.Config)):
        assert [f[0] for f in mirror._fields_] == names
        src = tmp_path / (struct + ".c")
        body = "".join('  printf("%%s %%zu\\n", "%s", offsetof(%s, %s));\n' % (n, struct, cname(n)) for n in names)
        src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "%s"\nint main(void)\n{\n  printf("sizeof %%zu\\n", sizeof(%s));\n%s  return 0;\n}\n'
                       % (os.path.join(root, header), struct, body))
        exe = tmp_path / struct
        subprocess.run(["gcc", "-o", str(exe), str(src)], check=True)
        out = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
        assert int(out["sizeof"]) == C.sizeof(mirror), struct
        for n in names:
            assert int(out[n]) == getattr(mirror, n).offset, (struct, n)


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
