"""The parameters our Configurable shims request, against what the reference's own classes request (no GPU).

A drop-in behind GRL's plugin API must accept the reference's YAML: same parameter names, types (incl. the role suffix that
drives path resolution, e.g. `int.memory`, `projector.pair`), defaults, mutability and optional flags.  The reference side
comes from the reference's own sources compiled into oracle/_ref/libgrlref_agent.so (refag_requests: ConfigurableFactory::create
+ Configurable::request), ours from grl_b200/host/plugin_probe --requests.  What we deliberately do not carry is listed in
NOT_CARRIED with the reason; what we add is listed in ADDED.
"""
import ctypes as C
import os
import re
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(ROOT, "oracle", "_ref", "libgrlref_agent.so")
PROBE = os.path.join(ROOT, "grl_b200", "host", "plugin_probe")

pytestmark = pytest.mark.skipif(not (os.path.exists(LIB) and os.path.exists(PROBE)),
                                reason="needs oracle/_ref/libgrlref_agent.so (built where /root/reference exists) and the host shims")

SAME_NAME = ["dynamics/pendulum", "dynamics/cart_pole", "dynamics/cart_double_pole", "model/dynamical", "task/pendulum/swingup",
             "task/cart_pole/swingup", "task/cart_double_pole/balancing", "discretizer/uniform", "sampler/greedy",
             "sampler/epsilon_greedy", "sampler/softmax", "trace/enumerated/replacing", "trace/enumerated/accumulating",
             "mapping/policy/discrete/value/q", "mapping/policy/action", "agent/td", "agent/fixed", "experiment/online_learning"]
PAIRS = [(t, t) for t in SAME_NAME] + [
    ("environment/model/cuda_batched", "environment/modeled"),
    ("projector/tile_coding/cuda", "projector/tile_coding"),
    ("representation/linear/cuda", "representation/parameterized/linear"),
    ("predictor/sarsa/cuda", "predictor/critic/sarsa"),
    ("predictor/q/cuda", "predictor/critic/q"),
    ("predictor/expected_sarsa/cuda", "predictor/critic/expected_sarsa"),
    ("predictor/critic/td/cuda", "predictor/critic/td"),
    ("predictor/ac/action/cuda", "predictor/ac/action"),
]
# reference parameters the shims do not carry, and why
EXPORT = "importers / exporters / signals are off the hot path (DESIGN.md section 2: out of scope)"
NOT_CARRIED = {
    "experiment/online_learning": {"exporter": EXPORT, "state": EXPORT, "action": EXPORT, "reward": EXPORT, "test_action": EXPORT,
                                   "curve": EXPORT},
    "environment/model/cuda_batched": {"exporter": EXPORT, "state": EXPORT, "action": EXPORT},
    "representation/linear/cuda": {"target": "target networks (interval > 0) are not built: the representation is its own target"},
    "predictor/sarsa/cuda": {"importer": EXPORT, "exporter": EXPORT},
    "predictor/q/cuda": {"importer": EXPORT, "exporter": EXPORT},
    "predictor/expected_sarsa/cuda": {"importer": EXPORT, "exporter": EXPORT},
    "predictor/critic/td/cuda": {"importer": EXPORT, "exporter": EXPORT},
    "predictor/ac/action/cuda": {"importer": EXPORT, "exporter": EXPORT},
}
# parameters only the shims have
ADDED = {
    "experiment/online_learning": {"agents", "checkpoint_agents", "steps_per_launch", "device", "devices", "sweep_alpha",
                                   "sweep_lambda", "sweep_epsilon"},
    "sampler/greedy": {"rand_max"},              # the reference's cfg files set it (cfg/pendulum/sarsa_tc.yaml); its request() forgot it
    "sampler/epsilon_greedy": {"rand_max"},
}


def _norm(v):
    v = v.strip()
    if v.startswith("["):
        return re.sub(r"\s+", "", v)
    try:
        return repr(float(v))
    except ValueError:
        return v


@pytest.fixture(scope="module")
def ours():
    out = subprocess.run([PROBE, "--requests"] + sorted({p[0] for p in PAIRS}), capture_output=True, text=True).stdout
    assert "MISSING" not in out, out
    d = {p[0]: {} for p in PAIRS}
    for line in out.splitlines():
        if ": " in line and "|" in line:
            t, rest = line.split(": ", 1)
            f = rest.split("|")
            d[t][f[0]] = f
    return d


@pytest.fixture(scope="module")
def theirs():
    L = C.CDLL(LIB)
    buf = C.create_string_buffer(1 << 16)

    def get(t):
        n = L.refag_requests(t.encode(), buf, len(buf))
        assert n >= 0, "the reference has no type %s (%d)" % (t, n)
        return {f[0]: f for f in (ln.split("|") for ln in buf.value.decode().splitlines())}
    return get


@pytest.mark.parametrize("mine,ref", PAIRS, ids=[p[0] for p in PAIRS])
def test_shim_requests_what_the_reference_type_requests(ours, theirs, mine, ref):
    m, r = ours[mine], theirs(ref)
    skip = NOT_CARRIED.get(mine, {})
    for name, f in r.items():
        if name in skip:
            assert name not in m, "%s: %s is carried after all; drop it from NOT_CARRIED" % (mine, name)
            continue
        assert name in m, "%s lacks the reference's parameter %s" % (mine, f)
        g = m[name]
        # name | type (with role) | default | mutability | optional
        assert (g[1], _norm(g[2]), g[3], g[4]) == (f[1], _norm(f[2]), f[3], f[4]), (mine, g, f)
    assert set(m) - set(r) == ADDED.get(mine, set()), (mine, sorted(set(m) - set(r)))
