"""Pin the oracle against the reference's own golden vector for this path.

tests/golden/ref_pendulum-sarsa-tc-0.txt is the reference's seeded learning curve
(grld -s 1, tests/pendulum-sarsa-tc.yaml; bin/runtests.py byte-compares it).  Every row
depends on the whole hot path: RK4 pendulum dynamics, reward, tile hash, Q read,
epsilon-greedy draws (glibc drand48/lrand48 in construction order), SARSA(lambda) write
and replacing-trace sweep over 2000 trials.  The oracle in glibc-RNG mode reproduces all
181 rows to the 6 significant digits the file holds.
"""
import ctypes as C
import os

import _oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_pendulum-sarsa-tc-0.txt")


def test_pendulum_sarsa_tc_golden_curve(orc):
    golden = [l.split() for l in open(GOLD) if l.strip()]
    assert len(golden) == 181
    c = O.pendulum_config(memory=8388608, rng=O.RNG_GLIBC, test_interval=10)
    orc.orc_glibc_seed(1)                       # grld -s 1 -> srand48(1), deployer.cpp:70-74
    a = orc.orc_create(C.byref(c), 0)
    rec = O.Record()
    rows, trial = [], 0
    try:
        while len(rows) < len(golden):
            orc.orc_step(a, C.byref(rec))
            if rec.episode_done:
                trial += 1
                if rec.test:
                    st = C.c_uint64()
                    orc.orc_stats(a, C.byref(st), None, None, None, None)
                    # online_learning.cpp:243: tt+1-(tt+1)/(test_interval+1), ss, total_reward
                    rows.append((str(trial - trial // 11), str(st.value), "%g" % rec.episode_return))
    finally:
        orc.orc_destroy(a)
    assert rows == [tuple(g) for g in golden]


REF_CLASSES = os.path.join(os.path.dirname(__file__), "golden", "ref_classes_closed_loop.json")


def _ref_class_cases():
    import json
    with open(REF_CLASSES) as f:
        return json.load(f)["cases"]


import pytest  # noqa: E402


@pytest.mark.parametrize("case", _ref_class_cases(), ids=lambda c: c["name"])
def test_oracle_reproduces_reference_class_vectors(orc, case):
    """tests/golden/ref_classes_closed_loop.json was written by tools/make_ref_agent_golden.py from a closed loop of the
    reference's OWN task / model / agent classes (compiled from /root/reference against a stand-in Eigen).
    The oracle in glibc-RNG mode, from the same srand48 seed, must arrive at the same step and trial counts, the same
    environment state bits and the same weight tables (SHA-256 of their bytes) -- on any machine, with no reference around."""
    import hashlib
    import numpy as np
    cfg = getattr(O, case["config"])(**case["overrides"])
    cfg.rng = O.RNG_GLIBC
    orc.orc_glibc_seed(case["seed"])
    a = orc.orc_create(C.byref(cfg), 0)
    try:
        for _ in range(case["steps"]):
            orc.orc_step(a, None)
        n_steps, n_eps = C.c_uint64(), C.c_uint64()
        orc.orc_stats(a, C.byref(n_steps), C.byref(n_eps), None, None, None)
        assert (n_steps.value, n_eps.value) == (case["learning_steps"], case["trials"])
        state = (C.c_double * O.MAX_STATE)()
        orc.orc_get_state.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_void_p, C.c_void_p, C.c_void_p]
        orc.orc_get_state(a, state, None, None, None)
        S = orc.orc_state_dims(cfg.dynamics)
        assert ["%016x" % b for b in np.array(list(state)[:S]).view(np.uint64)] == case["state_bits"]
        tab = np.ctypeslib.as_array(orc.orc_table(a), shape=(cfg.memory,))
        assert hashlib.sha256(tab.tobytes()).hexdigest() == case["table_sha256"]
        if "actor_table_sha256" in case:
            act = np.ctypeslib.as_array(orc.orc_actor_table(a), shape=(cfg.memory,))
            assert hashlib.sha256(act.tobytes()).hexdigest() == case["actor_table_sha256"]
    finally:
        orc.orc_destroy(a)


GRLD_CURVES = os.path.join(os.path.dirname(__file__), "golden", "grld_ref_curves.json")


def _grld_curve_cases():
    import json
    with open(GRLD_CURVES) as f:
        return json.load(f)["cases"]


@pytest.mark.parametrize("case", _grld_curve_cases(), ids=lambda c: c["name"])
def test_oracle_reproduces_reference_deployer_curves(orc, case):
    """tests/golden/grld_ref_curves.json holds the rows the reference's OWN deployer wrote (oracle/_ref/grld_ref: grld built from
    the reference's translation units; generator: tools/make_grld_ref_golden.py).  The oracle in glibc-RNG mode
    from the same seed must write the same rows: trial, learning steps so far, return to the 3 decimals printed, episode time."""
    from test_oracle_vs_grld_ref import oracle_rows
    cfg = getattr(O, case["config"])(**case["overrides"])
    want = [(int(r[0]), int(r[1]), float(r[2]), float(r[3])) for r in case["rows"]]
    got = oracle_rows(orc, cfg, case["seed"], case["trials"])
    assert len(got) == len(want) >= 10
    for g, w in zip(got, want):
        assert (g[0], g[1], g[3]) == (w[0], w[1], w[3]) and abs(g[2] - w[2]) <= 0.00051, (g, w)
