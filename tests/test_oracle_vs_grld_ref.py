"""The oracle's learning curves against the reference's own deployer (no GPU).

oracle/_ref/grld_ref (oracle/Makefile, target `ref`) is `grld` built from the reference's own translation units --
deployer.cpp, grl.cpp, configurable.cpp (YAML configurator), experiments/online_learning.cpp, environments/modeled.cpp,
the dynamics / tasks, agents, policies, predictors, samplers, representation, projector, traces -- with its own headers,
its own flags and the yaml-cpp it vendors; the one stand-in is Eigen (oracle/ref_shim_eigen/).  Here it runs YAML files
of the same shape as the reference's cfg/*/ (live type names) with `-s seed`, exactly as bin/runtests.py runs `grld -s 1`,
and the rows it writes -- test trial, learning steps so far, episode return (3 decimals), episode time -- are compared
with the rows of the oracle free-running in glibc RNG mode from the same seed.  This covers what no other pin does: the call
order inside ModeledEnvironment::step and the episode loop of OnlineLearningExperiment::run as the reference executes them.
"""
import ctypes as C
import os
import subprocess

import pytest

import _oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
GRLD = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "grld_ref")

pytestmark = pytest.mark.skipif(not os.path.exists(GRLD),
                                reason="oracle/_ref/grld_ref is built where /root/reference exists (make -C oracle ref)")

yaml_for = O.grld_yaml


def oracle_rows(orc, c, seed, trials):
    """Rows as online_learning.cpp:238-262 writes them: one per test trial, or one per trial with test_interval < 0."""
    c.rng = O.RNG_GLIBC
    orc.orc_glibc_seed(seed)
    a = orc.orc_create(C.byref(c), 0)
    rec = O.Record()
    rows, tt, ss, ep_steps = [], 0, 0, 0
    try:
        while tt < trials:
            ss += orc.orc_step(a, C.byref(rec))
            ep_steps += 1
            if rec.episode_done:
                ti = c.test_interval
                if ti >= 0:
                    if rec.test:
                        rows.append((tt + 1 - (tt + 1) // (ti + 1), ss, rec.episode_return, ep_steps))
                else:
                    rows.append((tt, ss, rec.episode_return, ep_steps))
                tt += 1
                ep_steps = 0
    finally:
        orc.orc_destroy(a)
    return rows


def grld_rows(c, seed, trials, tmp_path):
    (tmp_path / "cfg.yaml").write_text(yaml_for(c, trials, "curve"))
    r = subprocess.run([GRLD, "-s", str(seed), "cfg.yaml"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and os.path.exists(tmp_path / "curve-0.txt"), r.stdout[-2000:] + r.stderr[-2000:]
    rows = []
    for line in open(tmp_path / "curve-0.txt"):
        f = line.split()
        if f:
            rows.append((int(f[0]), int(f[1]), float(f[2]), float(f[3])))
    return rows


CASES = [
    ("pendulum_sarsa_shipped", "pendulum_config", dict(memory=8388608), 1, 330),
    ("pendulum_sarsa_randomized", "pendulum_config", dict(memory=65536, randomization=1.0), 2, 440),
    ("pendulum_sarsa_every_trial", "pendulum_config", dict(memory=65536, test_interval=-1), 3, 150),
    ("cart_pole_sarsa", "cart_pole_config", dict(memory=262144, randomization=1.0), 4, 330),
    ("cart_double_pole_sarsa", "cart_double_pole_config", dict(memory=131072), 5, 550),
    ("cart_double_pole_ac", "cart_double_pole_ac_config", dict(memory=131072), 6, 660),
    ("pendulum_ac", "pendulum_ac_config", dict(memory=65536), 7, 220),
    ("cart_pole_ac", "cart_pole_ac_config", dict(memory=262144), 8, 110),
    ("pendulum_q_learning", "pendulum_config", dict(memory=65536, algo=O.ALGO_Q, randomization=1.0), 9, 330),
    ("pendulum_expected_sarsa", "pendulum_config", dict(memory=65536, algo=O.ALGO_EXPECTED_SARSA, target_epsilon=0.05,
                                                        randomization=1.0), 13, 220),
    ("pendulum_softmax", "pendulum_config", dict(memory=65536, sampler=O.SAMPLER_SOFTMAX, softmax_tau=0.5), 10, 220),
    ("pendulum_accumulating_trace", "pendulum_config", dict(memory=65536, trace=O.TRACE_ACCUMULATING), 11, 220),
    ("pendulum_safe_2", "pendulum_config", dict(memory=65536, safe=2), 12, 220),
]


@pytest.mark.parametrize("name,fn,kw,seed,trials", CASES, ids=[c[0] for c in CASES])
def test_learning_curve_equals_the_reference_deployer(orc, tmp_path, name, fn, kw, seed, trials):
    c = getattr(O, fn)(**kw)
    want = oracle_rows(orc, c, seed, trials)
    got = grld_rows(c, seed, trials, tmp_path)
    assert len(got) == len(want) and len(got) >= 10
    for g, w in zip(got, want):
        assert g[0] == w[0] and g[1] == w[1], (g, w)                 # trial, learning steps so far
        assert abs(g[2] - w[2]) <= 0.00051, (g, w)                   # episode return, printed with 3 decimals
        assert g[3] == w[3], (g, w)                                  # episode time = steps (discrete_time: 1)


def test_checkpoint_written_by_the_reference_deployer(orc, tmp_path):
    """`save_every: run` of the reference's own deployer (representation.h:201-263, online_learning.cpp:281-302): the file
    name the host shims must use to be interchangeable with the reference -- note the '_' that the root of the object path
    contributes -- its size, and its content: the oracle's Q table after the same seeded run, bit for bit."""
    import numpy as np
    c = O.pendulum_config(memory=4096, test_interval=-1, randomization=1.0)
    trials, seed = 25, 3
    (tmp_path / "cfg.yaml").write_text(O.grld_yaml(c, trials, "ck").replace("save_every: never", "save_every: run"))
    r = subprocess.run([GRLD, "-s", str(seed), "cfg.yaml"], cwd=tmp_path, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout[-1500:]
    name = tmp_path / "ck-run0-_experiment_agent_policy_representation.dat"
    assert name.exists(), sorted(p.name for p in tmp_path.iterdir())
    table = np.fromfile(str(name), dtype=np.float64)
    assert table.size == c.memory
    c.rng = O.RNG_GLIBC
    orc.orc_glibc_seed(seed)
    a = orc.orc_create(C.byref(c), 0)
    rec = O.Record()
    done = 0
    while done < trials:
        orc.orc_step(a, C.byref(rec))
        done += rec.episode_done
    want = np.ctypeslib.as_array(orc.orc_table(a), shape=(c.memory,)).copy()
    orc.orc_destroy(a)
    assert np.array_equal(table.view(np.uint64), want.view(np.uint64))
    # the name the C++ shim composes (grl_b200/host/grl_cuda_objects.cpp: file + path with '/' -> '_' + ".dat", path rooted at '/')
    src = open(os.path.join(os.path.dirname(HERE), "grl_b200", "host", "grl_cuda_objects.cpp")).read()
    assert 'cfg_path = "/" + cfg_path' in src
