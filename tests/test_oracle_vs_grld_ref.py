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

DYN = {O.DYN_PENDULUM: ("dynamics/pendulum", "task/pendulum/swingup"),
       O.DYN_CART_POLE: ("dynamics/cart_pole", "task/cart_pole/swingup"),
       O.DYN_CART_DOUBLE_POLE: ("dynamics/cart_double_pole", "task/cart_double_pole/balancing")}


def _vec(a, n):
    return "[" + ", ".join(repr(float(a[i])) for i in range(n)) + "]"


def _environment(c):
    dyn, task = DYN[c.dynamics]
    extra = ""
    if c.dynamics == O.DYN_CART_POLE:
        extra = ", end_stop_penalty: %d" % c.end_stop_penalty
    # (task/cart_pole/swingup declares `randomization` as an integer, task/pendulum/swingup as a real number)
    rand = "" if c.dynamics == O.DYN_CART_DOUBLE_POLE and c.randomization == 0 else ", randomization: %s" % (
        "%d" % int(c.randomization) if c.dynamics == O.DYN_CART_POLE else repr(c.randomization))
    return """  environment:
    type: environment/modeled
    model:
      type: model/dynamical
      control_step: %r
      integration_steps: %d
      dynamics: {type: %s}
    task: {type: %s, timeout: %r%s%s}
""" % (c.control_step, c.integration_steps, dyn, task, c.timeout, rand, extra)


def _trace(c):
    return {O.TRACE_REPLACING: "{type: trace/enumerated/replacing}", O.TRACE_ACCUMULATING: "{type: trace/enumerated/accumulating}",
            O.TRACE_NONE: "0"}[c.trace]


def yaml_for(c, trials, output):
    """A configuration file of the shape of the reference's cfg/pendulum/sarsa_tc.yaml / cfg/cart_double_pole/ac.yaml
    (visualisation left out), generated from the oracle's configuration so that both sides always run the same thing."""
    head = """experiment:
  type: experiment/online_learning
  runs: 1
  trials: %d
  steps: 0
  rate: 0
  test_interval: %d
  output: %s
%s""" % (trials, c.test_interval, output, _environment(c))
    if c.algo == O.ALGO_AC:
        n = c.obs_dims
        return head + """  agent:
    type: agent/td
    policy:
      type: mapping/policy/action
      sigma: [%r]
      output_min: experiment/environment/task/action_min
      output_max: experiment/environment/task/action_max
      projector:
        type: projector/tile_coding
        tilings: %d
        memory: %d
        safe: 0
        resolution: %s
        wrapping: %s
      representation:
        type: representation/parameterized/linear
        init_min: [%r]
        init_max: [%r]
        memory: experiment/agent/policy/projector/memory
        outputs: experiment/environment/task/action_dims
        output_min: experiment/environment/task/action_min
        output_max: experiment/environment/task/action_max
    predictor:
      type: predictor/ac/action
      alpha: %r
      update_method: proportional
      step_limit: []
      critic:
        type: predictor/critic/td
        alpha: %r
        gamma: %r
        lambda: %r
        projector: {type: projector/tile_coding, tilings: %d, memory: %d, safe: 0, resolution: %s, wrapping: %s}
        representation:
          type: representation/parameterized/linear
          init_min: [%r]
          init_max: [%r]
          memory: experiment/agent/predictor/critic/projector/memory
          outputs: 1
          output_min: []
          output_max: []
        trace: %s
      projector: experiment/agent/policy/projector
      representation: experiment/agent/policy/representation
  test_agent:
    type: agent/fixed
    policy:
      type: mapping/policy/action
      sigma: []
      output_min: experiment/environment/task/action_min
      output_max: experiment/environment/task/action_max
      projector: experiment/agent/policy/projector
      representation: experiment/agent/policy/representation
  save_every: never
""" % (c.sigma, c.tilings, c.memory, _vec(c.actor_resolution, n), _vec(c.actor_wrapping, n), c.init_min, c.init_max,
       c.actor_alpha, c.alpha, c.gamma, c.lambda_, c.tilings, c.memory, _vec(c.resolution, n), _vec(c.wrapping, n),
       c.init_min, c.init_max, _trace(c))
    n = c.obs_dims + 1
    sampler = ("{type: sampler/softmax, tau: %r}" % c.softmax_tau if c.sampler == O.SAMPLER_SOFTMAX
               else "{type: sampler/epsilon_greedy, epsilon: %r}" % c.epsilon)
    predictor = {O.ALGO_SARSA: "predictor/critic/sarsa", O.ALGO_Q: "predictor/critic/q"}[c.algo]
    extra = "      discretizer: experiment/agent/policy/discretizer\n" if c.algo == O.ALGO_Q else ""
    return head + """  agent:
    type: agent/td
    policy:
      type: mapping/policy/discrete/value/q
      discretizer:
        type: discretizer/uniform
        min: experiment/environment/task/action_min
        max: experiment/environment/task/action_max
        steps: [%d]
      projector:
        type: projector/tile_coding
        tilings: %d
        memory: %d
        safe: %d
        resolution: %s
        wrapping: %s
      representation:
        type: representation/parameterized/linear
        init_min: [%r]
        init_max: [%r]
        memory: experiment/agent/policy/projector/memory
        outputs: 1
        output_min: []
        output_max: []
      sampler: %s
    predictor:
      type: %s
      alpha: %r
      gamma: %r
      lambda: %r
%s      projector: experiment/agent/policy/projector
      representation: experiment/agent/policy/representation
      trace: %s
  test_agent:
    type: agent/fixed
    policy:
      type: mapping/policy/discrete/value/q
      discretizer: experiment/agent/policy/discretizer
      projector: experiment/agent/policy/projector
      representation: experiment/agent/policy/representation
      sampler: {type: sampler/greedy}
  save_every: never
""" % (c.n_actions, c.tilings, c.memory, c.safe, _vec(c.resolution, n), _vec(c.wrapping, n), c.init_min, c.init_max,
       sampler, predictor, c.alpha, c.gamma, c.lambda_, extra, _trace(c))


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
