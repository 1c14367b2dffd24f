"""The Configurable shims (grl_b200/host): YAML -> object tree -> grlcu_config, error behaviour, and (GPU) that
grld_b200 driving the C ABI from C++ gives the same per-agent results as the ctypes harness."""
import json
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "grl_b200", "host")
GRLD = os.path.join(HOST, "grld_b200")
REF_CFG = "/root/reference/cfg"


@pytest.fixture(scope="module")
def grld():
    from grl_b200 import build
    build.build()
    subprocess.run(["make", "-s"], cwd=HOST, check=True)
    return GRLD


def check(grld, path, *extra):
    out = subprocess.run([grld, "-c", *extra, path], capture_output=True, text=True)
    return out.returncode, out.stdout, out.stderr


def test_shipped_configs_resolve_to_the_harness_configs(grld):
    import grl_b200
    for yaml, mk, agents in (("cfg/pendulum/sarsa_tc_cuda.yaml", grl_b200.pendulum_config, 65536),
                             ("cfg/cart_pole/sarsa_tc_cuda.yaml", grl_b200.cart_pole_config, 32768),
                             ("cfg/cart_double_pole/sarsa_tc_cuda.yaml", grl_b200.cart_double_pole_config, 65536)):
        rc, out, err = check(grld, os.path.join(ROOT, yaml))
        assert rc == 0, err
        got = json.loads(out)
        want = mk(memory=got["memory"], randomization=got["randomization"], test_interval=got["test_interval"])
        assert got["randomization"] == (0.0 if "double" in yaml else 1.0)
        assert got["agents"] == agents
        for k in ("dynamics", "integration_steps", "control_step", "timeout", "randomization", "tilings", "memory",
                  "obs_dims", "n_actions", "alpha", "gamma", "epsilon", "trace", "init_min", "init_max", "seed", "algo"):
            assert got[k] == getattr(want, k), k
        assert got["lambda"] == want.lambda_
        dims = got["obs_dims"] + 1
        assert got["resolution"] == list(want.resolution)[:dims]
        assert got["wrapping"] == list(want.wrapping)[:dims]
        assert got["actions"] == list(want.actions)[:3]


def test_actor_critic_config_resolves_to_the_harness_config(grld):
    import grl_b200
    rc, out, err = check(grld, os.path.join(ROOT, "cfg/cart_double_pole/ac_tc_cuda.yaml"))
    assert rc == 0, err
    got = json.loads(out)
    want = grl_b200.cart_double_pole_ac_config(memory=got["memory"], test_interval=10)
    assert got["agents"] == 4096 and got["algo"] == grl_b200.capi.ALGO_AC
    for k in ("dynamics", "integration_steps", "control_step", "timeout", "randomization", "tilings", "memory", "obs_dims",
              "n_actions", "alpha", "gamma", "epsilon", "trace", "init_min", "init_max", "seed", "algo",
              "sigma", "actor_alpha", "action_min", "action_max"):
        assert got[k] == getattr(want, k), k
    assert got["lambda"] == want.lambda_
    assert got["resolution"][:6] == list(want.resolution)[:6] == got["actor_resolution"] == list(want.actor_resolution)[:6]
    assert got["actor_wrapping"] == [0.0] * 6


@pytest.mark.parametrize("yaml,mk", [("cfg/pendulum/ac_tc_cuda.yaml", "pendulum_ac_config"),
                                     ("cfg/cart_pole/ac_tc_cuda.yaml", "cart_pole_ac_config")])
def test_sibling_actor_critic_configs_resolve_to_the_harness_configs(grld, yaml, mk):
    """The reference's other tile-coded actor-critic configs (cfg/pendulum/ac_tc.yaml, cfg/cart_pole/ac_tc.yaml)."""
    import grl_b200
    rc, out, err = check(grld, os.path.join(ROOT, yaml))
    assert rc == 0, err
    got = json.loads(out)
    want = getattr(grl_b200, mk)(memory=got["memory"], test_interval=10, randomization=got["randomization"])
    n = want.obs_dims
    for k in ("dynamics", "integration_steps", "control_step", "timeout", "tilings", "memory", "obs_dims", "n_actions",
              "alpha", "gamma", "epsilon", "trace", "init_min", "init_max", "seed", "algo", "sigma", "actor_alpha",
              "action_min", "action_max", "end_stop_penalty"):
        assert got[k] == getattr(want, k), k
    assert got["resolution"][:n] == list(want.resolution)[:n] == got["actor_resolution"]
    assert got["wrapping"][:n] == list(want.wrapping)[:n] == got["actor_wrapping"]
    assert got["end_stop_penalty"] == (0 if "cart_pole" in yaml else 1)


def swap_types(text):
    """The only edit a GRL user makes to an existing YAML: the four type names (+ a table size that fits)."""
    text = text.replace("type: environment/modeled", "type: environment/model/cuda_batched")
    text = re.sub(r"type: projector/tile_coding\b", "type: projector/tile_coding/cuda", text)
    text = text.replace("type: representation/parameterized/linear", "type: representation/linear/cuda")
    text = text.replace("type: predictor/critic/sarsa", "type: predictor/sarsa/cuda")
    text = text.replace("type: predictor/critic/q", "type: predictor/q/cuda")
    text = text.replace("type: predictor/critic/expected_sarsa", "type: predictor/expected_sarsa/cuda")
    # actor-critic (cfg/cart_double_pole/ac.yaml): the two predictors; the policy's type name in that file is stale
    # in the reference itself (TYPEINFO is mapping/policy/action, policies/action.h:44)
    text = text.replace("type: predictor/ac/action", "type: predictor/ac/action/cuda")
    text = text.replace("type: predictor/critic/td", "type: predictor/critic/td/cuda")
    text = text.replace("type: policy/parameterized/action", "type: mapping/policy/action")
    return text


@pytest.mark.skipif(not os.path.isdir(REF_CFG), reason="reference tree not present (GPU box)")
@pytest.mark.parametrize("name,dyn,mem", [("pendulum/sarsa_tc.yaml", 0, 8388608), ("cart_pole/sarsa.yaml", 1, 67108864),
                                          ("cart_double_pole/sarsa.yaml", 2, 67108864)])
def test_reference_yaml_loads_with_only_the_type_names_swapped(grld, tmp_path, name, dyn, mem):
    src = open(os.path.join(REF_CFG, name)).read()
    dst = tmp_path / "swapped.yaml"
    dst.write_text(swap_types(src))
    rc, out, err = check(grld, str(dst), "-s", "7")
    assert rc == 0, err
    got = json.loads(out)
    assert got["dynamics"] == dyn and got["memory"] == mem and got["seed"] == 7
    assert got["test_interval"] == 10 and got["trace"] == 1 and got["agents"] == 1
    assert got["actions"] == [[-3.0, 0.0, 3.0], [-15.0, 0.0, 15.0], [-20.0, 0.0, 20.0]][dyn]   # discretizer via task/action_min,max


@pytest.mark.skipif(not os.path.isdir(REF_CFG), reason="reference tree not present (GPU box)")
def test_reference_actor_critic_yaml_loads_with_the_type_names_swapped(grld, tmp_path):
    src = open(os.path.join(REF_CFG, "cart_double_pole/ac.yaml")).read()
    dst = tmp_path / "swapped.yaml"
    dst.write_text(swap_types(src))
    rc, out, err = check(grld, str(dst))
    assert rc == 0, err
    got = json.loads(out)
    assert got["dynamics"] == 2 and got["algo"] == 1 and got["memory"] == 67108864 and got["obs_dims"] == 6
    assert (got["sigma"], got["actor_alpha"], got["alpha"], got["lambda"]) == (5.0, 0.01, 0.2, 0.65)
    assert (got["action_min"], got["action_max"], got["timeout"], got["test_interval"]) == (-20.0, 20.0, 4.99, 10)


@pytest.mark.skipif(not os.path.isdir(REF_CFG), reason="reference tree not present (GPU box)")
@pytest.mark.parametrize("name,dyn,obs,sigma,amax,esp", [("pendulum/ac_tc.yaml", 0, 2, 1.0, 3.0, 1),
                                                         ("cart_pole/ac_tc.yaml", 1, 4, 5.0, 15.0, 0)])
def test_reference_sibling_actor_critic_yamls_load_with_the_type_names_swapped(grld, tmp_path, name, dyn, obs, sigma, amax, esp):
    """pendulum/ac_tc.yaml reaches the critic's memory through a parameter that is itself a reference
    (critic/projector names the policy's projector); cart_pole/ac_tc.yaml switches the end-stop penalty off."""
    dst = tmp_path / "swapped.yaml"
    dst.write_text(swap_types(open(os.path.join(REF_CFG, name)).read()))
    rc, out, err = check(grld, str(dst))
    assert rc == 0, err
    got = json.loads(out)
    assert (got["dynamics"], got["algo"], got["memory"], got["obs_dims"]) == (dyn, 1, 8388608, obs)
    assert (got["sigma"], got["actor_alpha"], got["alpha"], got["lambda"]) == (sigma, 0.01, 0.2, 0.65)
    assert (got["action_min"], got["action_max"], got["end_stop_penalty"]) == (-amax, amax, esp)
    assert got["actor_resolution"] == got["resolution"][:obs] and got["actor_wrapping"] == got["wrapping"][:obs]


def test_actor_critic_error_behaviour(grld, tmp_path):
    base = open(os.path.join(ROOT, "cfg/cart_double_pole/ac_tc_cuda.yaml")).read()
    cases = [
        # the critic's projector must be the actor's (one index set addresses both tables)
        ("          resolution: experiment/agent/policy/projector/resolution", "          resolution: [ 2.5, 0.1, 0.1, 2.5, 1, 1 ]", "critic's projector"),
        ("      update_method: proportional", "      update_method: cacla", "predictor/ac/action/cuda:update_method"),
        ("      sigma: [  ]", "      sigma: [ 1 ]", "must not explore"),
        ("        type: predictor/critic/td/cuda", "        type: predictor/sarsa/cuda", "critic"),
        ("        output_min: experiment/environment/task/action_min\n", "        output_min: [ -1 ]\n", "output_min/max"),
    ]
    for old, new, msg in cases:
        assert old in base, old
        p = tmp_path / "bad.yaml"
        p.write_text(base.replace(old, new, 1))
        rc, out, err = check(grld, str(p))
        assert rc == 1 and msg in err, (old, err)


@pytest.mark.skipif(not os.path.isdir(REF_CFG), reason="reference tree not present (GPU box)")
def test_reference_q_learning_yaml_and_expected_sarsa(grld, tmp_path):
    """cfg/pendulum/q_tc.yaml with the types swapped -> GRLCU_ALGO_Q; the same tree with an expected-SARSA predictor
    whose target policy is the greedy test policy / the epsilon-greedy learning policy."""
    src = swap_types(open(os.path.join(REF_CFG, "pendulum/q_tc.yaml")).read())
    dst = tmp_path / "q.yaml"
    dst.write_text(src)
    rc, out, err = check(grld, str(dst))
    assert rc == 0, err
    got = json.loads(out)
    assert (got["algo"], got["dynamics"], got["memory"], got["trace"], got["epsilon"]) == (2, 0, 8388608, 1, 0.05)
    es = src.replace("type: predictor/q/cuda", "type: predictor/expected_sarsa/cuda") \
            .replace("      discretizer: experiment/agent/policy/discretizer\n      projector: experiment/agent/policy/projector",
                     "      policy:\n        type: mapping/policy/discrete/value/q\n"
                     "        discretizer: experiment/agent/policy/discretizer\n        projector: experiment/agent/policy/projector\n"
                     "        representation: experiment/agent/policy/representation\n        sampler:\n          type: sampler/greedy\n"
                     "      projector: experiment/agent/policy/projector", 1)
    assert "type: sampler/greedy\n      projector" in es
    dst.write_text(es)
    rc, out, err = check(grld, str(dst))
    assert rc == 0, err
    assert (json.loads(out)["algo"], json.loads(out)["target_epsilon"]) == (3, -1.0)
    a = es.index("      policy:\n        type: mapping/policy/discrete/value/q\n        discretizer: experiment/agent")
    b = es.index("      projector: experiment/agent/policy/projector", a)
    dst.write_text(es[:a] + "      policy: experiment/agent/policy\n" + es[b:])
    rc, out, err = check(grld, str(dst))
    assert rc == 0, err
    assert (json.loads(out)["algo"], json.loads(out)["target_epsilon"]) == (3, 0.05)


def test_softmax_sampler_from_yaml(grld, tmp_path):
    """sampler/softmax (softmax.cpp:32-35: tau in [1e-6, 100]) in place of sampler/epsilon_greedy."""
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml")).read()
    old = "        type: sampler/epsilon_greedy\n        epsilon: 0.05"
    assert old in base
    p = tmp_path / "softmax.yaml"
    p.write_text(base.replace(old, "        type: sampler/softmax\n        tau: 0.25"))
    rc, out, err = check(grld, str(p))
    assert rc == 0, err
    got = json.loads(out)
    assert (got["sampler"], got["softmax_tau"], got["algo"]) == (1, 0.25, 0)
    p.write_text(base.replace(old, "        type: sampler/softmax\n        tau: 0"))
    rc, out, err = check(grld, str(p))
    assert rc == 1 and "tau" in err
    rc, out, err = check(grld, os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml"))
    assert json.loads(out)["sampler"] == 0


def test_error_behaviour(grld, tmp_path):
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml")).read()
    cases = {
        "type: projector/tile_coding/cuda": ("type: projector/tile_coding", "Unknown object type"),
        "wrapping: [6.283, 0, 0]": ("wrapping: [6.0, 0, 0]", "projector/tile_coding:wrapping"),
        "memory: experiment/agent/policy/projector/memory": ("memory: 1024", "representation/linear/cuda:memory"),
        "type: trace/enumerated/replacing": ("type: enumerated/nonesuch", "Unknown object type"),
        "tilings: 16": ("tilings: 8", "projector/tile_coding/cuda:tilings"),
        "projector: experiment/agent/policy/projector": ("projector: experiment/agent/policy/nothing", "does not name an object"),
    }
    for old, (new, msg) in cases.items():
        assert old in base
        p = tmp_path / "bad.yaml"
        p.write_text(base.replace(old, new, 1))
        rc, out, err = check(grld, str(p))
        assert rc == 1 and msg in err, (old, err)
    # unique type suffixes are accepted (configurable.cpp:106-125)
    p = tmp_path / "suffix.yaml"
    p.write_text(base.replace("type: trace/enumerated/replacing", "type: enumerated/accumulating"))
    rc, out, err = check(grld, str(p))
    assert rc == 0 and json.loads(out)["trace"] == 2


def test_sweep_grid_parameters(grld, tmp_path):
    """sweep_alpha / sweep_lambda / sweep_epsilon: the trace ring is sized for the grid's largest lambda."""
    rc, out, err = check(grld, os.path.join(ROOT, "cfg/pendulum/sarsa_tc_sweep_cuda.yaml"))
    assert rc == 0, err
    got = json.loads(out)
    assert (got["agents"], got["memory"], got["lambda"], got["alpha"], got["epsilon"]) == (131072, 32768, 0.9, 0.2, 0.05)
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_sweep_cuda.yaml")).read()
    for old, new, msg in (("sweep_lambda: [0.0, 0.2,", "sweep_lambda: [1.5, 0.2,", "experiment/online_learning:sweep_lambda"),
                          ("sweep_alpha: [0.05,", "sweep_alpha: [0.0,", "experiment/online_learning:sweep_alpha"),
                          ("sweep_epsilon: [0.01,", "sweep_epsilon: [-0.01,", "experiment/online_learning:sweep_epsilon")):
        assert old in base
        p = tmp_path / "bad.yaml"
        p.write_text(base.replace(old, new, 1))
        rc, out, err = check(grld, str(p))
        assert rc == 1 and msg in err, err


def test_run_without_gpu_fails_loudly(grld):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    out = subprocess.run([grld, os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml")], capture_output=True, text=True)
    assert out.returncode == 1 and "no CPU fallback" in out.stderr


@pytest.mark.gpu
def test_grld_matches_ctypes_harness(grld, tmp_path):
    import grl_b200
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml")).read()
    y = base.replace("agents: 65536", "agents: 96").replace("steps: 10000", "steps: 700") \
            .replace("memory: 65536", "memory: 4096").replace("test_interval: -1", "test_interval: 2") \
            .replace("  runs: 1", "  runs: 1\n  output: %s\n  save_every: run" % (tmp_path / "out"))
    y = y.replace("      trace:\n        type: trace/enumerated/replacing",
                  "      trace:\n        type: trace/enumerated/replacing\n  test_agent:\n    type: agent/fixed\n    policy:\n"
                  "      type: mapping/policy/discrete/value/q\n      discretizer: experiment/agent/policy/discretizer\n"
                  "      projector: experiment/agent/policy/projector\n      representation: experiment/agent/policy/representation\n"
                  "      sampler:\n        type: sampler/greedy")
    p = tmp_path / "small.yaml"
    p.write_text(y)
    out = subprocess.run([grld, "-s", "3", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    rows = [l.split() for l in open(str(tmp_path / "out") + "-0.txt") if l.startswith("# ") and l[2].isdigit()]
    assert len(rows) == 96
    cfg = grl_b200.pendulum_config(memory=4096, randomization=1.0, test_interval=2, seed=3)
    b = grl_b200.Batch(cfg, 96)
    launches = 0
    while True:
        b.run(100)
        launches += 1
        r = b.results()
        if r["steps"].min() >= 700:
            break
    for a, row in enumerate(rows):
        assert int(row[2]) == r["steps"][a] and int(row[3]) == r["episodes"][a]
        assert float(row[4]) == r["last_return"][a] and float(row[5]) == r["last_test_return"][a]
    # checkpoint in the reference's format (representation.h:201-263): raw doubles, path-derived name
    ck = np.fromfile(str(tmp_path / "out") + "-run0-_experiment_agent_policy_representation.dat", dtype=np.float64)
    assert np.array_equal(ck, b.get_table(0))
    b.close()


@pytest.mark.gpu
@pytest.mark.parametrize("yaml,mk,amax,kw", [
    ("cfg/cart_double_pole/ac_tc_cuda.yaml", "cart_double_pole_ac_config", 20.0, {}),
    ("cfg/pendulum/ac_tc_cuda.yaml", "pendulum_ac_config", 3.0, dict(randomization=1.0)),
    ("cfg/cart_pole/ac_tc_cuda.yaml", "cart_pole_ac_config", 15.0, {}),
])
def test_grld_actor_critic_matches_ctypes_harness(grld, tmp_path, yaml, mk, amax, kw):
    """cfg 4 (and its pendulum / cart-pole siblings) from YAML: same per-agent results as the ctypes harness, and both
    tables checkpointed in the reference's format."""
    import grl_b200
    base = open(os.path.join(ROOT, yaml)).read()
    y = base.replace("agents: 4096", "agents: 64").replace("steps: 5000", "steps: 600") \
            .replace("steps_per_launch: 500", "steps_per_launch: 100") \
            .replace("test_interval: 10", "test_interval: 3") \
            .replace("  save_every: never", "  output: %s\n  save_every: run" % (tmp_path / "ac"))
    y = re.sub(r"        memory: \d+\n", "        memory: 8192\n", y)
    p = tmp_path / "small.yaml"
    p.write_text(y)
    out = subprocess.run([grld, "-s", "5", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    rows = [l.split() for l in open(str(tmp_path / "ac") + "-0.txt") if l.startswith("# ") and l[2].isdigit()]
    assert len(rows) == 64
    cfg = getattr(grl_b200, mk)(memory=8192, test_interval=3, seed=5, **kw)
    b = grl_b200.Batch(cfg, 64)
    try:
        while True:
            b.run(100)
            r = b.results()
            if r["steps"].min() >= 600:
                break
        for a, row in enumerate(rows):
            assert int(row[2]) == r["steps"][a] and int(row[3]) == r["episodes"][a]
            assert float(row[4]) == r["last_return"][a] and float(row[5]) == r["last_test_return"][a]
        stem = str(tmp_path / "ac") + "-run0-_experiment_agent_"
        actor = np.fromfile(stem + "policy_representation.dat", dtype=np.float64)
        critic = np.fromfile(stem + "predictor_critic_representation.dat", dtype=np.float64)
        assert np.array_equal(actor, b.get_actor_table(0)) and np.array_equal(critic, b.get_table(0))
        assert np.abs(actor).max() <= amax and not np.array_equal(actor, critic)
        a7 = np.fromfile(stem + "policy_representation-agent7.dat", dtype=np.float64)
        assert np.array_equal(a7, b.get_actor_table(7))
    finally:
        b.close()


@pytest.mark.gpu
def test_grld_sweep_matches_ctypes_harness(grld, tmp_path):
    """The grid on the agents axis: agent g runs cell g mod cells (alpha fastest); per-cell curves are written."""
    import grl_b200
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_sweep_cuda.yaml")).read()
    y = base.replace("agents: 131072", "agents: 160").replace("steps: 10000", "steps: 400") \
            .replace("memory: 32768", "memory: 4096") \
            .replace("sweep_alpha: [0.05, 0.1, 0.2, 0.4]", "sweep_alpha: [0.1, 0.4]") \
            .replace("sweep_lambda: [0.0, 0.2, 0.4, 0.5, 0.65, 0.75, 0.85, 0.9]", "sweep_lambda: [0.0, 0.65, 0.9]") \
            .replace("sweep_epsilon: [0.01, 0.02, 0.05, 0.1]", "sweep_epsilon: [0.02, 0.3]") \
            .replace("  runs: 1", "  runs: 1\n  output: %s" % (tmp_path / "sw"))
    p = tmp_path / "sweep.yaml"
    p.write_text(y)
    out = subprocess.run([grld, "-s", "9", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    rows = [l.split() for l in open(str(tmp_path / "sw") + "-0.txt") if l.startswith("# ") and l[2].isdigit()]
    assert len(rows) == 160
    n, cells = 160, 12
    gid = np.arange(n)
    blk, k = gid // 32, (gid // 32 // 3) * 32 + gid % 32          # lambda constant per block of 32 agents
    ia, il, ie = k % 2, blk % 3, (k // 2) % 2
    g = ia + 2 * (il + 3 * ie)                                    # cell id
    al, la, ep = np.array([0.1, 0.4])[ia], np.array([0.0, 0.65, 0.9])[il], np.array([0.02, 0.3])[ie]
    cfg = grl_b200.pendulum_config(memory=4096, randomization=1.0, test_interval=-1, seed=9, lambda_=0.9)
    b = grl_b200.Batch(cfg, n)
    try:
        b.set_hyper(al, la, ep)
        reports = 0
        while True:
            b.run(100)
            reports += 1
            r = b.results()
            if r["steps"].min() >= 400:
                break
        for a, row in enumerate(rows):
            assert int(row[2]) == r["steps"][a] and int(row[3]) == r["episodes"][a] and float(row[4]) == r["last_return"][a]
    finally:
        b.close()
    cell_rows = [l.split() for l in open(str(tmp_path / "sw") + "-0-cells.txt") if not l.startswith("#")]
    assert len(cell_rows) == reports * len(np.unique(g))
    last = [c for c in cell_rows if int(c[0]) == reports - 1]
    for c in last:
        cell = int(c[1])
        members = r["last_return"][g == cell]
        first = int(np.argmax(g == cell))
        assert (float(c[2]), float(c[3]), float(c[4])) == (al[first], la[first], ep[first]) and int(c[5]) == members.size
        assert abs(float(c[6]) - members.mean()) <= 1e-6 * abs(members.mean())
        tol = 1e-6 * np.abs(members).max()                      # the file holds 9 significant digits
        assert members.min() - tol <= float(c[7]) <= float(c[8]) <= float(c[9]) <= members.max() + tol


@pytest.mark.gpu
@pytest.mark.parametrize("yaml,mk,kw", [
    ("cfg/pendulum/sarsa_tc_cuda.yaml", "pendulum_config", dict(randomization=1.0, test_interval=-1)),
    ("cfg/cart_pole/sarsa_tc_cuda.yaml", "cart_pole_config", dict(randomization=1.0)),
    ("cfg/cart_double_pole/ac_tc_cuda.yaml", "cart_double_pole_ac_config", dict(test_interval=10)),
    ("cfg/cart_double_pole/sarsa_tc_cuda.yaml", "cart_double_pole_config", dict()),
    ("cfg/pendulum/ac_tc_cuda.yaml", "pendulum_ac_config", dict(randomization=1.0)),
    ("cfg/cart_pole/ac_tc_cuda.yaml", "cart_pole_ac_config", dict()),
])
def test_role_interfaces_of_the_shims(grld, tmp_path, yaml, mk, kw):
    """`grld_b200 -i`: Environment::start/step, Projector::project, Representation::read/write/update of the CUDA
    objects (one-agent views, the reference's L2 signatures) against the same calls through the ctypes harness."""
    import grl_b200
    text = re.sub(r"agents: \d+", "agents: 4", open(os.path.join(ROOT, yaml)).read())
    p = tmp_path / "iface.yaml"
    p.write_text(text)
    out = subprocess.run([grld, "-i", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    got = json.loads(out.stdout.strip().splitlines()[-1])
    cfgj = json.loads(subprocess.run([grld, "-c", str(p)], capture_output=True, text=True).stdout)
    cfg = getattr(grl_b200, mk)(memory=cfgj["memory"], **kw)
    ac = cfg.algo == grl_b200.capi.ALGO_AC
    b = grl_b200.Batch(cfg, 4)
    try:
        st, obs, _ = b.task_start(np.array([0.5]), test=0)
        assert got["obs"] == list(obs[0])
        act = 1.5 if ac else cfg.actions[cfg.n_actions - 1]
        x = list(obs[0]) + ([] if ac else [act])
        _, idx = b.project(np.array([x]))
        assert got["index"] == [float(i) for i in idx[0]]
        q0 = b.read(0, idx[0], actor=ac)
        b.update(0, idx[0], 0.5 * ((q0 + 1.0) - q0), actor=ac)
        q1 = b.read(0, idx[0], actor=ac)
        b.update(0, idx[0], -0.25, actor=ac)
        q2 = b.read(0, idx[0], actor=ac)
        assert got["q"] == [q0, q1, q2] and q1 > q0 and q2 < q1
        nxt, obs2, rew, term = b.model_step(st, np.array([act]))
        assert got["next_obs"] == list(obs2[0]) and got["reward"] == rew[0] and got["terminal"] == term[0] and got["tau"] == 1.0
    finally:
        b.close()


# ---- round 2: registration by dlopen, experiment/multi, load_file / save_every, several GPUs --------------------------
def test_addon_registers_its_types_when_dlopened(grld):
    """GRL finds addons by globbing libaddon*.so and dlopen(RTLD_NOW|RTLD_GLOBAL) (grl.cpp:73-111); the static
    REGISTER_CONFIGURABLE objects do the rest (factory.h:81-91).  plugin_probe links the framework library only."""
    probe = os.path.join(HOST, "plugin_probe")
    ldd = subprocess.run(["ldd", probe], capture_output=True, text=True).stdout
    assert "libgrl_b200.so" in ldd and "addon" not in ldd and "grlcu" not in ldd
    types = ["environment/model/cuda_batched", "projector/tile_coding/cuda", "representation/linear/cuda",
             "predictor/sarsa/cuda", "experiment/online_learning", "experiment/multi"]
    out = subprocess.run([probe] + types + ["no/such/type"], capture_output=True, text=True)
    lines = [l for l in out.stdout.splitlines() if not l.startswith("[")]      # (log lines: [NTC] ...)
    assert lines[0].startswith("plugins 1 types_before 0 types_after ")
    assert int(lines[0].split()[-1]) >= 20
    for t, l in zip(types, lines[1:]):
        assert l == "%s %s" % (t, t)
    assert lines[-1].startswith("no/such/type MISSING") and out.returncode == 1


def _multi_yaml(instances, agents):
    """cfg/pendulum/sarsa_tc_cuda.yaml wrapped in an experiment/multi: the policy is configured INSIDE the prototype."""
    body = [l for l in open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml")).read().split("\n") if not l.startswith("#")]
    assert body[0] == "experiment:"
    inner = "\n".join("  " + l for l in body[1:] if l.strip())
    inner = inner.replace("experiment/agent", "experiment/experiment/agent").replace("experiment/environment", "experiment/experiment/environment")
    y = "experiment:\n  type: experiment/multi\n  instances: %d\n  experiment:\n%s\n" % (instances, inner)
    return y.replace("agents: 65536", "agents: %d" % agents)


def test_experiment_multi_maps_instances_onto_agents(grld, tmp_path):
    """experiment/multi (experiments/multi.cpp:38-75): instances become agents when every clone would own its policy;
    the reference's own example shares ONE policy between the clones and is rejected with an explanation."""
    p = tmp_path / "multi.yaml"
    p.write_text(_multi_yaml(4, 24))
    rc, out, err = check(grld, str(p))
    assert rc == 0, err
    assert json.loads(out)["agents"] == 96
    if os.path.isdir(REF_CFG):
        q = tmp_path / "shared.yaml"
        q.write_text(swap_types(open(os.path.join(REF_CFG, "pendulum/multi_sarsa_tc.yaml")).read()))
        rc, out, err = check(grld, str(q))
        assert rc == 1 and "share one policy" in err and "multi_sarsa_tc.yaml" in err


def test_save_every_values(grld, tmp_path):
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml")).read()
    for v, ok in (("never", True), ("run", True), ("test", True), ("trial", True), ("sometimes", False)):
        p = tmp_path / ("se_%s.yaml" % v)
        p.write_text(base.replace("  runs: 1", "  runs: 1\n  save_every: %s" % v))
        rc, out, err = check(grld, str(p))
        assert (rc == 0) == ok, (v, err)


@pytest.mark.gpu
def test_load_file_round_trip_from_reference_format(grld, tmp_path, orc):
    """f1: tables written by fwrite in the reference's layout (representation.h:201-263: raw little-endian doubles,
    `file` + path with '/' -> '_' + ".dat") -- here the ORACLE's learned tables -- load through `load_file`
    (online_learning.cpp:141-150, with $run substitution) and learning continues bit-exactly: grld_b200's per-agent results
    equal the ctypes harness started from the same tables, and a replay from the loaded tables stays in lock-step with
    the oracle."""
    import ctypes as C
    import grl_b200
    import _oracle as O
    n, mem = 32, 4096
    oc = O.pendulum_config(memory=mem, randomization=1.0, test_interval=-1, seed=9)
    L = O.lib()
    L.orc_run_total.restype = None
    L.orc_run_total.argtypes = [C.c_void_p, C.c_uint64]
    tables = []
    stem = str(tmp_path / "pol-0")                       # load_file: pol-$run
    for i in range(n):
        a = L.orc_create(C.byref(oc), 500 + i)
        L.orc_run_total(a, 400)
        t = np.ctypeslib.as_array(L.orc_table(a), shape=(mem,)).copy()
        L.orc_destroy(a)
        tables.append(t)
        t.tofile("%s-_experiment_agent_policy_representation%s.dat" % (stem, "" if i == 0 else "-agent%d" % i))
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml")).read()
    y = base.replace("agents: 65536", "agents: %d" % n).replace("steps: 10000", "steps: 300") \
            .replace("memory: 65536", "memory: %d" % mem) \
            .replace("  runs: 1", "  runs: 1\n  output: %s\n  load_file: %s\n  save_every: trial" % (tmp_path / "out", tmp_path / "pol-$run"))
    p = tmp_path / "load.yaml"
    p.write_text(y)
    out = subprocess.run([grld, "-s", "9", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    rows = [l.split() for l in open(str(tmp_path / "out") + "-0.txt") if l.startswith("# ") and l[2].isdigit()]
    assert len(rows) == n
    cfg = grl_b200.pendulum_config(memory=mem, randomization=1.0, test_interval=-1, seed=9)
    b = grl_b200.Batch(cfg, n)
    try:
        for i in range(n):
            b.set_table(i, tables[i])
        for _ in range(3):
            b.run(100)
        r = b.results()
        for a, row in enumerate(rows):
            assert int(row[2]) == r["steps"][a] and float(row[4]) == r["last_return"][a]
        # save_every: trial wrote a checkpoint after every report, named by the finished episodes (online_learning.cpp:281-291)
        ck = np.fromfile(str(tmp_path / "out") + "-run0-trial3-_experiment_agent_policy_representation-agent5.dat", dtype=np.float64)
        assert np.array_equal(ck, b.get_table(5))
        # a reserved bit pattern must not get in through a checkpoint
        bad = tables[0].copy()
        bad.view(np.uint64)[17] = np.uint64(0x7FF4C0DE00000003)
        with pytest.raises(grl_b200.GrlcuError):
            b.set_table(0, bad)
    finally:
        b.close()
    # replay from the loaded tables: fresh agents on both sides, the oracle's tables overwritten with the same files
    b = grl_b200.Batch(cfg, n, first_agent_id=500)
    ob = O.OracleBatch(oc, n, first_id=500)
    try:
        for i in range(n):
            b.set_table(i, tables[i])
            ob.table(i)[:] = tables[i]
        for t in range(150):
            d = ob.step()
            inp = dict(d)
            inp["force_action"] = np.maximum(d["action_index"], 0).astype(np.int32)
            o = b.replay_step(inp, want_trace=False)
            assert np.array_equal(o["index"], d["index"]) and np.array_equal(o["action_index"], d["action_index"]), t
            assert np.array_equal(o["q"], d["q"]) and np.array_equal(o["delta"], d["delta"]), t
        for i in range(n):
            assert np.array_equal(b.get_table(i), ob.table(i))
    finally:
        b.close(); ob.close()


@pytest.mark.gpu
def test_grld_on_two_gpus_equals_one_gpu(grld, tmp_path):
    """Agents shard over the GPUs of the box behind the plugin API (`devices`), one NCCL all-gather of the per-agent
    results per report: the output is keyed by global agent id and must not depend on the number of devices.  (The last
    column of a report row is wall time and is not compared.)"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_sweep_cuda.yaml")).read()
    outs = {}
    for tag, devs in (("one", "[0]"), ("two", "[0, 1]")):
        y = re.sub(r"agents: \d+", "agents: 256", base)
        y = re.sub(r"steps: \d+", "steps: 500", y, count=1)
        y = re.sub(r"memory: \d+\n", "memory: 32768\n", y, count=1)
        y = y.replace("  runs: 1", "  runs: 1\n  devices: %s\n  output: %s" % (devs, tmp_path / tag))
        p = tmp_path / (tag + ".yaml")
        p.write_text(y)
        out = subprocess.run([grld, "-v", "-s", "4", str(p)], capture_output=True, text=True)
        assert out.returncode == 0, out.stderr + out.stdout
        outs[tag] = out.stdout + out.stderr
        assert os.path.exists(str(tmp_path / tag) + "-0-cells.txt")
    strip = lambda path: [l if l.startswith("#") else " ".join(l.split()[:-1]) for l in open(path)]
    assert strip(str(tmp_path / "one") + "-0.txt") == strip(str(tmp_path / "two") + "-0.txt")
    assert open(str(tmp_path / "one") + "-0-cells.txt").read() == open(str(tmp_path / "two") + "-0-cells.txt").read()
    assert "NCCL all-gather" in outs["two"] and "NCCL all-gather" not in outs["one"]


def test_safe_parameter_reaches_the_c_abi(grld, tmp_path):
    """projector/tile_coding `safe` (tile_coding.cpp:35-43): 0, 1, 2 accepted and handed to the C ABI; 3 is illegal."""
    base = open(os.path.join(ROOT, "cfg/pendulum/sarsa_tc_cuda.yaml")).read()
    for v, ok in ((0, True), (1, True), (2, True), (3, False)):
        p = tmp_path / ("safe%d.yaml" % v)
        p.write_text(base.replace("        tilings: 16\n", "        tilings: 16\n        safe: %d\n" % v))
        rc, out, err = check(grld, str(p))
        assert (rc == 0) == ok, (v, err)
        if ok:
            assert json.loads(out)["safe"] == v
