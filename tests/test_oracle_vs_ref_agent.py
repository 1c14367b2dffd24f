"""The oracle's agent half against the reference's OWN agent-side sources, bit for bit (no GPU).

oracle/_ref/libgrlref_agent.so (oracle/Makefile, target `ref`; built here, where /root/reference exists) holds the
reference's tile_coding.cpp, linear.cpp, enumerated.cpp, uniform.cpp, greedy.cpp, q.cpp, action.cpp, predictor.cpp,
sarsa.cpp, td.cpp, ac.cpp, agents/td.cpp and agents/fixed.cpp compiled with the reference's own headers; only Eigen
(absent here) and one YAML type are stand-ins (oracle/ref_shim_eigen/).  oracle/ref_agent.cpp wires the objects the way
grld does for cfg/pendulum/sarsa_tc.yaml and cfg/cart_double_pole/ac.yaml.

Method: the oracle runs free in its glibc RNG mode (srand48(seed); per-object drand48_r streams seeded from lrand48() in
grld's construction order, utils.h:84-137) and records every step; then srand48(seed) again, the reference objects are
built and fed the SAME observations, rewards and terminal flags; they draw their own random numbers.  Compared: every
action (episode starts included), and the weight tables at the end -- equality of the final tables means every read,
target, TD error, write, trace sweep and trace update along the way was the same.
"""
import ctypes as C
import os

import numpy as np
import pytest

import _oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libgrlref_agent.so")

pytestmark = pytest.mark.skipif(not os.path.exists(LIB),
                                reason="oracle/_ref/libgrlref_agent.so is built where /root/reference exists (make -C oracle ref)")


@pytest.fixture(scope="module")
def ref():
    L = C.CDLL(LIB)
    dp = C.POINTER(C.c_double)
    L.refag_prepare.restype = None
    L.refag_create.restype = C.c_void_p
    L.refag_create.argtypes = [C.POINTER(O.Config), C.c_long]
    L.refag_destroy.argtypes = [C.c_void_p]
    L.refag_start.argtypes = [C.c_void_p, C.c_int, dp, dp]
    L.refag_step.argtypes = [C.c_void_p, C.c_int, dp, C.c_double, C.c_int, dp]
    L.refag_table.argtypes = [C.c_void_p, C.c_int, dp]
    L.refag_project.argtypes = [C.c_void_p, C.c_int, dp, C.c_int, C.POINTER(C.c_uint32)]
    L.refag_prepare()
    return L


def _bits(x):
    return np.asarray(x, dtype=np.float64).view(np.uint64)


def _run_both(ref, cfg, seed, steps):
    L = O.lib()
    ac = cfg.algo == O.ALGO_AC
    cfg.rng = O.RNG_GLIBC
    # 1. the oracle, free-running, recorded
    L.orc_glibc_seed(seed)
    a = L.orc_create(C.byref(cfg), 0)
    recs = []
    for _ in range(steps):
        r = O.Record()
        L.orc_step(a, C.byref(r))
        recs.append(r)
    table = np.ctypeslib.as_array(L.orc_table(a), shape=(cfg.memory,)).copy()
    actor = np.ctypeslib.as_array(L.orc_actor_table(a), shape=(cfg.memory,)).copy() if ac else None
    L.orc_destroy(a)

    # 2. the reference's objects on the oracle's observations
    h = ref.refag_create(C.byref(cfg), seed)        # srand48(seed) inside, at the point grld's construction starts
    assert h, "the reference's configure() threw"
    OBS = cfg.obs_dims
    act = C.c_double()
    n_actions_checked = n_starts = n_ends = n_test = 0
    for k, r in enumerate(recs):
        if r.started:
            obs0 = (C.c_double * O.MAX_DIMS)()
            term = C.c_int32()
            L.orc_observe(C.byref(cfg), r.start_state, obs0, C.byref(term))
            assert ref.refag_start(h, r.test, obs0, C.byref(act)) == 0
            want = r.start_q[3] if ac else cfg.actions[r.start_action_index]
            assert _bits(act.value) == _bits(want), "start action at step %d: %r vs %r" % (k, act.value, want)
            n_starts += 1
        assert _bits(r.action) == _bits(act.value) or r.started, "step %d applies another action than the reference chose" % k
        obs = (C.c_double * O.MAX_DIMS)(*r.obs)
        assert ref.refag_step(h, r.test, obs, r.reward, r.terminal, C.byref(act)) == 0
        n_test += r.test
        if r.terminal == 2:
            n_ends += 1
            continue
        want = r.q[3] if ac else cfg.actions[r.action_index]
        assert _bits(act.value) == _bits(want), "action at step %d: %r vs %r" % (k, act.value, want)
        n_actions_checked += 1
    out = np.empty(cfg.memory)
    assert ref.refag_table(h, 0, out.ctypes.data_as(C.POINTER(C.c_double))) == cfg.memory
    assert np.array_equal(_bits(out), _bits(table)), "%d entries of the %s table differ" % (
        int((_bits(out) != _bits(table)).sum()), "critic" if ac else "Q")
    if ac:
        assert ref.refag_table(h, 1, out.ctypes.data_as(C.POINTER(C.c_double))) == cfg.memory
        assert np.array_equal(_bits(out), _bits(actor)), "%d entries of the actor table differ" % int((_bits(out) != _bits(actor)).sum())
    ref.refag_destroy(h)
    return dict(actions=n_actions_checked, starts=n_starts, absorbing_ends=n_ends, test_steps=n_test,
                changed=int((table != table[0]).sum()))


@pytest.mark.parametrize("seed", [1, 7])
def test_pendulum_sarsa_agent_equals_the_reference_sources(ref, seed):
    # cfg/pendulum/sarsa_tc.yaml: epsilon-greedy Q policy, SARSA(lambda) with the replacing trace, greedy test episodes
    cfg = O.pendulum_config(memory=65536, test_interval=3, randomization=1.0)
    s = _run_both(ref, cfg, seed, 4000)
    assert s["actions"] > 3900 and s["starts"] >= 40 and s["test_steps"] > 500


def test_pendulum_sarsa_small_table_collisions_and_accumulating_trace(ref):
    cfg = O.pendulum_config(memory=257, test_interval=-1, randomization=1.0)
    _run_both(ref, cfg, 3, 1500)
    cfg = O.pendulum_config(memory=4096, test_interval=-1, randomization=1.0, trace=O.TRACE_ACCUMULATING)
    _run_both(ref, cfg, 3, 1500)
    cfg = O.pendulum_config(memory=4096, test_interval=-1, randomization=1.0, trace=O.TRACE_NONE)
    _run_both(ref, cfg, 3, 800)


@pytest.mark.parametrize("safe,memory", [(1, 32768), (2, 32768), (2, 65536)])
def test_collision_claim_table_equals_the_reference_sources(ref, safe, memory):
    # projector/tile_coding safe = 1 | 2: the claim table of tile_coding.h:116-148 (linear probing; safe = 2 claims on reads too).
    # The table must not fill up: the reference then probes forever (tile_coding.h:127-137 has no lap guard).
    cfg = O.pendulum_config(memory=memory, test_interval=4, randomization=1.0, safe=safe)
    _run_both(ref, cfg, 9, 2500)


@pytest.mark.parametrize("algo,kw", [(O.ALGO_Q, {}), (O.ALGO_EXPECTED_SARSA, dict(target_epsilon=-1.0)),
                                     (O.ALGO_EXPECTED_SARSA, dict(target_epsilon=0.1))])
def test_q_learning_and_expected_sarsa_equal_the_reference_sources(ref, algo, kw):
    # predictor/critic/q (advantage.cpp:71-110) and predictor/critic/expected_sarsa (sarsa.cpp:167-194; greedy or
    # epsilon-greedy target policy)
    cfg = O.pendulum_config(memory=65536, test_interval=5, randomization=1.0, algo=algo, **kw)
    _run_both(ref, cfg, 4, 3000)


def test_softmax_sampler_equals_the_reference_sources(ref):
    # sampler/softmax as the learning policy's sampler (softmax.cpp:48-101, utils.h:227-240)
    cfg = O.pendulum_config(memory=65536, test_interval=5, randomization=1.0, sampler=O.SAMPLER_SOFTMAX, softmax_tau=0.5)
    _run_both(ref, cfg, 6, 3000)


def test_cart_pole_sarsa_agent_equals_the_reference_sources(ref):
    # absorbing end-stops: Agent::end with an empty action (sarsa.cpp:104-109)
    cfg = O.cart_pole_config(memory=262144, test_interval=5, randomization=1.0)
    s = _run_both(ref, cfg, 2, 4000)
    assert s["absorbing_ends"] >= 1


@pytest.mark.parametrize("which", ["cart_double_pole", "pendulum", "cart_pole"])
def test_actor_critic_agent_equals_the_reference_sources(ref, which):
    # cfg/cart_double_pole/ac.yaml (and its pendulum / cart-pole siblings): policy/parameterized/action with Gaussian
    # exploration drawn from the thread's RandGen (Box-Muller, utils.h:120-125), predictor/ac/action around
    # predictor/critic/td with a replacing trace -- the path the reference's golden vector no longer pins
    cfg = {"cart_double_pole": O.cart_double_pole_ac_config, "pendulum": O.pendulum_ac_config,
           "cart_pole": O.cart_pole_ac_config}[which](test_interval=4)
    s = _run_both(ref, cfg, 5, 3000)
    assert s["actions"] > 2000 and s["starts"] >= 10


def test_projector_alone_equals_the_reference_sources(ref):
    # TileCodingProjector::project incl. wrapping and negative coordinates, against orc_project
    L = O.lib()
    cfg = O.pendulum_config(memory=65536)
    cfg.rng = O.RNG_GLIBC
    h = ref.refag_create(C.byref(cfg), 11)
    rng = np.random.default_rng(0)
    dims, T = cfg.obs_dims + 1, cfg.tilings
    res = (C.c_double * dims)(*[cfg.resolution[i] for i in range(dims)])
    wrp = (C.c_double * dims)(*[cfg.wrapping[i] for i in range(dims)])
    hs, want, got = (C.c_uint32 * T)(), (C.c_uint32 * T)(), (C.c_uint32 * T)()
    for _ in range(2000):
        x = [rng.uniform(-7, 14), rng.uniform(-60, 60), rng.choice([-3.0, 0.0, 3.0])]
        v = (C.c_double * dims)(*x)
        L.orc_project(res, wrp, dims, T, cfg.memory, v, hs, want)
        assert ref.refag_project(h, 0, v, dims, got) == T
        assert list(got) == list(want), x
    ref.refag_destroy(h)


@pytest.mark.parametrize("dyn,S,tau,umax", [(O.DYN_PENDULUM, 3, 0.03, 3.0), (O.DYN_CART_POLE, 5, 0.05, 15.0),
                                            (O.DYN_CART_DOUBLE_POLE, 7, 0.05, 20.0)])
def test_rk4_model_step_equals_the_reference_sources(ref, dyn, S, tau, umax):
    # DynamicalModel::step (modeled.cpp:254-276) of the reference's own modeled.cpp around its own dynamics, against orc_rk4:
    # same libm, same order of operations -> the same bits, for all three systems (the 7-state one has no golden vector)
    L = O.lib()
    dp = C.POINTER(C.c_double)
    ref.refag_model_step.restype = C.c_double
    ref.refag_model_step.argtypes = [C.c_int, C.c_double, C.c_int, dp, C.c_double, dp]
    rng = np.random.default_rng(dyn)
    worst = 0
    for k in range(3000):
        x = rng.normal(0, 1.5, S)
        if dyn != O.DYN_PENDULUM:
            x[0] = rng.uniform(-2.6, 2.6)           # both sides of the end stops at +-2.4
        x[S - 1] = rng.uniform(0, 5)                # time
        u = float(rng.uniform(-umax, umax))
        xs = (C.c_double * S)(*x)
        want, got = (C.c_double * S)(), (C.c_double * S)()
        L.orc_rk4(dyn, tau, 5, xs, u, want)
        assert ref.refag_model_step(dyn, tau, 5, xs, u, got) == tau
        assert np.array_equal(_bits(list(got)), _bits(list(want))), (k, list(x), u, list(got), list(want))


@pytest.mark.parametrize("name,steps", [("pendulum_config", 6000), ("cart_pole_config", 6000), ("cart_double_pole_config", 4000),
                                        ("cart_double_pole_ac_config", 6000), ("pendulum_ac_config", 4000),
                                        ("cart_pole_ac_config", 4000)])
def test_closed_loop_of_reference_pieces_equals_the_free_running_oracle(ref, name, steps):
    # The whole loop with nothing fed: the reference's own task (start / actuate / observe / evaluate), its own
    # DynamicalModel::step and its own agents in a closed loop (refag_run; only the call order of ModeledEnvironment::step and of
    # the experiment's episode loop is restated in the harness) against orc_step free-running in glibc RNG mode from the same
    # seed.  Same libm on both sides: environment states, counters and weight tables must be the same bits.
    L = O.lib()
    cfg = getattr(O, name)(test_interval=6, randomization=1.0)
    cfg.rng = O.RNG_GLIBC
    ac = cfg.algo == O.ALGO_AC
    seed = 21
    L.orc_glibc_seed(seed)
    a = L.orc_create(C.byref(cfg), 0)
    for _ in range(steps):
        L.orc_step(a, None)
    n_steps, n_eps = C.c_uint64(), C.c_uint64()
    L.orc_stats(a, C.byref(n_steps), C.byref(n_eps), None, None, None)
    state = (C.c_double * O.MAX_STATE)()
    L.orc_get_state.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_get_state(a, state, None, None, None)
    table = np.ctypeslib.as_array(L.orc_table(a), shape=(cfg.memory,)).copy()
    actor = np.ctypeslib.as_array(L.orc_actor_table(a), shape=(cfg.memory,)).copy() if ac else None
    L.orc_destroy(a)

    dp = C.POINTER(C.c_double)
    ref.refag_run.argtypes = [C.c_void_p, C.c_long, dp, dp]
    h = ref.refag_create(C.byref(cfg), seed)
    assert h
    out, st = (C.c_double * 3)(), (C.c_double * O.MAX_STATE)()
    assert ref.refag_run(h, steps, out, st) == 0
    S = L.orc_state_dims(cfg.dynamics)
    assert int(out[0]) == n_steps.value and int(out[1]) == n_eps.value, (list(out), n_steps.value, n_eps.value)
    assert np.array_equal(_bits(list(st)[:S]), _bits(list(state)[:S])), (list(st)[:S], list(state)[:S])
    got = np.empty(cfg.memory)
    ref.refag_table(h, 0, got.ctypes.data_as(dp))
    assert np.array_equal(_bits(got), _bits(table)), int((_bits(got) != _bits(table)).sum())
    if ac:
        ref.refag_table(h, 1, got.ctypes.data_as(dp))
        assert np.array_equal(_bits(got), _bits(actor))
    assert n_eps.value >= 5
    ref.refag_destroy(h)


@pytest.mark.parametrize("lam,alpha,eps", [(0.0, 0.4, 0.1), (0.2, 0.05, 0.01), (0.9, 0.1, 0.02), (0.75, 0.2, 0.05)])
def test_sweep_corner_hyperparameters_equal_the_reference_sources(ref, lam, alpha, eps):
    # BASELINE's configs[4] sweeps alpha x lambda x epsilon: lambda = 0 makes the trace's decay 0 < 0.01, i.e. clear() and a
    # one-entry trace swept with e = 0 (trace.h:220-226); lambda = 0.9 is the longest ring (gamma*lambda = 0.873, 34 entries)
    cfg = O.pendulum_config(memory=32768, test_interval=-1, randomization=1.0, lambda_=lam, alpha=alpha, epsilon=eps)
    _run_both(ref, cfg, 12, 3000)
