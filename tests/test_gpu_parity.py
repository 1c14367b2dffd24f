"""GPU parity tests proper: libgrlcu (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): tile hashes / indices, action choices and trace index sets bit-exact;
dynamics and Q-values within 1e-9 relative per step (the agent half is in fact bit-exact in FP64 too,
because it is fed identical inputs and keeps the reference's summation order)."""
import ctypes as C

import numpy as np
import pytest

import _oracle as O

pytestmark = pytest.mark.gpu

RTOL = 1e-9


@pytest.fixture(scope="module")
def G():
    import grl_b200
    grl_b200.load()
    return grl_b200


def _cfgs(G):
    return [("pendulum", G.pendulum_config, O.pendulum_config, 65536),
            ("cart_pole", G.cart_pole_config, O.cart_pole_config, 262144),
            ("cart_double_pole", G.cart_double_pole_config, O.cart_double_pole_config, 131072)]


def _rel(a, b):
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)) if a.size else 0.0


@pytest.mark.parametrize("which", [0, 1, 2])
def test_tile_hash_bit_exact(G, orc, which):
    name, gcfg, ocfg, mem = _cfgs(G)[which]
    rng = np.random.default_rng(7 + which)
    cfg = gcfg(memory=mem)
    dims = cfg.obs_dims + 1
    n = 20000
    x = rng.uniform(-30, 30, size=(n, dims))
    x[: n // 4] *= 1e-3                       # near the origin (negative/positive boundary)
    x[n // 4: n // 2] = np.round(x[n // 4: n // 2])   # exactly on tile boundaries
    b = G.Batch(cfg, 1)
    try:
        h, idx = b.project(x)
    finally:
        b.close()
    oc = ocfg(memory=mem)
    dp, u32p = C.POINTER(C.c_double), C.POINTER(C.c_uint32)
    oh, oi = np.empty(16, np.uint32), np.empty(16, np.uint32)
    for k in range(n):
        row = np.ascontiguousarray(x[k])
        orc.orc_project(oc.resolution, oc.wrapping, dims, 16, mem, row.ctypes.data_as(dp),
                        oh.ctypes.data_as(u32p), oi.ctypes.data_as(u32p))
        assert np.array_equal(h[k], oh), (name, k)
        assert np.array_equal(idx[k], oi), (name, k)


@pytest.mark.parametrize("which", [0, 1])
def test_model_step_within_tolerance(G, orc, which):
    name, gcfg, ocfg, mem = _cfgs(G)[which]
    rng = np.random.default_rng(11 + which)
    cfg, oc = gcfg(memory=1024), ocfg(memory=1024)
    S = 3 if which == 0 else 5
    n = 4000
    st = rng.uniform(-3, 3, size=(n, S))
    if which == 0:
        st[:, 0] = rng.uniform(-20, 20, n); st[:, 1] = rng.uniform(-40, 40, n); st[:, 2] = rng.uniform(0, 3.1, n)
    else:
        st[:, 0] = rng.uniform(-2.6, 2.6, n); st[:, 1] = rng.uniform(-10, 10, n)
        st[:, 2] = rng.uniform(-5, 5, n); st[:, 3] = rng.uniform(-15, 15, n); st[:, 4] = rng.uniform(0, 10.2, n)
    act = rng.choice(np.array(list(cfg.actions)[:3]), n)
    b = G.Batch(cfg, 1)
    try:
        nxt, obs, rew, term = b.model_step(st, act)
    finally:
        b.close()
    dp = C.POINTER(C.c_double)
    on, oo, orr, ot = np.empty(7), np.empty(8), C.c_double(), C.c_int32()
    for k in range(n):
        row = np.ascontiguousarray(st[k])
        orc.orc_env_step(C.byref(oc), row.ctypes.data_as(dp), float(act[k]), on.ctypes.data_as(dp),
                         oo.ctypes.data_as(dp), C.byref(orr), C.byref(ot))
        assert term[k] == ot.value
        assert _rel(nxt[k], on[:S]) <= RTOL, (name, k, nxt[k], on[:S])
        assert _rel(obs[k], oo[:cfg.obs_dims]) <= RTOL
        assert abs(rew[k] - orr.value) <= RTOL * max(abs(orr.value), 1e-300)


def test_table_init_bit_exact(G, orc):
    cfg, oc = G.pendulum_config(memory=4099, seed=5), O.pendulum_config(memory=4099, seed=5)
    b = G.Batch(cfg, 3, first_agent_id=1000)
    ob = O.OracleBatch(oc, 3, first_id=1000)
    try:
        for i in range(3):
            assert np.array_equal(b.get_table(i), ob.table(i))
    finally:
        b.close(); ob.close()


def _replay(G, gcfg, ocfg, n, steps, mem, check_tables=True, hyper=None, **kw):
    cfg, oc = gcfg(memory=mem, **kw), ocfg(memory=mem, **kw)
    b = G.Batch(cfg, n)
    ob = O.OracleBatch(oc, n)
    mism_actions = 0
    try:
        if hyper is not None:
            b.set_hyper(*hyper)
            ob.set_hyper(*hyper)
        for t in range(steps):
            d = ob.step()
            inp = dict(d)
            inp["force_action"] = np.where(d["action_index"] < 0, 0, d["action_index"]).astype(np.int32)
            out = b.replay_step(inp)
            live = d["terminal"] != 2
            # environment half: 1e-9 relative
            assert _rel(out["next"], d["next"]) <= RTOL, t
            assert _rel(out["obs"], d["obs"]) <= RTOL, t
            assert _rel(out["reward"], d["reward"]) <= RTOL, t
            assert np.array_equal(out["terminal"], d["terminal"]), t
            # agent half: integers bit-exact
            assert np.array_equal(out["hash"][:, :, live], d["hash"][:, :, live]), t
            assert np.array_equal(out["index"][:, :, live], d["index"][:, :, live]), t
            assert np.array_equal(out["action_index"], d["action_index"]), t
            st = d["started"] != 0
            assert np.array_equal(out["start_action_index"][st], d["start_action_index"][st]), t
            assert np.array_equal(out["start_q"][:, st], d["start_q"][:, st]), t
            assert np.array_equal(out["trace_len"], d["trace_len"]), t
            assert np.array_equal(out["trace_index"], ob.trace(b.trace_cap)), t
            learn = d["test"] == 0
            assert np.array_equal(out["swept"][learn], d["swept"][learn]), t
            # FP of the agent half: identical inputs and order -> identical bits; bar is 1e-9
            assert _rel(out["q"][:, live], d["q"][:, live]) <= RTOL, t
            assert np.array_equal(out["q"][:, live], d["q"][:, live]), t
            for k in ("q_prev", "target", "delta"):
                assert np.array_equal(out[k][learn], d[k][learn]), (t, k)
        if check_tables:
            for i in range(n):
                assert np.array_equal(b.get_table(i), ob.table(i)), i
        res = b.results()
        for i, a in enumerate(ob.agents):
            s, e, w = C.c_uint64(), C.c_uint64(), C.c_uint64()
            orc_l = O.lib()
            orc_l.orc_stats(a, C.byref(s), C.byref(e), C.byref(w), None, None)
            assert (res["steps"][i], res["episodes"][i], res["swept"][i]) == (s.value, e.value, w.value)
    finally:
        b.close(); ob.close()
    return mism_actions


def test_replay_pendulum(G, orc):
    # 3 learning episodes + part of a 4th incl. a test episode (test_interval=2)
    _replay(G, G.pendulum_config, O.pendulum_config, n=48, steps=450, mem=65536,
            randomization=1.0, test_interval=2)


def test_replay_pendulum_shipped_memory(G, orc):
    """The reference's shipped table size (cfg/pendulum/sarsa_tc.yaml:32, memory 8,388,608 = 64 MiB per agent) at a
    reduced agent count, for index parity with the shipped configuration."""
    _replay(G, G.pendulum_config, O.pendulum_config, n=8, steps=300, mem=8388608,
            randomization=1.0, test_interval=2)


def test_replay_pendulum_small_table_collisions(G, orc):
    # memory 257: many intra-projection and cross-tiling collisions (duplicate indices)
    _replay(G, G.pendulum_config, O.pendulum_config, n=32, steps=330, mem=257,
            randomization=1.0, test_interval=-1)


def test_replay_cart_pole(G, orc):
    _replay(G, G.cart_pole_config, O.cart_pole_config, n=32, steps=500, mem=262144,
            randomization=1.0, test_interval=3, epsilon=0.3)


def test_replay_cart_double_pole_sarsa(G, orc):
    """The reference's cfg/cart_double_pole/sarsa.yaml: 6-D observations (8 ints per hash), lambda = 0.92 with
    gamma = 0.99 (a 51-entry trace ring, the 64-bit duplicate mask), optimistic initial values 20..21; early episodes
    are short, so starts, absorbing ends and test episodes are frequent."""
    _replay(G, G.cart_double_pole_config, O.cart_double_pole_config, n=32, steps=320, mem=131072, test_interval=3, epsilon=0.2)
    _replay(G, G.cart_double_pole_config, O.cart_double_pole_config, n=16, steps=200, mem=521, test_interval=-1,
            algo=O.ALGO_Q)


def test_replay_per_agent_hyperparameters(G, orc):
    """The sweep workload (MultiExperiment / grlo grids): alpha, lambda, epsilon per agent; lambda = 0.9 needs a
    35-entry trace ring (more than 32: the fast update paths carry a 64-bit duplicate mask)."""
    n = 64
    i = np.arange(n)
    hyper = (np.array([0.05, 0.1, 0.2, 0.4])[i % 4],
             np.array([0.0, 0.2, 0.4, 0.5, 0.65, 0.75, 0.85, 0.9])[(i // 4) % 8],
             np.array([0.01, 0.3])[i // 32])
    _replay(G, G.pendulum_config, O.pendulum_config, n=n, steps=320, mem=65536, hyper=hyper,
            randomization=1.0, test_interval=4, lambda_=0.9)
    _replay(G, G.pendulum_config, O.pendulum_config, n=n, steps=200, mem=307, hyper=hyper,
            randomization=1.0, test_interval=-1, lambda_=0.9)


def test_replay_q_learning_and_expected_sarsa(G, orc):
    """Sibling predictors on the same kernels: predictor/critic/q and predictor/critic/expected_sarsa
    (greedy and epsilon-greedy target policy) differ from SARSA only in what the target bootstraps from."""
    for algo, eps in ((O.ALGO_Q, 0.0), (O.ALGO_EXPECTED_SARSA, -1.0), (O.ALGO_EXPECTED_SARSA, 0.05)):
        _replay(G, G.pendulum_config, O.pendulum_config, n=32, steps=260, mem=16384,
                randomization=1.0, test_interval=3, algo=algo, target_epsilon=eps)
    _replay(G, G.cart_pole_config, O.cart_pole_config, n=16, steps=300, mem=262144,
            randomization=1.0, test_interval=-1, epsilon=0.2, algo=O.ALGO_Q)
    # small table: equal Q-values (ties) and duplicate indices
    _replay(G, G.pendulum_config, O.pendulum_config, n=32, steps=200, mem=67,
            randomization=1.0, test_interval=-1, algo=O.ALGO_EXPECTED_SARSA, target_epsilon=0.3)


def test_replay_softmax_sampler(G, orc):
    """sampler/softmax as the learning policy's sampler: own action choices from the fed uniform are
    bit-exact against the oracle (device exp vs glibc exp can only matter within an ulp of a bucket boundary), test
    episodes stay greedy; Q-learning on top of it as well."""
    _replay(G, G.pendulum_config, O.pendulum_config, n=48, steps=330, mem=65536,
            randomization=1.0, test_interval=2, sampler=O.SAMPLER_SOFTMAX, softmax_tau=0.5)
    _replay(G, G.cart_pole_config, O.cart_pole_config, n=32, steps=300, mem=4099,
            randomization=1.0, test_interval=-1, sampler=O.SAMPLER_SOFTMAX, softmax_tau=20.0, algo=O.ALGO_Q)
    with pytest.raises(G.GrlcuError):
        G.Batch(G.pendulum_config(memory=1024, sampler=O.SAMPLER_SOFTMAX, softmax_tau=0.0), 1)


def test_replay_accumulating_and_no_trace(G, orc):
    _replay(G, G.pendulum_config, O.pendulum_config, n=16, steps=150, mem=8192,
            randomization=1.0, test_interval=-1, trace=2)
    _replay(G, G.pendulum_config, O.pendulum_config, n=16, steps=150, mem=8192,
            randomization=1.0, test_interval=-1, trace=0)


def test_free_running_matches_oracle_statistically(G, orc):
    """Seeded Philox streams on both sides: identical draws; trajectories agree until a last-bit
    sin/cos difference flips a tile, so compare per-agent learning curves statistically."""
    n, episodes = 256, 60
    cfg = G.pendulum_config(memory=65536, randomization=1.0, test_interval=-1)
    oc = O.pendulum_config(memory=65536, randomization=1.0, test_interval=-1)
    b = G.Batch(cfg, n)
    try:
        b.run(100 * episodes)
        res = b.results()
    finally:
        b.close()
    assert np.all(res["steps"] == 100 * episodes) and np.all(res["episodes"] == episodes)
    L = O.lib()
    orc_ret, orc_last = [], []
    for i in range(32):
        a = L.orc_create(C.byref(oc), i)
        L.orc_run(a, 100 * episodes)
        lr = C.c_double()
        L.orc_stats(a, None, None, None, C.byref(lr), None)
        orc_last.append(lr.value)
        L.orc_destroy(a)
    g_mean, o_mean = float(np.mean(res["last_return"])), float(np.mean(orc_last))
    g_sd = float(np.std(res["last_return"]))
    # same distribution: means within 4 standard errors of the smaller sample
    assert abs(g_mean - o_mean) < 4 * g_sd / np.sqrt(32) + 1e-9, (g_mean, o_mean, g_sd)
    # learning happened: mean return of the summed curve improves on the first episode's level
    assert g_mean > -4000


def test_one_agent_views(G, orc):
    """grlcu_task_start / grlcu_read / grlcu_update: the single-agent views the L2-signature shims use."""
    for name, gcfg, ocfg, mem in _cfgs(G):
        rand = {} if name == "cart_double_pole" else dict(randomization=1.0)
        cfg = gcfg(memory=4096, **rand)
        oc = ocfg(memory=4096, **rand)
        b = G.Batch(cfg, 3)
        try:
            rng = np.random.default_rng(5)
            u = rng.uniform(0, 1, 64)
            st, obs, term = b.task_start(u, test=0)
            st_t, obs_t, _ = b.task_start(u, test=1)
            S = b.S
            # Task::start (pendulum.cpp:97-105, cart_pole.cpp:155-166, cart_double_pole.cpp:250-262): the randomised
            # coordinate only
            k = 0 if cfg.dynamics == 0 else 1
            want = [np.pi + u * 2 * np.pi, np.pi + (0.1 * u - 0.05), u * 0.01 - 0.005][cfg.dynamics]
            assert np.array_equal(st[:, k], want), name
            # pendulum test episodes are not randomised (pendulum.cpp:100); the others are (cart_pole.cpp:160)
            assert np.array_equal(st_t[:, k], np.full(64, np.pi) if cfg.dynamics == 0 else want), name
            mask = np.ones(S, bool); mask[k] = False
            assert not st[:, mask].any()
            dp, i32p = C.POINTER(C.c_double), C.POINTER(C.c_int32)
            for i in range(64):
                o = np.empty(8); t = C.c_int32()
                row = np.ascontiguousarray(np.concatenate([st[i], np.zeros(8 - S)]))
                orc.orc_observe(C.byref(oc), row.ctypes.data_as(dp), o.ctypes.data_as(dp), C.byref(t))
                assert np.allclose(obs[i], o[:cfg.obs_dims], rtol=1e-12, atol=0) and term[i] == t.value
            # read / update on agent 1's table, duplicates and an invalidated index included
            tbl = b.get_table(1)
            idx = rng.integers(0, 4096, 16).astype(np.uint64)
            idx[5] = idx[2]
            want = 0.0
            for x in idx:
                want += tbl[int(x)]
            assert b.read(1, idx) == want / 16
            upd = idx.copy(); upd[7] = np.uint64(0xFFFFFFFFFFFFFFFF)
            b.update(1, upd, 0.125)
            ref = tbl.copy()
            for x in upd:
                if x != np.uint64(0xFFFFFFFFFFFFFFFF):
                    ref[int(x)] += 0.125
            assert np.array_equal(b.get_table(1), ref)
            assert np.array_equal(b.get_table(0), b.get_table(0)) and not np.array_equal(b.get_table(0), ref)
            with pytest.raises(G.GrlcuError):
                b.read(1, np.array([4096], np.uint64))
        finally:
            b.close()


def test_random_access_probe(G):
    """grlcu_probe_random_access: the live denominators of bench.py's random-access roofline."""
    r, w, m = G.capi.probe_random_access(0, 1 << 31)
    assert 1e10 < m < w < r < 1e12            # B200: about 37 G reads/s, 20 G writes/s, 13.5 G read-modify-writes/s
    with pytest.raises(G.GrlcuError):
        G.capi.probe_random_access(0, 3 << 20)


def test_run_host_start_states_apply_to_the_next_episode_an_agent_begins(G):
    """grlcu_run_host: the uploaded start-state row replaces Task::start for the next episode the agent BEGINS; an episode
    in progress continues (so the end-to-end leg of bench.py runs the same trajectory phase as the device-resident one)."""
    cfg = G.pendulum_config(memory=4096, randomization=1.0, test_interval=-1)
    n = 8
    b = G.Batch(cfg, n)
    try:
        s1 = np.zeros((n, 3)); s1[:, 0] = 1.0 + 0.1 * np.arange(n)
        b.run_host(s1, 1)
        st = np.array([b.get_state(i)[0] for i in range(n)])
        assert np.allclose(st[:, 2], 0.03) and np.all(np.abs(st[:, 0] - s1[:, 0]) < 0.2)      # began from the rows
        s2 = np.zeros((n, 3)); s2[:, 0] = -2.0
        b.run_host(s2, 1)
        st2 = np.array([b.get_state(i)[0] for i in range(n)])
        assert np.allclose(st2[:, 2], 0.06) and np.all(np.abs(st2[:, 0] - st[:, 0]) < 0.5)     # episode in progress continues
        b.run_host(None, 98)                                                                    # ... to its end (100 steps)
        r = b.results()
        assert np.all(r["episodes"] == 1)
        b.run_host(None, 1)                                                                     # next episode: from the pending rows
        st3 = np.array([b.get_state(i)[0] for i in range(n)])
        assert np.allclose(st3[:, 2], 0.03) and np.all(np.abs(st3[:, 0] + 2.0) < 0.2)
    finally:
        b.close()


@pytest.mark.parametrize("safe,mem,steps", [(1, 8209, 300), (2, 16411, 300), (1, 65536, 320)])
def test_replay_collision_claim_table(G, orc, safe, mem, steps):
    """projector/tile_coding `safe` > 0 (tile_coding.h:116-148): per-agent claim table with linear probing; the policy's
    multi-projection claims only with safe = 2, the predictor's own projections of (s, a) and (s', a') always do, so a
    location can move between the act-time read and the update.  Indices, actions, traces, Q-values and tables bit-exact in
    replay, and the pressure counter (longest probe walk) equal.  Table sizes leave free slots for every claim."""
    kw = dict(randomization=1.0, test_interval=3, safe=safe)
    _replay(G, G.pendulum_config, O.pendulum_config, n=32, steps=steps, mem=mem, **kw)
    # pressure: run a second, shorter pair by hand to read both counters
    cfg, oc = G.pendulum_config(memory=mem, **kw), O.pendulum_config(memory=mem, **kw)
    b, ob = G.Batch(cfg, 16), O.OracleBatch(oc, 16)
    try:
        for t in range(120):
            d = ob.step()
            inp = dict(d)
            inp["force_action"] = np.maximum(d["action_index"], 0).astype(np.int32)
            b.replay_step(inp, want_trace=False)
        orc.orc_pressure.restype = C.c_uint64
        orc.orc_pressure.argtypes = [C.c_void_p]
        want = np.array([orc.orc_pressure(a) for a in ob.agents], dtype=np.uint32)
        assert np.array_equal(b.pressure(), want)
    finally:
        b.close(); ob.close()


def test_collision_claim_table_free_running_and_limits(G, orc):
    """The fused kernel with a claim table against the free-running oracle, agent by agent; safe is SARSA-only."""
    kw = dict(memory=65536, randomization=1.0, test_interval=4, safe=1, epsilon=0.2)
    cfg, oc = G.pendulum_config(**kw), O.pendulum_config(**kw)
    n, steps = 128, 260
    orc.orc_run_total.restype = None
    orc.orc_run_total.argtypes = [C.c_void_p, C.c_uint64]
    b = G.Batch(cfg, n, first_agent_id=77)
    try:
        b.run(100); b.run(steps - 100)
        res = b.results()
        for i in range(n):
            a = orc.orc_create(C.byref(oc), 77 + i)
            orc.orc_run_total(a, steps)
            s, e, w, lr = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_double()
            orc.orc_stats(a, C.byref(s), C.byref(e), C.byref(w), C.byref(lr), None)
            t = np.ctypeslib.as_array(orc.orc_table(a), shape=(cfg.memory,))
            assert (int(res["steps"][i]), int(res["episodes"][i]), int(res["swept"][i])) == (s.value, e.value, w.value), i
            assert abs(res["last_return"][i] - lr.value) <= RTOL * abs(lr.value), i
            assert np.allclose(b.get_table(i), t, rtol=RTOL, atol=0), i
            orc.orc_destroy(a)
    finally:
        b.close()
    with pytest.raises(G.GrlcuError):
        G.Batch(G.pendulum_config(memory=4096, safe=1, algo=O.ALGO_Q), 2)
    with pytest.raises(G.GrlcuError):
        G.Batch(G.pendulum_config(memory=4096, safe=3), 2)


@pytest.mark.parametrize("which,A", [(0, 2), (0, 5), (1, 4)])
def test_replay_other_action_counts(G, orc, which, A):
    """discretizer/uniform with `steps` other than 3 (uniform.cpp:60-94: min + k (max - min) / (steps - 1)): 2, 4 and 5
    actions run the uncached kernels; replay bit-exact, and the fused kernel per agent against the free-running oracle."""
    name, gcfg, ocfg, mem = _cfgs(G)[which]
    amax = (3.0, 15.0)[which]
    acts = [-amax + k * (2 * amax) / (A - 1) for k in range(A)] + [0.0] * (8 - A)
    def mk(f):
        def g(**kw):
            c = f(**kw)
            c.n_actions = A
            for i, v in enumerate(acts):
                c.actions[i] = v
            return c
        return g
    _replay(G, mk(gcfg), mk(ocfg), n=32, steps=260, mem=16384, randomization=1.0, test_interval=3, epsilon=0.2)
    cfg = mk(gcfg)(memory=16384, randomization=1.0, test_interval=-1)
    oc = mk(ocfg)(memory=16384, randomization=1.0, test_interval=-1)
    orc.orc_run_total.restype = None
    orc.orc_run_total.argtypes = [C.c_void_p, C.c_uint64]
    b = G.Batch(cfg, 64, first_agent_id=9)
    try:
        b.run(150)
        res = b.results()
        for i in range(64):
            a = orc.orc_create(C.byref(oc), 9 + i)
            orc.orc_run_total(a, 150)
            s, e, w, lr = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_double()
            orc.orc_stats(a, C.byref(s), C.byref(e), C.byref(w), C.byref(lr), None)
            assert (int(res["steps"][i]), int(res["episodes"][i]), int(res["swept"][i])) == (s.value, e.value, w.value), i
            assert np.allclose(b.get_table(i), np.ctypeslib.as_array(orc.orc_table(a), shape=(16384,)), rtol=RTOL, atol=0), i
            orc.orc_destroy(a)
    finally:
        b.close()
    with pytest.raises(G.GrlcuError):
        c = gcfg(memory=1024); c.n_actions = 7
        G.Batch(c, 2)


def test_replay_visit_ordered_layout(G, orc):
    """Contexts of 64 agents or more whose tiles do not fit the table get the visit-ordered physical layout (a pilot run
    inside grlcu_create): replay parity -- logical indices, traces, Q-values, tables -- must not notice."""
    b = G.Batch(G.cart_pole_config(memory=65536, randomization=1.0), 64)
    try:
        info = b.layout_info()
        assert info["kind"].startswith("visit-ordered") and info["inbox_words"] > 0, info
    finally:
        b.close()
    _replay(G, G.cart_pole_config, O.cart_pole_config, n=64, steps=220, mem=65536, randomization=1.0, test_interval=3, epsilon=0.3)
    _replay(G, G.cart_double_pole_config, O.cart_double_pole_config, n=64, steps=120, mem=131072, test_interval=-1, epsilon=0.2)
    # the pendulum with a table that its box fills more than half: visit-ordered as well
    _replay(G, G.pendulum_config, O.pendulum_config, n=64, steps=220, mem=32768, randomization=1.0, test_interval=2)
