"""Actor-critic path (BASELINE configs[3]: cart-double-pole, two tile-coded tables per agent) against the
oracle in replay mode: states, actions, observations, rewards and the Gaussian exploration term are fed to
both; tile hashes/indices and trace sets must be bit-exact, the environment half within 1e-9 relative, and
everything the agent half computes (pi, V, u, a', target, delta, both tables) bit-exact as well."""
import ctypes as C

import numpy as np
import pytest

import _oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def _rel(a, b):
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)) if a.size else 0.0


@pytest.fixture(scope="module")
def G():
    import grl_b200
    grl_b200.load()
    return grl_b200


def test_double_pole_projection_and_dynamics(G, orc):
    cfg = G.cart_double_pole_ac_config(memory=131072)
    oc = O.cart_double_pole_ac_config(memory=131072)
    rng = np.random.default_rng(11)
    b = G.Batch(cfg, 1)
    try:
        # projection of 6-D observations: 7 ints per hash (tile_coding.cpp:103-149)
        n = 5000
        x = rng.uniform(-3, 3, size=(n, 6))
        x[: n // 4] *= 1e-3
        h, idx = b.project(x)
        dp, u32p = C.POINTER(C.c_double), C.POINTER(C.c_uint32)
        oh, oi = np.empty(16, np.uint32), np.empty(16, np.uint32)
        for k in range(n):
            row = np.ascontiguousarray(x[k])
            orc.orc_project(oc.resolution, oc.wrapping, 6, 16, 131072, row.ctypes.data_as(dp),
                            oh.ctypes.data_as(u32p), oi.ctypes.data_as(u32p))
            assert np.array_equal(h[k], oh) and np.array_equal(idx[k], oi), k
        # RK4 over the cart-double-pole equations of motion incl. the 3x3 inverse and the end stops
        m = 4000
        st = np.zeros((m, 7))
        st[:, 0] = rng.uniform(-2.6, 2.6, m)
        st[:, 1:3] = rng.uniform(-0.9, 0.9, (m, 2))
        st[:, 3:6] = rng.uniform(-3, 3, (m, 3))
        st[:, 6] = rng.uniform(0, 5.2, m)
        act = rng.uniform(-20, 20, m)
        nxt, obs, rew, term = b.model_step(st, act)
        for k in range(m):
            on, oo, orw, ot = np.empty(7), np.empty(8), C.c_double(), C.c_int32()
            s = np.ascontiguousarray(st[k])
            orc.orc_env_step(C.byref(oc), s.ctypes.data_as(dp), C.c_double(act[k]), on.ctypes.data_as(dp),
                             oo.ctypes.data_as(dp), C.byref(orw), C.byref(ot))
            assert _rel(nxt[k], on) <= RTOL, k
            assert _rel(obs[k], oo[:6]) <= RTOL and term[k] == ot.value, k
            assert abs(rew[k] - orw.value) <= RTOL * max(abs(orw.value), 1e-300), k
    finally:
        b.close()


def _replay_ac(G, n, steps, mem, which="cart_double_pole", **kw):
    cfg, oc = getattr(G, which + "_ac_config")(memory=mem, **kw), getattr(O, which + "_ac_config")(memory=mem, **kw)
    b = G.Batch(cfg, n)
    ob = O.OracleBatch(oc, n)
    try:
        for t in range(steps):
            d = ob.step()
            out = b.replay_step(dict(d))
            live = d["terminal"] != 2
            assert _rel(out["next"], d["next"]) <= RTOL, t
            assert _rel(out["obs"], d["obs"]) <= RTOL, t
            assert _rel(out["reward"], d["reward"]) <= RTOL, t
            assert np.array_equal(out["terminal"], d["terminal"]), t
            # one projection serves actor and critic (same projector parameters, unseeded hash)
            assert np.array_equal(d["hash"][0], d["hash"][1])
            assert np.array_equal(out["hash"][0][:, live], d["hash"][0][:, live]), t
            assert np.array_equal(out["index"][0][:, live], d["index"][0][:, live]), t
            assert np.array_equal(out["trace_len"], d["trace_len"]), t
            assert np.array_equal(out["trace_index"], ob.trace(b.trace_cap)), t
            learn = d["test"] == 0
            assert np.array_equal(out["swept"][learn], d["swept"][learn]), t
            # pi(s'), V(s'), a' where the agent acted; u, V(p), target, delta where it learned
            for row in (0, 1, 3):
                assert np.array_equal(out["q"][row][live], d["q"][row][live]), (t, row)
            assert np.array_equal(out["q"][2][learn], d["q"][2][learn]), t
            st = d["started"] != 0
            assert np.array_equal(out["start_q"][0][st], d["start_q"][0][st]), t
            assert np.array_equal(out["start_q"][3][st], d["start_q"][3][st]), t
            for k in ("q_prev", "target", "delta"):
                assert np.array_equal(out[k][learn], d[k][learn]), (t, k)
        for i in range(n):
            assert np.array_equal(b.get_table(i), ob.table(i)), i
            assert np.array_equal(b.get_actor_table(i), ob.actor_table(i)), i
        res = b.results()
        for i, a in enumerate(ob.agents):
            s, e, w = C.c_uint64(), C.c_uint64(), C.c_uint64()
            O.lib().orc_stats(a, C.byref(s), C.byref(e), C.byref(w), None, None)
            assert (res["steps"][i], res["episodes"][i], res["swept"][i]) == (s.value, e.value, w.value)
    finally:
        b.close(); ob.close()


def test_replay_ac(G, orc):
    # early episodes are short (the poles fall within ~10 steps): many starts, absorbing ends and test episodes
    _replay_ac(G, n=48, steps=400, mem=131072, test_interval=3)


def test_replay_ac_small_table_and_saturating_actor(G, orc):
    # memory 509: duplicate indices inside projections; a large actor step drives weights into the +-20 clamp
    _replay_ac(G, n=32, steps=300, mem=509, test_interval=-1, actor_alpha=0.5, sigma=15.0)


def test_replay_ac_pendulum(G, orc):
    # the reference's cfg/pendulum/ac_tc.yaml: same agent on the pendulum swing-up (2-D observations, wrapping angle)
    _replay_ac(G, n=48, steps=330, mem=65536, which="pendulum", test_interval=2, randomization=1.0)
    _replay_ac(G, n=16, steps=250, mem=251, which="pendulum", test_interval=-1, actor_alpha=0.3, sigma=4.0)


def test_replay_ac_cart_pole(G, orc):
    # the reference's cfg/cart_pole/ac_tc.yaml: end_stop_penalty 0 (the cart rides the end stop, episodes end by
    # timeout only); and the task's default end_stop_penalty 1 (absorbing end, reward -10000) with the same agent
    _replay_ac(G, n=32, steps=450, mem=262144, which="cart_pole", test_interval=1, randomization=1.0)
    _replay_ac(G, n=32, steps=300, mem=4099, which="cart_pole", test_interval=3, end_stop_penalty=1, sigma=15.0)


def test_end_stop_penalty_off(G, orc):
    """task/cart_pole/swingup end_stop_penalty = 0 (cart_pole.cpp:177,196): beyond the end stop the episode goes on and
    the reward is the plain potential; with 1 the observation is absorbing and the reward carries -10000."""
    rng = np.random.default_rng(5)
    m = 2000
    st = np.zeros((m, 5))
    st[:, 0] = rng.uniform(-2.6, 2.6, m)
    st[:, 1] = rng.uniform(-7, 7, m)
    st[:, 2:4] = rng.uniform(-3, 3, (m, 2))
    st[:, 4] = rng.uniform(0, 10.2, m)
    act = rng.uniform(-15, 15, m)
    dp = C.POINTER(C.c_double)
    seen = set()
    for esp in (0, 1):
        cfg, oc = G.cart_pole_ac_config(memory=1024, end_stop_penalty=esp), O.cart_pole_ac_config(memory=1024, end_stop_penalty=esp)
        b = G.Batch(cfg, 1)
        try:
            nxt, obs, rew, term = b.model_step(st, act)
        finally:
            b.close()
        for k in range(m):
            on, oo, orw, ot = np.empty(7), np.empty(8), C.c_double(), C.c_int32()
            s = np.ascontiguousarray(st[k])
            orc.orc_env_step(C.byref(oc), s.ctypes.data_as(dp), C.c_double(act[k]), on.ctypes.data_as(dp),
                             oo.ctypes.data_as(dp), C.byref(orw), C.byref(ot))
            assert _rel(nxt[k], on[:5]) <= RTOL and term[k] == ot.value, (esp, k)
            assert abs(rew[k] - orw.value) <= RTOL * max(abs(orw.value), 1e-300), (esp, k)
            seen.add((esp, int(term[k])))
        out = np.abs(nxt[:, 0]) > 2.4
        assert out.any()
        assert np.all(term[out] == 2) if esp else np.all(term[out] != 2)
        assert np.all(rew[out] < -9000) if esp else np.all(rew[out] > -1000)
    assert seen == {(0, 0), (0, 1), (1, 0), (1, 1), (1, 2)}
    with pytest.raises(G.GrlcuError):
        G.Batch(G.cart_pole_ac_config(memory=1024, end_stop_penalty=2), 1)


def test_ac_free_running_statistics(G, orc):
    """Seeded Philox on both sides; libm differences make trajectories diverge, so compare populations."""
    n, steps = 256, 3000
    cfg = G.cart_double_pole_ac_config(memory=131072, test_interval=-1)
    oc = O.cart_double_pole_ac_config(memory=131072, test_interval=-1)
    b = G.Batch(cfg, n)
    try:
        b.run(steps)
        r = b.results()
    finally:
        b.close()
    L = O.lib()
    eps, rets = [], []
    for i in range(64):
        a = L.orc_create(C.byref(oc), i)
        L.orc_run(a, steps)
        s, e, lr = C.c_uint64(), C.c_uint64(), C.c_double()
        L.orc_stats(a, C.byref(s), C.byref(e), None, C.byref(lr), None)
        eps.append(e.value); rets.append(lr.value)
        L.orc_destroy(a)
    assert np.all(r["steps"] == steps)
    # episodes finished within 3000 steps and last returns: same population on both sides
    assert abs(np.mean(r["episodes"]) - np.mean(eps)) < 4 * (np.std(eps) / np.sqrt(64) + np.std(r["episodes"]) / np.sqrt(n)) + 1
    assert abs(np.mean(r["last_return"]) - np.mean(rets)) < 4 * (np.std(rets) / np.sqrt(64) + np.std(r["last_return"]) / np.sqrt(n)) + 1


def test_actor_table_views(G, orc):
    """grlcu_read_actor / grlcu_update_actor: ActionPolicy's representation (limits = action_min/max, limit = 1)."""
    cfg = G.cart_double_pole_ac_config(memory=2048)
    b = G.Batch(cfg, 2)
    try:
        idx = np.array([5, 17, 5, 900, 2047, 33, 1, 0, 12, 13, 14, 15, 16, 18, 19, 20], dtype=np.uint64)
        t = b.get_actor_table(1)
        want = 0.0
        for i in idx:
            want += t[int(i)]
        assert b.read(1, idx, actor=True) == min(max(want / 16, -20.0), 20.0)
        # update: sequential += delta, every written weight clamped (linear.cpp:207-211); duplicates accumulate
        b.update(1, idx, 15.0, actor=True)
        t2 = t.copy()
        for i in idx:
            t2[int(i)] = min(max(t2[int(i)] + 15.0, -20.0), 20.0)
        assert np.array_equal(b.get_actor_table(1), t2)
        assert t2[5] == 20.0                                   # 0..1 + 15 + 15 saturates
        assert b.read(1, idx, actor=True) <= 20.0
        assert np.array_equal(b.get_actor_table(0), b.get_actor_table(0)) and not np.array_equal(b.get_actor_table(0), t2)
        # the critic views are not limited
        c = b.get_table(1)
        b.update(1, idx[:4], 1e6)
        assert b.get_table(1)[17] == c[17] + 1e6
        with pytest.raises(G.GrlcuError):
            b.read(1, np.array([2048], dtype=np.uint64), actor=True)
    finally:
        b.close()
    s = G.Batch(G.pendulum_config(memory=2048), 1)
    try:
        with pytest.raises(G.GrlcuError):
            s.read(0, idx, actor=True)                         # no actor table on a SARSA context
    finally:
        s.close()
