"""CPU tests of the oracle's primitives against independent formulations."""
import ctypes as C

import numpy as np

import _oracle as O


def test_hash_matches_independent_murmur2(orc):
    # GRL's createHashSum == Appleby's MurmurHash2 over the little-endian int bytes with
    # seed' = 449 ^ n ^ 4n (tile_coding.h:94 xors the COUNT of ints, the original the byte length).
    rng = np.random.default_rng(3)
    for n in (1, 3, 4, 6, 7, 9):
        for _ in range(200):
            ints = rng.integers(-2**31, 2**31 - 1, size=n, dtype=np.int64).astype(np.int32)
            h = orc.orc_hash(ints.ctypes.data_as(C.POINTER(C.c_int32)), n)
            m = orc.orc_murmur2_bytes(ints.ctypes.data_as(C.c_void_p), 4 * n, (449 ^ n ^ (4 * n)) & 0xFFFFFFFF)
            assert h == m


def test_murmur2_python_restatement(orc):
    def murmur(ints):
        m, h = 0x5bd1e995, 449 ^ len(ints)
        for v in ints:
            k = (int(v) & 0xFFFFFFFF) * m & 0xFFFFFFFF
            k ^= k >> 24
            k = k * m & 0xFFFFFFFF
            h = (h * m & 0xFFFFFFFF) ^ k
        h ^= h >> 13
        h = h * m & 0xFFFFFFFF
        return h ^ (h >> 15)
    for ints in ([0, 0, 0, 0], [1, -1, 5, 15], [-320, 7, -3, 0], [2**31 - 1, -2**31, 0, 3, 2, 1]):
        a = np.array(ints, dtype=np.int32)
        assert orc.orc_hash(a.ctypes.data_as(C.POINTER(C.c_int32)), len(ints)) == murmur(ints)


def test_philox_known_answer(orc):
    # Random123 kat_vectors: philox4x32-10, counter/key all zero and all ones
    out = (C.c_uint32 * 4)()
    orc.orc_philox(0, 0, 0, 0, 0, 0, out)
    assert list(out) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    orc.orc_philox(0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, out)
    assert list(out) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]


def test_projection_wrapping_and_tiling_structure(orc):
    c = O.pendulum_config(memory=1 << 20)
    dp, u32p = C.POINTER(C.c_double), C.POINTER(C.c_uint32)

    def proj(x):
        x = np.array(x, dtype=np.float64)
        h, i = np.empty(16, np.uint32), np.empty(16, np.uint32)
        orc.orc_project(c.resolution, c.wrapping, 3, 16, c.memory, x.ctypes.data_as(dp),
                        h.ctypes.data_as(u32p), i.ctypes.data_as(u32p))
        return h, i
    h0, i0 = proj([0.1, 0.5, 0.0])
    # angle wraps at 6.283: same tiles one period later (wrapping 6.283 * 16/0.31415 = 320 exactly)
    h1, _ = proj([0.1 + 6.283, 0.5, 0.0])
    assert np.array_equal(h0, h1)
    assert np.array_equal(i0, h0 % np.uint32(c.memory))
    # a tiny move keeps most tilings' tiles, a big one none
    h2, _ = proj([0.1 + 0.31415 / 16 * 1.01, 0.5, 0.0])
    assert 1 <= int(np.sum(h0 != h2)) <= 2
    h3, _ = proj([2.0, -3.0, 3.0])
    assert int(np.sum(h0 == h3)) == 0


def test_rk4_pendulum_energy_and_time(orc):
    dp = C.POINTER(C.c_double)
    x = np.array([np.pi, 0.0, 0.0])
    nxt = np.empty(3)
    orc.orc_rk4(0, 0.03, 5, x.ctypes.data_as(dp), 0.0, nxt.ctypes.data_as(dp))
    assert abs(nxt[2] - 0.03) < 1e-15 and nxt[2] != 1.0
    assert abs(nxt[0] - np.pi) < 1e-9 and abs(nxt[1]) < 1e-6     # hanging equilibrium stays put
    x = np.array([0.3, 0.0, 0.0])
    orc.orc_rk4(0, 0.03, 5, x.ctypes.data_as(dp), 3.0, nxt.ctypes.data_as(dp))
    assert nxt[1] > 0            # positive torque + gravity away from upright accelerate positively


def test_trace_length_bound_and_V(orc):
    """Replacing trace with gamma*lambda = 0.6305 holds <= 9 entries; V <= 16*9."""
    c = O.pendulum_config(memory=65536, randomization=1.0, test_interval=-1)
    a = orc.orc_create(C.byref(c), 7)
    rec = O.Record()
    mx_len = mx_v = 0
    for _ in range(600):
        orc.orc_step(a, C.byref(rec))
        mx_len, mx_v = max(mx_len, rec.trace_len), max(mx_v, rec.swept)
    orc.orc_destroy(a)
    assert mx_len == 9 and 0 < mx_v <= 144


def test_actor_critic_oracle_learns_and_respects_limits(orc):
    """cfg 4 (cart-double-pole AC): the oracle's actor-critic balances longer with training, the actor table
    stays inside [action_min, action_max] (linear.cpp:207-211) and test episodes draw no noise."""
    import ctypes as C
    import numpy as np
    import _oracle as O
    c = O.cart_double_pole_ac_config(memory=1 << 20, test_interval=10)
    a = orc.orc_create(C.byref(c), 3)
    rec = O.Record()
    tests, lens, cur = [], [], 0
    try:
        while len(tests) < 60:
            orc.orc_step(a, C.byref(rec))
            cur += 1
            if rec.test:
                assert rec.noise == 0.0 and rec.start_noise == 0.0
            assert -20.0 <= rec.action <= 20.0
            if rec.episode_done:
                if rec.test:
                    tests.append(rec.episode_return)
                lens.append(cur); cur = 0
        actor = np.ctypeslib.as_array(orc.orc_actor_table(a), shape=(1 << 20,))
        assert actor.min() >= -20.0 and actor.max() <= 20.0
        assert np.mean(tests[-15:]) > 1.5 * np.mean(tests[:15])      # 6.2 per balanced step: longer episodes
    finally:
        orc.orc_destroy(a)


def test_sibling_actor_critic_configs_in_the_oracle(orc):
    """cfg/pendulum/ac_tc.yaml and cfg/cart_pole/ac_tc.yaml: actions stay in the task's range; with end_stop_penalty 0
    (cart_pole.cpp:177,196) no episode is absorbing, every episode runs to the timeout (200 steps) and no reward carries
    the -10000 penalty even when the cart sits beyond the end stop; with 1 the same agent sees absorbing ends."""
    import ctypes as C
    import _oracle as O
    rec = O.Record()
    c = O.pendulum_ac_config(memory=1 << 16, test_interval=3)
    a = orc.orc_create(C.byref(c), 1)
    try:
        n_done = 0
        for t in range(1000):
            orc.orc_step(a, C.byref(rec))
            assert -3.0 <= rec.action <= 3.0 and rec.terminal in (0, 1)
            n_done += rec.episode_done
        assert n_done == 10                          # 100 steps per episode (timeout 2.99 at 0.03)
    finally:
        orc.orc_destroy(a)
    seen = {}
    for esp in (0, 1):
        c = O.cart_pole_ac_config(memory=1 << 16, test_interval=-1, end_stop_penalty=esp, sigma=15.0)
        a = orc.orc_create(C.byref(c), 2)
        lens, cur, beyond, terms, rmin = [], 0, 0, set(), 0.0
        try:
            for t in range(4000):
                orc.orc_step(a, C.byref(rec))
                cur += 1
                assert -15.0 <= rec.action <= 15.0
                beyond += abs(rec.next[0]) > 2.4
                terms.add(rec.terminal)
                rmin = min(rmin, rec.reward)
                if rec.episode_done:
                    lens.append(cur); cur = 0
        finally:
            orc.orc_destroy(a)
        seen[esp] = (lens, beyond, terms, rmin)
    lens, beyond, terms, rmin = seen[0]
    assert set(lens) == {200} and beyond > 0 and terms == {0, 1} and rmin > -1000
    lens, beyond, terms, rmin = seen[1]
    assert 2 in terms and min(lens) < 200 and rmin < -9000


def test_softmax_sampler_in_the_oracle(orc):
    """sampler/softmax (softmax.cpp:48-101): the learning policy draws one uniform and picks from the Boltzmann
    distribution of Q/tau; empirical frequencies match an independent numpy restatement, a tiny tau is greedy, test
    episodes stay greedy (sampler/greedy) and the agent still learns."""
    import ctypes as C
    import numpy as np
    import _oracle as O
    rec = O.Record()
    for tau in (0.5, 1e-6):
        c = O.pendulum_config(memory=1 << 16, test_interval=1, randomization=1.0, sampler=O.SAMPLER_SOFTMAX, softmax_tau=tau)
        a = orc.orc_create(C.byref(c), 4)
        try:
            n = hit = 0
            for t in range(3000):
                orc.orc_step(a, C.byref(rec))
                if rec.terminal == 2 or rec.action_index < 0:
                    continue
                q = np.array(rec.q[:3])
                if rec.test:
                    assert rec.action_index == int(np.argmax(q)) or (q == q.max()).sum() > 1
                    continue
                # independent restatement: softmax of q / tau, cumulative pick on u
                z = q / tau
                pr = np.exp(z - z.max()); pr /= pr.sum()
                pick = int(np.searchsorted(np.cumsum(pr), rec.u, side="right"))
                n += 1
                hit += min(pick, 2) == rec.action_index
                if tau < 1e-3 and (q.max() - np.sort(q)[-2]) / tau > 50:
                    assert rec.action_index == int(np.argmax(q))
            assert n > 1000 and hit >= n - 2          # (u within an ulp of a boundary may differ)
        finally:
            orc.orc_destroy(a)


def test_q_learning_and_expected_sarsa_targets(orc):
    """predictor/critic/q (advantage.cpp:84-90) and predictor/critic/expected_sarsa (sarsa.cpp:172-173 ->
    q.cpp:56-68, greedy.cpp:88-98,220-238): the oracle's targets against a direct restatement on its own
    recorded Q(s', .), and Q-learning still learns the pendulum swing-up."""
    import ctypes as C
    import numpy as np
    import _oracle as O

    def value(q, eps):
        mx = max(q)
        man = sum(1 for x in q if x == mx)
        v = 0.0
        for x in q:
            d = 1.0 / man if x == mx else 0.0
            if eps >= 0:
                if d == 1:
                    d = 1 - 1.0 * eps
                d += 1.0 * eps / len(q)
            v += x * d
        return v

    for algo, eps in ((O.ALGO_Q, 0.0), (O.ALGO_EXPECTED_SARSA, -1.0), (O.ALGO_EXPECTED_SARSA, 0.05), (O.ALGO_SARSA, 0.0)):
        c = O.pendulum_config(memory=4096, randomization=1.0, test_interval=-1, algo=algo, target_epsilon=eps)
        a = orc.orc_create(C.byref(c), 11)
        rec = O.Record()
        checked = 0
        try:
            for _ in range(400):
                orc.orc_step(a, C.byref(rec))
                if rec.terminal == 2 or rec.test:
                    continue
                q = [rec.q[i] for i in range(3)]
                if algo == O.ALGO_Q:
                    want = rec.reward + c.gamma * max(q)
                elif algo == O.ALGO_EXPECTED_SARSA:
                    want = rec.reward + c.gamma * value(q, eps)
                else:
                    want = rec.reward + c.gamma * q[rec.action_index]
                assert rec.target == want
                checked += 1
        finally:
            orc.orc_destroy(a)
        assert checked > 300
    # learning: Q-learning reaches the same return band as SARSA on the swing-up
    c = O.pendulum_config(memory=65536, randomization=0.0, test_interval=-1, algo=O.ALGO_Q)
    a = orc.orc_create(C.byref(c), 2)
    try:
        orc.orc_run(a, 100 * 300)
        lr = C.c_double()
        orc.orc_stats(a, None, None, None, C.byref(lr), None)
        assert lr.value > -2500
    finally:
        orc.orc_destroy(a)
