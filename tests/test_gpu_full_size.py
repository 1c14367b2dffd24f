"""BASELINE.json's full sizes (one GPU's share of each config) through size-independent properties: results are
keyed by the global agent id, so any slice of a full-size batch must equal, bit for bit, a small batch created
with that slice's first_agent_id (this is also what makes the multi-GPU sharding exact), tables and traces
included; the trace value cache must leave no directory sentinel behind; step and episode counters must add up.
A 32 GiB table array is 2^35 bytes: every 32-bit slip in the index arithmetic shows up here and nowhere else."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _full(G, mk, n, memory, steps, slices, hyper=None, **kw):
    import torch
    free, _ = torch.cuda.mem_get_info(0)
    tables = 2 if mk(memory=16).algo == G.capi.ALGO_AC else 1
    need = n * memory * 8 * tables * 1.08 + (4 << 30)
    if free < need:
        pytest.skip("needs %.0f GiB of free HBM" % (need / 2**30))
    cfg = mk(memory=memory, **kw)
    big = G.Batch(cfg, n)
    try:
        if hyper is not None:
            big.set_hyper(*[h(np.arange(n)) for h in hyper])
        for k in steps:
            big.run(k)
        res = big.results()
        total = sum(steps)
        assert np.all(res["steps"] <= total)
        if kw.get("test_interval", -1) < 0:
            assert int(res["steps"].sum()) == n * total                       # every loop iteration is a learning step
        assert np.all(res["episodes"] >= 1)
        picked = {}
        for first in slices:
            picked[first] = ([big.get_table(first + i) for i in (0, 63)],
                             [big.get_actor_table(first + i) for i in (0, 63)] if tables == 2 else None,
                             [big.get_state(first + i) for i in (0, 63)])
    finally:
        big.close()
    for first in slices:
        small = G.Batch(cfg, 64, first_agent_id=first)
        try:
            if hyper is not None:
                small.set_hyper(*[h(np.arange(first, first + 64)) for h in hyper])
            for k in steps:
                small.run(k)
            r = small.results()
            for key in r:
                assert np.array_equal(r[key], res[key][first:first + 64]), (first, key)
            for j, i in enumerate((0, 63)):
                t = small.get_table(i)
                assert not np.isnan(t).any()                                 # no directory sentinel left behind
                assert np.array_equal(t, picked[first][0][j]), (first, i)
                if tables == 2:
                    assert np.array_equal(small.get_actor_table(i), picked[first][1][j]), (first, i)
                a, b = small.get_state(i), picked[first][2][j]
                assert np.array_equal(a[0], b[0]) and a[1:] == b[1:]
        finally:
            small.close()


def test_pendulum_65536_agents_x_65536_entries():
    import grl_b200 as G
    _full(G, G.pendulum_config, 65536, 65536, [150, 1, 99], [0, 32768 - 32, 65536 - 64], randomization=1.0, test_interval=-1)


def test_cart_pole_32768_agents_x_262144_entries():
    import grl_b200 as G
    _full(G, G.cart_pole_config, 32768, 262144, [260], [0, 20000, 32768 - 64], randomization=1.0, test_interval=9)


def test_actor_critic_65536_agents_x_2_x_131072_entries():
    import grl_b200 as G
    _full(G, G.cart_double_pole_ac_config, 65536, 131072, [180], [0, 65536 - 64], test_interval=10)


def test_sibling_configurations_at_full_size():
    """The reference's other tile-coded configs on these tasks at the bench sizes: actor-critic on the cart-pole
    (2 x 64 GiB of tables: indices beyond 2^33 entries) and SARSA on the cart-double-pole (51-entry trace ring)."""
    import grl_b200 as G
    _full(G, G.cart_pole_ac_config, 32768, 262144, [230], [0, 32768 - 64], randomization=1.0, test_interval=4)
    _full(G, G.cart_double_pole_config, 65536, 131072, [160, 60], [0, 40000 - 40000 % 64, 65536 - 64], test_interval=-1)


def test_sweep_131072_agents_x_32768_entries():
    import grl_b200 as G
    lam = np.array([0.0, 0.2, 0.4, 0.5, 0.65, 0.75, 0.85, 0.9])
    hyper = (lambda g: np.array([0.05, 0.1, 0.2, 0.4])[g % 4], lambda g: lam[(g // 4) % 8],
             lambda g: np.array([0.01, 0.02, 0.05, 0.1])[(g // 32) % 4])
    _full(G, G.pendulum_config, 131072, 32768, [220], [0, 77777 - 77777 % 64, 131072 - 64], hyper=hyper,
          randomization=1.0, test_interval=-1, lambda_=0.9)
