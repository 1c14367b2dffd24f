"""The trace value cache (grlcu_cached.cuh) and the tile-major physical table layout (struct Geo, grlcu_layout.h)
are internal optimisations of the fused kernel: they must leave tables (in the reference's logical order), traces,
counters and returns bit-identical to the uncached path on the reference's raw hash order, which the replay tests
pin to the oracle step by step.  Both paths run the same device dynamics and Philox streams, so every
bit has to agree, across launch boundaries, test episodes, absorbing ends and index collisions."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(G, mk, mode, n, launches, sweep=False, layout=1, **kw):
    os.environ["GRLCU_MODE"] = str(mode)
    os.environ["GRLCU_LAYOUT"] = str(layout)
    try:
        b = G.Batch(mk(**kw), n)
    finally:
        os.environ.pop("GRLCU_MODE", None)
        os.environ.pop("GRLCU_LAYOUT", None)
    try:
        if sweep:
            # per-agent hyperparameters (MultiExperiment / grlo sweeps): trace lengths differ inside a warp
            i = np.arange(n)
            lam = np.array([0.0, 0.2, 0.4, 0.5, 0.65, 0.75, 0.85, 0.9])[i % 8]
            b.set_hyper(alpha=np.array([0.05, 0.1, 0.2, 0.4])[(i // 8) % 4], lambda_=lam,
                        epsilon=np.array([0.01, 0.05, 0.1, 0.3])[(i // 32) % 4])
        for k in launches:
            b.run(k)
        res = b.results()
        tables = np.stack([b.get_table(i) for i in range(n)])
        if b.cfg.algo == G.capi.ALGO_AC:
            tables = np.concatenate([tables, np.stack([b.get_actor_table(i) for i in range(n)])], axis=1)
        states = [b.get_state(i) for i in range(n)]
    finally:
        b.close()
    return res, tables, states


@pytest.mark.parametrize("case", [
    dict(which="pendulum", n=96, memory=65536, launches=[37, 100, 63, 1, 250], test_interval=3),
    dict(which="pendulum", n=64, memory=257, launches=[150, 75], test_interval=-1),      # heavy collisions
    # tiles of the box / table entries = 0.76: a third of the tile words share an earlier tile's slot (tile-major layout)
    dict(which="pendulum", n=64, memory=32768, launches=[100, 200], test_interval=3),
    dict(which="pendulum", n=16, memory=8388608, launches=[300], test_interval=-1),      # the shipped table size
    dict(which="pendulum", n=64, memory=4099, launches=[220], test_interval=2),          # non power of two
    dict(which="cart_pole", n=64, memory=262144, launches=[180, 333], test_interval=4, epsilon=0.2),
    dict(which="cart_pole", n=32, memory=1031, launches=[400], test_interval=-1, epsilon=0.5),
    dict(which="pendulum", n=32, memory=8192, launches=[130], test_interval=-1, lambda_=0.0),  # decay < 0.01
    dict(which="pendulum", n=32, memory=8192, launches=[260], test_interval=-1, lambda_=0.9),  # long trace
    # trace capacity 35 (> 32 entries: the duplicate mask is 64 bits wide) with per-agent hyperparameters
    dict(which="pendulum", n=128, memory=65536, launches=[120, 260], test_interval=5, lambda_=0.9, sweep=True),
    dict(which="pendulum", n=64, memory=521, launches=[300], test_interval=-1, lambda_=0.9, sweep=True),
    # Q-learning / expected SARSA targets through the cache
    dict(which="pendulum", n=64, memory=16384, launches=[150, 150], test_interval=3, algo=2),
    dict(which="cart_pole", n=32, memory=65536, launches=[300], test_interval=-1, algo=3, target_epsilon=0.1),
    # actor-critic (BASELINE configs[3]): the critic's trace goes through the cache, the actor table does not
    dict(which="ac", n=64, memory=131072, launches=[90, 1, 310], test_interval=3),
    dict(which="ac", n=32, memory=509, launches=[350], test_interval=-1, actor_alpha=0.5, sigma=15.0),
    # sampler/softmax as the learning policy's sampler
    dict(which="pendulum", n=64, memory=65536, launches=[140, 160], test_interval=3, sampler=1, softmax_tau=0.5),
    # SARSA on the cart-double-pole (cfg/cart_double_pole/sarsa.yaml): 51-entry ring
    dict(which="cart_double_pole", n=64, memory=131072, launches=[150, 1, 250], test_interval=3, epsilon=0.1),
    dict(which="cart_double_pole", n=32, memory=1021, launches=[300], test_interval=-1),
    # the same agent on the pendulum and on the cart-pole (cfg/pendulum/ac_tc.yaml, cfg/cart_pole/ac_tc.yaml)
    dict(which="pendulum_ac", n=64, memory=65536, launches=[120, 230], test_interval=2, randomization=1.0),
    dict(which="pendulum_ac", n=32, memory=263, launches=[300], test_interval=-1, randomization=1.0, sigma=3.0, actor_alpha=0.3),
    dict(which="cart_pole_ac", n=64, memory=262144, launches=[250, 200], test_interval=1, randomization=1.0),
    dict(which="cart_pole_ac", n=32, memory=2053, launches=[420], test_interval=-1, end_stop_penalty=1, sigma=15.0),
])
def test_cached_equals_uncached(case):
    import grl_b200 as G
    case = dict(case)
    which = case.pop("which")
    mk = {"pendulum": G.pendulum_config, "cart_pole": G.cart_pole_config, "ac": G.cart_double_pole_ac_config,
          "cart_double_pole": G.cart_double_pole_config,
          "pendulum_ac": G.pendulum_ac_config, "cart_pole_ac": G.cart_pole_ac_config}[which]
    n, launches = case.pop("n"), case.pop("launches")
    if which in ("pendulum", "cart_pole"):
        case["randomization"] = 1.0
    # the reference's literal form: uncached, raw hash order
    r1, t1, s1 = _run(G, mk, 1, n, launches, layout=0, **case)
    # mode 0: the value cache; layout 1: tile-major physical layout where the context is eligible (the pendulum cases
    # with tiles <= memory; everywhere else layout 1 is the raw order again and the variant repeats an earlier one)
    variants = [dict(mode=0, layout=1), dict(mode=0, layout=0), dict(mode=1, layout=1)]
    for v in variants:
        r0, t0, s0 = _run(G, mk, v["mode"], n, launches, layout=v["layout"], **case)
        assert not np.isnan(t0).any(), v                   # no directory sentinel left in a table
        assert np.array_equal(t0, t1), v
        for k in r0:
            if k != "table_writes":               # roofline accounting: differs between the paths by design
                assert np.array_equal(r0[k], r1[k]), (v, k)
        for a, b in zip(s0, s1):
            assert np.array_equal(a[0], b[0]) and a[1:] == b[1:], v
        assert int(r0["steps"].sum()) > 0 and int(r0["swept"].sum()) >= 0
