"""The TIMED kernel against the oracle, agent by agent.

bench.py times k_run<..., CACHED> (its own episode loop, Philox counter schedule and trace value cache); the replay
tests drive a different kernel (k_replay) from the oracle's loop.  Here the fused kernel runs free (grlcu_run) and the
oracle runs free (orc_run) on the same global agent ids and the same Philox key, and every agent is compared with
itself: learning steps, episodes and swept trace indices (V) must be EQUAL, returns within 1e-9 relative, final tables
within 1e-9 relative (north_star's bar for Q-values; the typical difference is a few ulp, from device sin/cos vs glibc).
An agent may legitimately leave the oracle's trajectory when a last-bit difference flips a decision; every such agent
is re-run alone, step by step, and the test fails unless its first diverging decision was a Q tie / near-tie or an
observation within 1e-9 of a tile boundary (online_learning.cpp:154-222, agents/td.cpp:50-81).
"""
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


def _oracle_agent(L, oc, gid, steps):
    a = L.orc_create(C.byref(oc), gid)
    L.orc_run_total(a, steps)
    s, e, w = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lr, ltr = C.c_double(), C.c_double()
    L.orc_stats(a, C.byref(s), C.byref(e), C.byref(w), C.byref(lr), C.byref(ltr))
    tbl = np.ctypeslib.as_array(L.orc_table(a), shape=(oc.memory,)).copy()
    act = None
    if oc.algo == O.ALGO_AC:
        act = np.ctypeslib.as_array(L.orc_actor_table(a), shape=(oc.memory,)).copy()
    L.orc_destroy(a)
    return (s.value, e.value, w.value, lr.value, ltr.value), tbl, act


def _close(a, b):
    return np.all(np.abs(a - b) <= RTOL * np.maximum(np.abs(b), 1e-300))


def _explain_divergence(G, gcfg_f, ocfg_f, kw, gid, steps):
    """Re-run agent `gid` alone on both sides one step at a time; return a description of the first diverging decision,
    or None when the divergence has no legitimate cause."""
    cfg, oc = gcfg_f(**kw), ocfg_f(**kw)
    L = O.lib()
    b = G.Batch(cfg, 1, first_agent_id=gid)
    a = L.orc_create(C.byref(oc), gid)
    rec = O.Record()
    S = b.S
    scaling = np.array([cfg.tilings / cfg.resolution[i] for i in range(cfg.obs_dims)])
    try:
        for counted in range(1, steps + 1):
            L.orc_step(a, C.byref(rec))
            b.run(1)                               # grlcu_run counts test-episode steps too, like this loop
            st, act, _ = b.get_state(0)
            ost = np.array(list(rec.next)[:S])
            oact = float(O.lib().orc_pending_action(a))
            same_state = _close(st, ost)
            if act != oact or not same_state:
                q = np.array(list(rec.q)[:max(cfg.n_actions, 1)])
                obs = np.array(list(rec.obs)[:cfg.obs_dims])
                frac = obs * scaling
                near_tile = float(np.min(np.abs(frac - np.round(frac))))
                qs = np.sort(q)[::-1]
                near_tie = float(abs(qs[0] - qs[1]) / max(abs(qs[0]), 1e-300)) if q.size > 1 else 1.0
                why = dict(step=counted, action_gpu=act, action_oracle=oact, q=q.tolist(), near_tie=near_tie,
                           near_tile=near_tile, state_gpu=st.tolist(), state_oracle=ost.tolist())
                if near_tie < 1e-9 or near_tile < 1e-9:
                    return why
                return dict(why, unexplained=True)
    finally:
        b.close()
        L.orc_destroy(a)
    return dict(step=None, note="no divergence when run alone")


CASES = [
    # name, config pair, kwargs, agents, steps
    ("pendulum", "pendulum_config", dict(memory=65536, randomization=1.0, test_interval=-1), 4096, 300),
    ("pendulum_test_episodes", "pendulum_config", dict(memory=65536, randomization=1.0, test_interval=2), 1024, 300),
    ("cart_pole", "cart_pole_config", dict(memory=262144, randomization=1.0, test_interval=-1), 4096, 300),
    ("cart_double_pole_ac", "cart_double_pole_ac_config", dict(memory=131072, test_interval=-1), 4096, 200),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_fused_kernel_equals_oracle_per_agent(G, orc, case):
    name, cfgname, kw, n, steps = case
    gcfg_f, ocfg_f = getattr(G, cfgname), getattr(O, cfgname)
    cfg, oc = gcfg_f(**kw), ocfg_f(**kw)
    first = 1000                                   # global ids need not start at zero
    orc.orc_pending_action.restype = C.c_double
    orc.orc_pending_action.argtypes = [C.c_void_p]
    orc.orc_run_total.restype = None
    orc.orc_run_total.argtypes = [C.c_void_p, C.c_uint64]
    b = G.Batch(cfg, n, first_agent_id=first)
    try:
        # three launches: the cache stays live across them, and the last one ends mid-episode
        b.run(steps // 3); b.run(steps // 3); b.run(steps - 2 * (steps // 3))
        res = b.results()
        outliers, worst_tbl, ulp_hist = [], 0.0, []
        for i in range(n):
            (s, e, w, lr, ltr), otbl, oact = _oracle_agent(orc, oc, first + i, steps)
            gt = b.get_table(i)
            ok = (int(res["steps"][i]) == s and int(res["episodes"][i]) == e and int(res["swept"][i]) == w
                  and _close(res["last_return"][i:i + 1], np.array([lr]))
                  and _close(res["last_test_return"][i:i + 1], np.array([ltr])) and _close(gt, otbl))
            if ok and oact is not None:
                ok = _close(b.get_actor_table(i), oact)
            if ok:
                d = np.abs(gt - otbl)
                worst_tbl = max(worst_tbl, float(np.max(d / np.maximum(np.abs(otbl), 1e-300))))
                ulp_hist.append(float(np.max(d / np.spacing(np.maximum(np.abs(otbl), 1e-300)))))
            else:
                outliers.append(i)
    finally:
        b.close()
    print("%s: %d/%d agents equal to the oracle; worst table rel. diff %.3g, median max-ulp %.1f; outliers %s"
          % (name, n - len(outliers), n, worst_tbl, float(np.median(ulp_hist)) if ulp_hist else -1, outliers[:8]))
    assert len(outliers) <= n // 100, (name, len(outliers))
    for i in outliers:
        why = _explain_divergence(G, gcfg_f, ocfg_f, kw, first + i, steps)
        print("outlier agent %d: %s" % (first + i, why))
        assert why is not None and not why.get("unexplained"), (name, first + i, why)
