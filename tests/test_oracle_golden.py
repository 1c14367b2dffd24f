"""Pin the oracle against the reference's own golden vector for this path.

tests/golden/ref_pendulum-sarsa-tc-0.txt is the reference's seeded learning curve
(grld -s 1, tests/pendulum-sarsa-tc.yaml; bin/runtests.py byte-compares it).  Every row
depends on the whole hot path: RK4 pendulum dynamics, reward, tile hash, Q read,
epsilon-greedy draws (glibc drand48/lrand48 in construction order), SARSA(lambda) write
and replacing-trace sweep over 2000 trials.  The oracle in glibc-RNG mode reproduces all
181 rows to the 6 significant digits the file holds.
"""
import ctypes as C
import os

import _oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_pendulum-sarsa-tc-0.txt")


def test_pendulum_sarsa_tc_golden_curve(orc):
    golden = [l.split() for l in open(GOLD) if l.strip()]
    assert len(golden) == 181
    c = O.pendulum_config(memory=8388608, rng=O.RNG_GLIBC, test_interval=10)
    orc.orc_glibc_seed(1)                       # grld -s 1 -> srand48(1), deployer.cpp:70-74
    a = orc.orc_create(C.byref(c), 0)
    rec = O.Record()
    rows, trial = [], 0
    try:
        while len(rows) < len(golden):
            orc.orc_step(a, C.byref(rec))
            if rec.episode_done:
                trial += 1
                if rec.test:
                    st = C.c_uint64()
                    orc.orc_stats(a, C.byref(st), None, None, None, None)
                    # online_learning.cpp:243: tt+1-(tt+1)/(test_interval+1), ss, total_reward
                    rows.append((str(trial - trial // 11), str(st.value), "%g" % rec.episode_return))
    finally:
        orc.orc_destroy(a)
    assert rows == [tuple(g) for g in golden]
