"""The oracle's restatement of the environment arithmetic against the REFERENCE'S OWN sources.

oracle/_ref/libgrlref_env.so is /root/reference/base/src/environments/{pendulum,cart_pole,cart_double_pole}.cpp
compiled where they lie (oracle/Makefile target `ref`) against stand-in headers (oracle/ref_shim/): Eigen 3 is
absent, so grl::Vector and Eigen::Matrix3d/Vector3d are stand-ins; every formula in eom / start / actuate / observe /
evaluate is the reference's own code.  Bit-exact agreement on 1e5 random (state, action) pairs per function pins rows
a2-a4 and a6-a8 of SURVEY.md section 8 to first-party reference code; what stays an assumption is Eigen's summation
order inside Matrix3d::inverse() and the 3x3 products (one function).

Also: the oracle's cart-pole equations of motion + 5-state RK4 reproduce all 10 rows of the reference-held vector
tests/template/cart_pole_balancing-pid-0.txt (grld -s 1 tests/cart_pole_balancing-pid.yaml).
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import _oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
REFSO = os.path.join(O.ORACLE_DIR, "_ref", "libgrlref_env.so")
N = 100000
dp, i32p = C.POINTER(C.c_double), C.POINTER(C.c_int32)


@pytest.fixture(scope="module")
def ref():
    if os.path.isdir("/root/reference/base/src"):
        subprocess.run(["make", "-s", "-C", O.ORACLE_DIR, "ref"], check=True)
    if not os.path.exists(REFSO):
        pytest.skip("oracle/_ref/libgrlref_env.so was not built (no /root/reference here)")
    L = C.CDLL(REFSO)
    L.ref_configure.argtypes = [C.c_double, C.c_double, C.c_int]
    L.ref_eom.argtypes = [C.c_int, dp, C.c_double, dp]
    L.ref_start.argtypes = [C.c_int, C.c_int, C.c_double, dp]
    L.ref_actuate.argtypes = [C.c_int, dp, C.c_double]
    L.ref_actuate.restype = C.c_double
    L.ref_observe.argtypes = [C.c_int, dp, dp, i32p, i32p]
    L.ref_evaluate.argtypes = [C.c_int, dp, C.c_double, dp]
    L.ref_evaluate.restype = C.c_double
    return L


def _states(dyn, rng, n):
    S = (3, 5, 7)[dyn]
    x = rng.uniform(-4, 4, size=(n, S))
    if dyn == 0:
        x[:, 0] = rng.uniform(-25, 25, n); x[:, 1] = rng.uniform(-45, 45, n); x[:, 2] = rng.uniform(0, 3.2, n)
    elif dyn == 1:
        x[:, 0] = rng.uniform(-2.7, 2.7, n); x[:, 1] = rng.uniform(-12, 12, n)
        x[:, 2] = rng.uniform(-6, 6, n); x[:, 3] = rng.uniform(-16, 16, n); x[:, 4] = rng.uniform(0, 10.3, n)
    else:
        x[:, 0] = rng.uniform(-2.7, 2.7, n); x[:, 1] = rng.uniform(-0.9, 0.9, n); x[:, 2] = rng.uniform(-0.9, 0.9, n)
        x[:, 3:6] = rng.uniform(-8, 8, size=(n, 3)); x[:, 6] = rng.uniform(0, 5.2, n)
    # a slice exactly on the end stops / thresholds the tasks test
    x[: n // 50, 0] = np.where(rng.uniform(size=n // 50) < 0.5, 2.4, -2.4) if dyn else x[: n // 50, 0]
    return np.ascontiguousarray(x)


@pytest.mark.parametrize("dyn", [0, 1, 2])
@pytest.mark.parametrize("esp", [1, 0])
def test_environment_arithmetic_equals_reference_sources(orc, ref, dyn, esp):
    if esp == 0 and dyn != 1:
        pytest.skip("end_stop_penalty only exists on task/cart_pole/swingup")
    cfgf = (O.pendulum_config, O.cart_pole_config, O.cart_double_pole_config)[dyn]
    oc = cfgf(memory=1024, randomization=1.0, end_stop_penalty=esp)
    assert ref.ref_configure(oc.timeout, oc.randomization, esp) == 0
    orc.orc_eom.argtypes = [C.c_int, dp, C.c_double, dp]
    orc.orc_observe.argtypes = [C.POINTER(O.Config), dp, dp, i32p]
    orc.orc_evaluate.argtypes = [C.POINTER(O.Config), dp, C.c_double, dp]
    orc.orc_evaluate.restype = C.c_double
    orc.orc_task_start.argtypes = [C.POINTER(O.Config), C.c_int, C.c_double, dp]
    rng = np.random.default_rng(100 + dyn)
    S, OBS = (3, 5, 7)[dyn], (2, 4, 6)[dyn]
    amax = (3.0, 15.0, 20.0)[dyn]
    x, nx = _states(dyn, rng, N), _states(dyn, rng, N)
    act = rng.choice(np.array([-amax, 0.0, amax, 1.7 * amax, -0.31 * amax]), N)
    u = rng.uniform(0, 1, N)
    a_xd, b_xd = np.empty(8), np.empty(8)
    a_o, b_o = np.zeros(8), np.zeros(8)
    a_s, b_s = np.zeros(8), np.zeros(8)
    ta, tb, ab = C.c_int32(), C.c_int32(), C.c_int32()
    for k in range(N):
        px, pn = x[k].ctypes.data_as(dp), nx[k].ctypes.data_as(dp)
        # Task::actuate then Dynamics::eom (modeled.cpp:160-170, 254-276)
        actuation = ref.ref_actuate(dyn, px, float(act[k]))
        want_act = min(max(act[k], -3.0), 3.0) if dyn == 0 else act[k]      # orc_env_step's actuation
        assert actuation == want_act
        assert ref.ref_eom(dyn, px, actuation, b_xd.ctypes.data_as(dp)) == 0
        orc.orc_eom(dyn, px, actuation, a_xd.ctypes.data_as(dp))
        assert np.array_equal(a_xd[:S], b_xd[:S]), (dyn, k, a_xd[:S], b_xd[:S])
        # Task::observe
        assert ref.ref_observe(dyn, px, b_o.ctypes.data_as(dp), C.byref(tb), C.byref(ab)) == OBS
        orc.orc_observe(C.byref(oc), px, a_o.ctypes.data_as(dp), C.byref(ta))
        assert ta.value == tb.value and np.array_equal(a_o[:OBS], b_o[:OBS]), (dyn, k)
        assert ab.value == (1 if tb.value == 2 else 0)
        # Task::evaluate
        ra = orc.orc_evaluate(C.byref(oc), px, float(act[k]), pn)
        rb = ref.ref_evaluate(dyn, px, float(act[k]), pn)
        assert ra == rb, (dyn, k, ra, rb)
        # Task::start, learning and test episodes
        if k < 2000:
            for test in (0, 1):
                assert ref.ref_start(dyn, test, float(u[k]), b_s.ctypes.data_as(dp)) == S
                orc.orc_task_start(C.byref(oc), test, float(u[k]), a_s.ctypes.data_as(dp))
                assert np.array_equal(a_s[:S], b_s[:S]), (dyn, k, test)


def test_cart_pole_balancing_pid_golden(orc):
    """eom_cart_pole + RK4 over 5 states against the reference-held vector (10 rows x 200 steps, 6 significant digits)."""
    golden = [l.split() for l in open(os.path.join(HERE, "golden", "ref_cart_pole_balancing-pid-0.txt")) if l.strip()]
    assert len(golden) == 10
    orc.orc_pin_cart_pole_pid.argtypes = [C.c_long, dp, C.c_double, C.c_int, dp, i32p]
    p = np.array([-10.0, -50.0, -6.0, -10.0])          # tests/cart_pole_balancing-pid.yaml
    rew, steps = np.empty(10), np.empty(10, np.int32)
    orc.orc_pin_cart_pole_pid(1, p.ctypes.data_as(dp), 9.99, 10, rew.ctypes.data_as(dp), steps.ctypes.data_as(i32p))
    # online_learning.cpp:243 (test_interval 0): trial column tt+1-(tt+1)/1 = 0, steps column ss = 0 (test steps do not count)
    rows = [["0", "0", "%g" % r] for r in rew]
    assert rows == golden
    assert np.all(steps == 200)
