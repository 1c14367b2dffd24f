#!/usr/bin/env python
"""Golden vectors from the reference's OWN classes (oracle/_ref/libgrlref_agent.so: the reference's agent / model / task
sources compiled from /root/reference against a stand-in Eigen), for machines without /root/reference.

For each configuration: the closed loop refag_run (tasks, DynamicalModel::step, agents: all the reference's) from srand48(seed)
for N agent-steps; recorded: learning steps, trials, the environment state (bit patterns) and the SHA-256 of the weight tables'
bytes.  tests/test_oracle_golden.py::test_oracle_reproduces_reference_class_vectors replays the same seed through the oracle
(glibc RNG mode) and must arrive at the same bits.  Run here (where the library can be built): python tools/make_ref_agent_golden.py
"""
import ctypes as C
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle as O  # noqa: E402

CASES = [
    # name, config function, overrides, seed, steps
    ("pendulum_sarsa", "pendulum_config", dict(test_interval=10, randomization=1.0), 1, 8000),
    ("pendulum_sarsa_shipped_memory", "pendulum_config", dict(memory=8388608, test_interval=10, randomization=0.0), 1, 3000),
    ("cart_pole_sarsa", "cart_pole_config", dict(test_interval=10, randomization=1.0), 2, 8000),
    ("cart_double_pole_sarsa", "cart_double_pole_config", dict(test_interval=10, randomization=1.0), 3, 6000),
    ("cart_double_pole_ac", "cart_double_pole_ac_config", dict(test_interval=10), 4, 8000),
    ("pendulum_ac", "pendulum_ac_config", dict(test_interval=10), 5, 6000),
    ("cart_pole_ac", "cart_pole_ac_config", dict(test_interval=10), 6, 6000),
    ("pendulum_q_learning", "pendulum_config", dict(test_interval=10, randomization=1.0, algo=O.ALGO_Q), 7, 5000),
    ("pendulum_expected_sarsa", "pendulum_config", dict(test_interval=10, randomization=1.0, algo=O.ALGO_EXPECTED_SARSA,
                                                        target_epsilon=0.05), 8, 5000),
    ("pendulum_softmax", "pendulum_config", dict(test_interval=10, randomization=1.0, sampler=O.SAMPLER_SOFTMAX,
                                                 softmax_tau=0.5), 9, 5000),
    ("pendulum_safe2", "pendulum_config", dict(memory=65536, test_interval=10, randomization=1.0, safe=2), 10, 4000),
]


def main():
    lib = os.path.join(ROOT, "oracle", "_ref", "libgrlref_agent.so")
    R = C.CDLL(lib)
    dp = C.POINTER(C.c_double)
    R.refag_create.restype = C.c_void_p
    R.refag_create.argtypes = [C.POINTER(O.Config), C.c_long]
    R.refag_run.argtypes = [C.c_void_p, C.c_long, dp, dp]
    R.refag_table.argtypes = [C.c_void_p, C.c_int, dp]
    R.refag_destroy.argtypes = [C.c_void_p]
    R.refag_prepare()
    L = O.lib()
    out = {"what": __doc__.strip().split("\n\n")[0], "cases": []}
    for name, fn, kw, seed, steps in CASES:
        cfg = getattr(O, fn)(**kw)
        cfg.rng = O.RNG_GLIBC
        h = R.refag_create(C.byref(cfg), seed)
        assert h, name
        res, st = (C.c_double * 3)(), (C.c_double * 8)()
        assert R.refag_run(h, steps, res, st) == 0
        S = L.orc_state_dims(cfg.dynamics)
        tab = np.empty(cfg.memory)
        R.refag_table(h, 0, tab.ctypes.data_as(dp))
        case = {"name": name, "config": fn, "overrides": {k: v for k, v in kw.items()}, "seed": seed, "steps": steps,
                "learning_steps": int(res[0]), "trials": int(res[1]),
                "state_bits": ["%016x" % b for b in np.array(list(st)[:S]).view(np.uint64)],
                "table_sha256": hashlib.sha256(tab.tobytes()).hexdigest()}
        if cfg.algo == O.ALGO_AC:
            R.refag_table(h, 1, tab.ctypes.data_as(dp))
            case["actor_table_sha256"] = hashlib.sha256(tab.tobytes()).hexdigest()
        R.refag_destroy(h)
        out["cases"].append(case)
        print(name, case["learning_steps"], case["trials"], case["table_sha256"][:12])
    with open(os.path.join(ROOT, "tests", "golden", "ref_classes_closed_loop.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
