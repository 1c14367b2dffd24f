#!/usr/bin/env python
"""Learning curves written by the reference's own deployer, as committed golden vectors.

oracle/_ref/grld_ref is grld built from the reference's translation units.  For every case of
tests/test_oracle_vs_grld_ref.py this script runs `grld_ref -s <seed> cfg.yaml` (the YAML comes from tests/_oracle.grld_yaml) and
stores the rows it wrote -- trial, learning steps so far, episode return as printed (3 decimals), episode time -- in
tests/golden/grld_ref_curves.json, so that the oracle can be held against the reference's output on machines that have neither
/root/reference nor the binary (tests/test_oracle_golden.py::test_oracle_reproduces_reference_deployer_curves).
Run where the binary can be built: python tools/make_grld_ref_golden.py
"""
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle as O  # noqa: E402
from test_oracle_vs_grld_ref import CASES, GRLD  # noqa: E402


# The reference's SHIPPED configuration files, run as they are (visualisation stripped -- no graphics here --, `trials` and `output`
# set), against the oracle's transcription of the same file (tests/_oracle.py: <fn>(memory=<the file's memory>)).
SHIPPED = [
    # file under /root/reference/cfg, oracle config function, extra overrides, trials
    ("pendulum/sarsa_tc.yaml", "pendulum_config", {}, 330),
    ("cart_pole/sarsa.yaml", "cart_pole_config", {}, 220),
    ("cart_double_pole/sarsa.yaml", "cart_double_pole_config", {}, 330),
    ("pendulum/ac_tc.yaml", "pendulum_ac_config", {}, 110),
    ("cart_pole/ac_tc.yaml", "cart_pole_ac_config", {}, 110),
    ("pendulum/q_tc.yaml", "pendulum_config", {"algo": O.ALGO_Q}, 110),
]


def shipped_cases(cfg_root="/root/reference/cfg"):
    import yaml

    def strip(n):
        if isinstance(n, dict):
            for k in list(n.keys()):
                if k in ("visualizer", "visualization"):
                    del n[k]
                else:
                    strip(n[k])

    cases = []
    for rel, fn, kw, trials in SHIPPED:
        d = yaml.safe_load(open(os.path.join(cfg_root, rel)))
        for k in list(d.keys()):
            if k != "experiment":
                del d[k]
        strip(d)
        d["experiment"]["trials"] = trials
        d["experiment"]["output"] = "curve"
        memory = d["experiment"]["agent"]["policy"]["projector"]["memory"]
        with tempfile.TemporaryDirectory() as t:
            with open(os.path.join(t, "cfg.yaml"), "w") as f:
                yaml.safe_dump(d, f, default_flow_style=None, sort_keys=False)
            subprocess.run([GRLD, "-s", "1", "cfg.yaml"], cwd=t, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            rows = [ln.split()[:4] for ln in open(os.path.join(t, "curve-0.txt")) if ln.strip()]
        cases.append({"name": "cfg/" + rel, "config": fn, "overrides": dict(kw, memory=memory,
                                                                             test_interval=d["experiment"]["test_interval"]),
                      "seed": 1, "trials": trials, "rows": rows})
        print("cfg/" + rel, memory, len(rows), rows[-1])
    return cases


def main():
    out = {"what": __doc__.strip().split("\n\n")[0], "cases": []}
    for name, fn, kw, seed, trials in CASES:
        cfg = getattr(O, fn)(**kw)
        with tempfile.TemporaryDirectory() as d:
            with open(os.path.join(d, "cfg.yaml"), "w") as f:
                f.write(O.grld_yaml(cfg, trials, "curve"))
            subprocess.run([GRLD, "-s", str(seed), "cfg.yaml"], cwd=d, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            rows = [ln.split()[:4] for ln in open(os.path.join(d, "curve-0.txt")) if ln.strip()]
        out["cases"].append({"name": name, "config": fn, "overrides": kw, "seed": seed, "trials": trials, "rows": rows})
        print(name, len(rows), rows[-1])
    out["cases"] += shipped_cases()
    with open(os.path.join(ROOT, "tests", "golden", "grld_ref_curves.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))


if __name__ == "__main__":
    main()
