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
    with open(os.path.join(ROOT, "tests", "golden", "grld_ref_curves.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))


if __name__ == "__main__":
    main()
