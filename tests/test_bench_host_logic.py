"""bench.py's host-side bookkeeping (no GPU): the committed DRAM-traffic cross-reference and the byte model.

The traffic figure of `roofline.traffic` comes from profiles/traffic.json (tools/update_traffic.py, one `ncu --set full`
capture per workload).  A key mismatch between the writer and the reader once left it null in every bench line without
anyone noticing, hence this test."""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

HEADLINE = [("pendulum", 65536, 65536), ("cart_pole", 32768, 262144),
            ("cart_double_pole_ac", 65536, 131072), ("sweep", 131072, 32768)]


@pytest.mark.parametrize("workload,agents,memory", HEADLINE)
def test_committed_traffic_is_found_for_every_headline_workload(workload, agents, memory):
    with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
        entry = json.load(f)["%s:%d:%d" % (workload, agents, memory)]
    steps = entry["agent_steps_per_launch"]
    traffic, capture, tree, current = bench.committed_traffic(workload, agents, memory, steps)
    # the same launch size as the capture's gives the capture's own byte count back
    assert traffic == pytest.approx(entry["dram_bytes_per_launch"], rel=1e-12)
    assert os.path.exists(os.path.join(ROOT, capture))
    assert tree == entry["kernel_sources_sha"]
    assert current == (tree == bench.kernel_sources_sha())
    # half a launch moves half the bytes
    assert bench.committed_traffic(workload, agents, memory, steps / 2)[0] == pytest.approx(traffic / 2, rel=1e-12)
    # sector-granular traffic is never below the algorithmic bytes of the same launch (V = 0 is the model's minimum)
    assert traffic > bench.table_bytes_per_step(workload, 0.0) * steps


def test_committed_traffic_unknown_workload_is_none():
    assert bench.committed_traffic("pendulum", 3, 5, 100) == (None, None, None, False)


def test_byte_model_matches_survey_8d():
    # SARSA: 8*A*T + 8*T + 8*T + 16*V = 640 + 16 V; actor-critic: 768 + 16 V  
    assert bench.table_bytes_per_step("pendulum", 0.0) == 640
    assert bench.table_bytes_per_step("pendulum", 144.0) == 2944
    assert bench.table_bytes_per_step("cart_pole", 10.0) == 800
    assert bench.table_bytes_per_step("cart_double_pole_ac", 0.0) == 768
    assert bench.table_bytes_per_step("cart_double_pole_ac", 5.0) == 848


def test_reference_deployer_rate_where_it_was_built():
    # bench.py --impl reference / cpu_baseline time oracle/_ref/grld_ref (the reference's own deployer) where it exists
    import argparse
    args = argparse.Namespace(workload="pendulum", agents=4, memory=65536, cohort=0)
    if not os.path.exists(bench.GRLD_REF):
        assert bench.grld_rate(args, 0.2) is None
        return
    r = bench.grld_rate(args, 0.3)
    assert r and r["kind"] == "reference" and r["value"] > 1e3 and r["cores"] >= 1
