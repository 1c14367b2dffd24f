#!/usr/bin/env python
"""Record the DRAM traffic of one launch of the run kernel from ncu summaries (tools/ncu_summary.py output) in
profiles/traffic.json, keyed by workload:agents:memory, together with the sha of the kernel sources the capture was
taken from (bench.py only turns the figure into `traffic_frac` when that sha is the current tree's).

usage: update_traffic.py <summary.txt> <workload> <agents> <memory> [<summary.txt> <workload> <agents> <memory> ...]
"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (kernel_sources_sha, STEPS_PER_PASS)


def parse(path):
    text = open(path).read()
    rd = float(re.search(r"dram__bytes_read\.sum\s+(\S+)\s+(\S+)", text).group(2))
    wr = float(re.search(r"dram__bytes_write\.sum\s+(\S+)\s+(\S+)", text).group(2))
    unit = re.search(r"dram__bytes_read\.sum\s+(\S+)", text).group(1)
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[unit]
    grid = int(re.search(r"grid \((\d+),", text).group(1))
    block = int(re.search(r"block \((\d+),", text).group(1))
    return (rd + wr) * scale, grid * block


def main():
    out = os.path.join(ROOT, "profiles", "traffic.json")
    data = {"what": "dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the run kernel from an `ncu --set full` "
                    "capture, per agent-step of that launch (launch = grid*block agents x 100 steps), keyed by "
                    "workload:agents:memory; kernel_sources_sha ties the capture to the tree it was taken from"}
    sha = bench.kernel_sources_sha()
    args = sys.argv[1:]
    for i in range(0, len(args), 4):
        path, workload, agents, memory = args[i], args[i + 1], int(args[i + 2]), int(args[i + 3])
        total, launch_agents = parse(path)
        data["%s:%d:%d" % (workload, agents, memory)] = {
            "dram_bytes_per_launch": total, "agents_in_launch": launch_agents,
            "agent_steps_per_launch": launch_agents * bench.STEPS_PER_PASS,
            "dram_bytes_per_agent_step": total / (launch_agents * bench.STEPS_PER_PASS),
            "capture": os.path.relpath(path, ROOT), "kernel_sources_sha": sha}
    with open(out, "w") as f:
        json.dump(data, f, indent=1)
    print(json.dumps(data, indent=1))


if __name__ == "__main__":
    main()
