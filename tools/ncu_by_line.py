#!/usr/bin/env python
"""Aggregate an ncu SASS source page per CUDA source line.

usage: ncu_by_line.py <report.ncu-rep> <lib.so> [top]

ncu's `--page source --csv` lists SASS instructions with their counters but not their source lines;
nvdisasm -g on the cubin inside the .so has the line table.  Both list the kernel's instructions in the
same order at 16-byte strides, so they are joined by offset.  The line table is taken from the ONE .text section
whose name is exactly the mangled name of the profiled instantiation (read from the report): the library holds
a dozen instantiations of each kernel template, and joining any other section attributes counters to lines
the kernel never executes (round-1 summaries did).
"""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def mangled_name(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--print-kernel-base", "mangled"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    names = {r[rows[0].index("Kernel Name")] for r in rows[2:] if len(r) == len(rows[0])}
    assert len(names) == 1, "expected one profiled kernel in the report, found %s" % sorted(names)
    return names.pop()


def line_table(so, mangled):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, check=True, stdout=subprocess.DEVNULL)
    cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
    table, cur, on, matched = {}, ("?", 0), False, 0
    for ln in dis.splitlines():
        if ln.startswith(".text."):
            on = ln.strip().rstrip(":") == ".text." + mangled
            matched += on
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*)", ln)
        if m:
            table[int(m.group(1), 16)] = (cur, m.group(2).strip())
    assert matched == 1, "%d .text sections named %s in %s" % (matched, mangled, so)
    return table


def main():
    rep, so = sys.argv[1:3]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    kernel = mangled_name(rep)
    print("kernel (mangled):", kernel)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    table = line_table(so, kernel)
    base = int(rows[2][0], 16)
    keys = ["# Samples", "Instructions Executed", "L2 Theoretical Sectors Global", "L2 Theoretical Sectors Global Ideal",
            "L2 Theoretical Sectors Local", "stall_long_sb"]
    agg = defaultdict(lambda: [0.0] * len(keys))
    tot = [0.0] * len(keys)
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        off = int(r[0], 16) - base
        loc = table.get(off, (("?", 0), ""))[0]
        for k, name in enumerate(keys):
            try:
                v = float(r[col[name]])
            except (ValueError, KeyError):
                v = 0.0
            agg[loc][k] += v
            tot[k] += v
    print("totals:", {k: t for k, t in zip(keys, tot)})
    for sortkey in (0, 1, 2):
        print("\n== top lines by %s" % keys[sortkey])
        print("%-28s" % "file:line" + "".join("%14s" % k[-13:] for k in keys))
        for loc, v in sorted(agg.items(), key=lambda kv: -kv[1][sortkey])[:top]:
            print("%-28s" % ("%s:%d" % loc) + "".join("%14.4g" % x for x in v))


if __name__ == "__main__":
    main()
