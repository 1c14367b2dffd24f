#!/usr/bin/env python
"""Static SASS size of one kernel by source region (10-line buckets): what makes a kernel outgrow the I-cache.
usage: sass_size_by_line.py <lib.so> <mangled-kernel-substring> [top]"""
import collections, os, re, subprocess, sys, tempfile
so, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, check=True, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
on, cur, cnt = False, None, collections.Counter()
for ln in dis.splitlines():
    if ln.startswith(".text."):
        on = kern in ln
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]+\*/", ln):
        cnt[cur] += 1
tot = sum(cnt.values())
print("total %d instructions, %.1f KB" % (tot, tot * 16 / 1024))
agg = collections.Counter()
for (f, l), c in cnt.items():
    agg[(f, l // 10 * 10)] += c
for (f, l), c in agg.most_common(top):
    print("%-22s %4d-%-4d %6d  %6.1f KB" % (f, l, l + 9, c, c * 16 / 1024))
