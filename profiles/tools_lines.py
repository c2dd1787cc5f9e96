"""Executed warp instructions per source line for one kernel of an ncu report.
    python profiles/tools_lines.py <report.ncu-rep> <kernel regex> <nvdisasm -g -c listing of the same build> [launch index]
The SASS page of the report carries the counts, the nvdisasm listing the file:line of every instruction; both list
the kernel's instructions in the same order."""
import csv, re, subprocess, sys
from collections import defaultdict

rep, kern, dis = sys.argv[1:4]
which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(out))
his = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hi = his[which]
end = his[which + 1] - 1 if which + 1 < len(his) else len(rows)
h = rows[hi]
ci, ti = h.index("Instructions Executed"), h.index("Thread Instructions Executed")
data = [r for r in rows[hi + 1:end] if r and r[0].startswith("0x")]
name = rows[hi - 1][1] if hi else ""
lines = open(dis).read().splitlines()
mangled = [i for i, l in enumerate(lines) if l.startswith(".text.") and re.search(kern, l)]
start = mangled[0]
cur, tags, stack = None, [], []
for l in lines[start + 1:]:
    if l.startswith(".text.") or l.startswith("//---"):
        if tags: break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        tags.append(cur)
assert len(tags) >= len(data), (len(tags), len(data))
agg = defaultdict(lambda: [0, 0])
tot = 0
for r, tg in zip(data, tags):
    c, t = int(float(r[ci] or 0)), int(float(r[ti] or 0))
    agg[tg][0] += c; agg[tg][1] += t; tot += c
print(name[:100], "sass", len(data), "warp inst", tot)
byfile = defaultdict(int)
for (f, ln), (c, t) in ((k, v) for k, v in agg.items() if k): byfile[f] += c
print(dict(byfile))
for tg, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:60]:
    if c: print(f"{100 * c / tot:5.1f}%  lanes {t / c:4.1f}  {tg}")
# coarse view: per file, buckets of 25 source lines
buck = defaultdict(lambda: [0, 0])
for tg, (c, t) in agg.items():
    if tg: buck[(tg[0], tg[1] // 25 * 25)][0] += c; buck[(tg[0], tg[1] // 25 * 25)][1] += t
print("-- buckets of 25 lines")
for k, (c, t) in sorted(buck.items()):
    if c > 0.01 * tot: print(f"{100 * c / tot:5.1f}%  lanes {t / c:4.1f}  {k[0]}:{k[1]}-{k[1] + 24}")
