#!/usr/bin/env python
"""Executed warp instructions of one kernel by (source file, line): joins the SASS page of an .ncu-rep with the cubin's
line table like ncu_regions.py, but keeps the file of every line (inlined helpers stay attributed to their own file).
usage: ncu_filelines.py <sass.csv> <dis.txt> <section regex> <units> [top]"""
import collections, csv, re, sys
sass, dis, sect, units = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
units = float(units)
lines = open(dis).read().splitlines()
start = next(i for i, l in enumerate(lines) if re.match(r'\s*\.section\s+\.text\.' + sect, l))
seq, last = [], None
for l in lines[start + 1:]:
    if re.match(r'\s*\.section\s+\.text\.', l): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        last = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        t = m.group(2).split()
        seq.append((int(m.group(1), 16), last, (t[1] if t[0].startswith('@') else t[0]).split('.')[0]))
rows = list(csv.reader(open(sass)))
ci = rows[1].index("Instructions Executed")
base, ex = None, {}
for r in rows[2:]:
    if not r[0].startswith("0x"): continue
    a = int(r[0], 16)
    if base is None: base = a
    ex[a - base] = int(r[ci])
tot = sum(ex.values())
agg, ops = collections.Counter(), collections.defaultdict(collections.Counter)
byfile = collections.Counter()
for a, le, op in seq:
    agg[le] += ex.get(a, 0); ops[le][op] += ex.get(a, 0)
    byfile[le[0] if le else None] += ex.get(a, 0)
print("total warp instructions %d (%.1f per unit)" % (tot, tot / units))
for f, v in byfile.most_common():
    print("file %-40s %5.1f %%  %7.1f per unit" % (f, 100.0 * v / tot, v / units))
for le, v in agg.most_common(top):
    print("%-28s %5.1f %%  %6.1f per unit   %s" % ("%s:%d" % le if le else "?", 100.0 * v / tot, v / units,
          ", ".join("%s %.1f" % (o, c / units) for o, c in ops[le].most_common(5))))
