#!/usr/bin/env python
"""Executed warp instructions of one kernel by CODE REGION of its own source file: joins the SASS page of an .ncu-rep
(ncu -i X.ncu-rep --page source --csv --kernel-name regex:K > sass.csv) with the cubin's line table
(cuobjdump -xelf all lib.so; nvdisasm -g -c X.cubin > dis.txt).  Inlined helper code (other files) is attributed to the
last line of <file> seen before it in SASS order, which is its call site up to instruction scheduling.
usage: ncu_regions.py <sass.csv> <dis.txt> <section regex> <file> <units> <first:last:name> [...]
  units = divisor for the per-unit column (e.g. warp-steps of the launch)
example: ncu_regions.py sass.csv dis.txt _Z16enumerate_kernelILi4ELb0E enumerate.cuh 8.19e6 0:185:setup 186:253:step 254:317:rescan"""
import collections, csv, re, sys
sass, dis, sect, fname, units = sys.argv[1:6]
units = float(units)
regions = [(int(a), int(b), n) for a, b, n in (x.split(":") for x in sys.argv[6:])]
lines = open(dis).read().splitlines()
start = next(i for i, l in enumerate(lines) if re.match(r'\s*\.section\s+\.text\.' + sect, l))
seq, cur, last = [], None, None
for l in lines[start + 1:]:
    if re.match(r'\s*\.section\s+\.text\.', l): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        if m.group(1).split('/')[-1] == fname: last = int(m.group(2))
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
for a, le, op in seq:
    name = next((n for lo, hi, n in regions if le is not None and lo <= le <= hi), "other")
    agg[name] += ex.get(a, 0); ops[name][op] += ex.get(a, 0)
print("total warp instructions %d (%.1f per unit)" % (tot, tot / units))
for name, v in agg.most_common():
    print("%-12s %5.1f %%  %7.1f per unit   top: %s" % (name, 100.0 * v / tot, v / units,
          ", ".join("%s %.1f" % (o, c / units) for o, c in ops[name].most_common(8))))
