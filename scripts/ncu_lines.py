#!/usr/bin/env python
"""Per-source-line share of executed warp instructions of one kernel: joins the SASS page of an .ncu-rep with the line
table of the cubin (nvdisasm -g).  usage: ncu_lines.py <src_sass.csv> <nvdisasm.txt> <section regex> [min_pct]"""
import collections, csv, re, sys
sass, dis, sect = sys.argv[1:4]
minp = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
lines = open(dis).read().splitlines()
start = next(i for i, l in enumerate(lines) if re.match(r'\s*\.section\s+\.text\.' + sect, l))
cur, seq = None, []
for l in lines[start + 1:]:
    if re.match(r'\s*\.section\s+\.text\.', l): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m: seq.append((int(m.group(1), 16), cur))
off2 = dict(seq)
rows = list(csv.reader(open(sass)))
hdr = rows[1]; ci = hdr.index("Instructions Executed"); cs = hdr.index("# Samples")
base = None; agg = collections.Counter(); smp = collections.Counter(); tot = tots = 0
for r in rows[2:]:
    if r[0] == "Kernel Name": break
    if not r[0].startswith("0x"): continue
    a = int(r[0], 16)
    if base is None: base = a
    k = off2.get(a - base); agg[k] += int(r[ci]); smp[k] += int(r[cs]); tot += int(r[ci]); tots += int(r[cs])
src = {}
print("total warp instructions", tot, "samples", tots)
for k in sorted(agg, key=lambda x: (x is None, x)):
    if agg[k] / tot * 100 < minp: continue
    t = ''
    if k:
        try:
            if k[0] not in src: src[k[0]] = open('brisk_b200/csrc/' + k[0]).read().splitlines()
            t = src[k[0]][k[1] - 1].strip()[:95]
        except OSError: pass
    print("%5.1f%% inst %5.1f%% samp  %s:%s  %s" % (100 * agg[k] / tot, 100 * smp[k] / tots, k[0] if k else None, k[1] if k else '', t))
