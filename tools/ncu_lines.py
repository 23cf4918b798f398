#!/usr/bin/env python
"""Aggregate an `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass` dump into stall-sample shares per
source file, per function region and per line (inlined code is attributed to the callee's lines).  Used to write
the summaries under profiles/.   usage: ncu_lines.py dump.csv [top_n]"""
import csv, sys, collections, re
path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
files = collections.OrderedDict()  # file -> {line: [src, samples, instr]}
cur = None; seen_kernel = None; col_s = col_i = None
for r in rows:
    if not r: continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        if seen_kernel is None: seen_kernel = r[1]
        elif r[1] != seen_kernel: cur = None
        continue
    if r[0] == "Line No":
        col_s = r.index("# Samples"); col_i = r.index("Instructions Executed"); continue
    if cur is None or not r[0].isdigit(): continue
    d = files.setdefault(cur, {})
    try:
        e = d.setdefault(int(r[0]), [r[1], 0, 0]); e[1] += int(r[col_s] or 0); e[2] += int(r[col_i] or 0)
    except ValueError:
        pass
tot_s = sum(e[1] for d in files.values() for e in d.values()); tot_i = sum(e[2] for d in files.values() for e in d.values())
print(f"kernel {seen_kernel}: {tot_s} stall samples, {tot_i} warp instructions (summed over captured launches)")
print("-- by file / function region (samples%, instructions%)")
allreg = []
for f, d in files.items():
    reg = collections.OrderedDict(); name = "(top)"
    for ln in sorted(d):
        if re.match(r"\s*(?:static\s+)?(?:__device__|__global__)", d[ln][0]):
            c = [x for x in re.findall(r"(\w+)\s*\(", d[ln][0]) if x != "__launch_bounds__"]
            if c: name = c[0]
        a = reg.setdefault(name, [0, 0]); a[0] += d[ln][1]; a[1] += d[ln][2]
    allreg += [(f, k, s, n) for k, (s, n) in reg.items()]
for f, k, s, n in sorted(allreg, key=lambda t: -t[2])[:top]:
    print(f"{f:14s} {k:22s} {100*s/max(tot_s,1):6.2f}% {100*n/max(tot_i,1):6.2f}%")
print(f"-- top {top} lines")
L = [(f, ln, e) for f, d in files.items() for ln, e in d.items()]
for f, ln, (src, s, n) in sorted(L, key=lambda t: -t[2][1])[:top]:
    print(f"{f:14s}:{ln:4d} {100*s/max(tot_s,1):5.2f}% {100*n/max(tot_i,1):5.2f}%  {src.strip()[:100]}")
