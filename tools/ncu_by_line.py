#!/usr/bin/env python3
"""Aggregate an `ncu -i x.ncu-rep --page source --csv --print-source sass,cuda` export by CUDA source line.
usage: ncu_by_line.py export.csv [top_n] [kernel-name substring]   (default: the kernel with most instructions)"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
want = sys.argv[3] if len(sys.argv) > 3 else None
agg = collections.defaultdict(collections.Counter)       # kernel -> (file, line) -> warp instructions
samples = collections.defaultdict(collections.Counter)
src = {}
cur, fn, hdr = None, None, None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur, hdr = r[1].split("/")[-1], None
        continue
    if r[0] == "Function Name":
        fn = r[1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < 3 or r[2] != "-":
        continue
    try:
        ln = int(r[0])
        v = float(r[hdr.index("Instructions Executed")] or 0)
        s = float(r[hdr.index("# Samples")] or 0)
    except ValueError:
        continue
    agg[fn][(cur, ln)] += v
    samples[fn][(cur, ln)] += s
    src[(cur, ln)] = r[1].strip()[:100]
names = [k for k in agg if want is None or want in k]
name = max(names, key=lambda k: sum(agg[k].values()))
a, sm = agg[name], samples[name]
tot, ts = sum(a.values()), sum(sm.values())
print("%s\ntotal warp instructions %d, stall samples %d" % (name[:100], tot, ts))
by_file = collections.Counter()
for (f, _), v in a.items():
    by_file[f] += v
print("by file: " + ", ".join("%s %.1f%%" % (f, 100 * v / tot) for f, v in by_file.most_common()))
for k, v in a.most_common(top):
    print("%-18s %4d inst=%5.2f%% samp=%5.2f%% | %s" % (k[0], k[1], 100 * v / tot, 100 * sm[k] / max(ts, 1), src[k]))
