#!/usr/bin/env python3
"""Aggregate an `ncu --page source --csv --print-source sass,cuda` export by CUDA source line.
usage: ncu_by_line.py export.csv [top_n]   (first captured launch only)"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg, samples, src = collections.Counter(), collections.Counter(), {}
cur, hdr, seen = None, None, set()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        if cur in seen and hdr is not None and cur == first:
            break
        if not seen:
            first = cur
        seen.add(cur)
        hdr = None
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or r[2] != "-":
        continue
    try:
        ln = int(r[0])
        v = float(r[hdr.index("Instructions Executed")] or 0)
        s = float(r[hdr.index("# Samples")] or 0)
    except ValueError:
        continue
    agg[(cur, ln)] += v
    samples[(cur, ln)] += s
    src[(cur, ln)] = r[1].strip()[:100]
tot, ts = sum(agg.values()), sum(samples.values())
print("total warp instructions %d, stall samples %d" % (tot, ts))
for k, v in agg.most_common(top):
    print("%-20s %4d inst=%5.2f%% samp=%5.2f%% | %s" % (k[0], k[1], 100 * v / tot, 100 * samples[k] / max(ts, 1), src[k]))
