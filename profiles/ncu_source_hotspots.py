#!/usr/bin/env python
"""Aggregates `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass` per CUDA source line.
usage: ncu_source_hotspots.py src_cs.csv [top_n]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = collections.OrderedDict()
fname, cur, hdr = None, None, None
seen = set()  # (source line, SASS address): the page can list a function twice
for r in rows:
    if not r:
        continue
    if r[0] in ("File Name", "File Path"):
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; ie = r.index("Instructions Executed"); ws = r.index("Warp Stall Sampling (All Samples)"); continue
    if hdr is None or len(r) <= ie:
        continue
    if r[0] != "":
        cur = (fname, int(r[0]), r[1].strip()[:100]); agg.setdefault(cur, [0, 0, collections.Counter()]); continue
    if cur is None or r[2] == "..." or (cur, r[2]) in seen:
        continue
    seen.add((cur, r[2]))
    try:
        n, s = int(r[ie] or 0), int(r[ws] or 0)
    except ValueError:
        continue
    agg[cur][0] += n; agg[cur][1] += s
    for name in ("stall_long_sb", "stall_short_sb", "stall_barrier", "stall_wait", "stall_math", "stall_branch_resolving", "stall_no_inst", "stall_mio", "stall_lg"):
        try:
            agg[cur][2][name] += int(r[hdr.index(name)] or 0)
        except ValueError:
            pass
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print(f"total warp-instructions {ti}  stall samples {ts}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][int(__import__("os").environ.get("BY_INST","0")) ^ 1])[:top]:
    st = ",".join(f"{n[6:]}={c}" for n, c in v[2].most_common(3) if c)
    print(f"{v[0]/max(ti,1)*100:5.1f}%inst {v[1]/max(ts,1)*100:5.1f}%smp {k[0]}:{k[1]:<4} {k[2][:70]:70s} [{st}]")
