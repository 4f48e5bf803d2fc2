#!/usr/bin/env python
"""Global-memory sectors per CUDA source line, from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name ...`.
usage: ncu_source_sectors.py src.csv [top_n] [file-name filter]
Sectors are 32 bytes; "theoretical" = requested by the instruction, before L1 / L2 hits.  An instruction inlined from another
file is listed under every source line of its inline stack: give the kernel's own file as the filter to count each once."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
only = sys.argv[3] if len(sys.argv) > 3 else ""
agg, cur, hdr, fname = collections.OrderedDict(), None, None, ""
seen = set()  # (source line, SASS address): the page can list a function twice
for r in rows:
    if not r:
        continue
    if r[0] in ("File Name", "File Path"):
        fname = r[1]
        cur = None
        continue
    if r[0] == "Line No":
        hdr = r
        c_sec, c_ideal, c_op, c_space, c_req = (r.index(n) for n in ("L2 Theoretical Sectors Global", "L2 Theoretical Sectors Global Ideal",
                                                                       "Access Operation", "Address Space", "L1 Tag Requests Global"))
        continue
    if hdr is None or len(r) <= c_sec:
        continue
    if r[0] != "":
        cur = (int(r[0]), r[1].strip()[:90]) if only in fname else None
        continue
    if cur is None or "Global" not in r[c_space] or (fname, cur, r[2]) in seen:
        continue
    seen.add((fname, cur, r[2]))
    try:
        sec, ideal, req = int(r[c_sec] or 0), int(r[c_ideal] or 0), int(r[c_req] or 0)
    except ValueError:
        continue
    a = agg.setdefault((cur, r[c_op]), [0, 0, 0])
    a[0] += sec
    a[1] += ideal
    a[2] += req
tot = {}
for (k, op), v in agg.items():
    tot[op] = tot.get(op, 0) + v[0]
print("theoretical global sectors by operation:", {k: f"{v * 32 / 1e9:.3f} GB" for k, v in tot.items()})
for (k, op), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{v[0] * 32 / 1e6:9.1f} MB  ideal {v[1] * 32 / 1e6:9.1f} MB  requests {v[2]:>10d}  {op:6s} :{k[0]:<4} {k[1]}")
