#!/usr/bin/env python3
"""One-line summary of bench.py JSON lines: value, ms/step, per-kernel-class ms, roofline fraction."""
import json, sys
for p in sys.argv[1:]:
    try:
        d = json.loads(open(p).read().strip().splitlines()[-1])
    except Exception as e:
        print(p, "unreadable:", e); continue
    k = {n: round(v["ms_per_step"], 2) for n, v in (d.get("kernels") or {}).items()}
    print("%s: %.1f M/s  %.2f ms/step  frac=%.3f (%s)  e2e=%s  %s" % (
        p.split("/")[-1], d["value"] / 1e6, d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel"],
        ("%.1f" % (d["e2e"]["value"] / 1e6)) if d.get("e2e") else None, k))
