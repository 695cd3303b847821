#!/usr/bin/env python3
"""Per CUDA source line: executed warp instructions of a kernel split into FP64 and other (from an ncu report captured with
--import-source on and code built with -lineinfo). usage: ncu_line_mix.py report.ncu-rep kernel-regex [top N]"""
import csv, io, re, subprocess, sys, collections
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
kern, fname, hdr, cur = None, None, None, (-1, "")
agg = collections.defaultdict(lambda: [0, 0, "", collections.Counter()])
seen = set()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        kern = r[1]; continue
    if r[0] == "Line No":
        hdr = r; ie = hdr.index("Instructions Executed"); continue
    if hdr is None or len(r) <= ie:
        continue
    line, src, addr, sass = r[0], r[1], r[2], r[3]
    if line.isdigit():
        cur = (int(line), src.strip()[:90]); continue
    if (kern, addr) in seen:
        continue
    seen.add((kern, addr))
    try:
        n = int(r[ie])
    except ValueError:
        continue
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", sass)
    op = m.group(2) if m else "?"
    a = agg[(kern, fname, cur[0])]
    if op in ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX"):
        a[0] += n
    else:
        a[1] += n; a[3][op] += n
    a[2] = cur[1]
kerns = sorted({k[0] for k in agg})
for kn in kerns:
    items = [(k, v) for k, v in agg.items() if k[0] == kn]
    f64 = sum(v[0] for _, v in items); oth = sum(v[1] for _, v in items)
    print("== %s\n   FP64 %d  other %d  (FP64 share %.1f%%)" % (kn[:110], f64, oth, 100.0 * f64 / max(f64 + oth, 1)))
    for k, v in sorted(items, key=lambda kv: -kv[1][1])[:top]:
        ops = " ".join("%s:%.1f%%" % (o, 100.0 * c / max(oth, 1)) for o, c in v[3].most_common(4))
        print("   %-18s:%-4d other %5.1f%%  fp64 %5.1f%%  | %s | %s" % (k[1], k[2], 100.0 * v[1] / max(oth, 1), 100.0 * v[0] / max(f64, 1), ops, v[2]))
