#!/usr/bin/env python3
"""Summarise `nvcc -Xptxas -v` output: kernel (demangled, shortened) -> registers, spills, stack."""
import re, subprocess, sys
txt = open(sys.argv[1]).read()
cur = None
rows = []
for line in txt.splitlines():
    m = re.search(r"Compiling entry function '(\S+)'", line)
    if m:
        cur = dict(name=m.group(1), regs=None, spill=0, stack=0); rows.append(cur); continue
    if cur is None: continue
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
    if m: cur["stack"] = max(cur["stack"], int(m.group(1))); cur["spill"] = max(cur["spill"], int(m.group(2)) + int(m.group(3)))
    m = re.search(r"Used (\d+) registers", line)
    if m: cur["regs"] = int(m.group(1))
names = subprocess.run(["c++filt"], input="\n".join(r["name"] for r in rows), capture_output=True, text=True).stdout.splitlines()
for r, n in zip(rows, names):
    n = re.sub(r"\(.*", "", n).replace("void awfl::", "")
    print("%-60s regs=%3s spill=%5d stack=%5d" % (n, r["regs"], r["spill"], r["stack"]))
