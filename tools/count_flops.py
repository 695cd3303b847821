#!/usr/bin/env python3
"""Builds tools/flopcount/count_flops.cpp for each reconstruction order and prints the instrumented operation count of one
Dycore::reconstruct -- the numbers frozen in pam_b200/roofline.py (RECON_OPS). `python tools/count_flops.py [out.json]`"""
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def count(order, tmpdir):
    exe = os.path.join(tmpdir, "count%d" % order)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    libdir = os.path.join(ROOT, "pam_b200")
    subprocess.check_call([cxx, "-std=c++17", "-O1", "-DMW_ORD=%d" % order, "-I", os.path.join(ROOT, "pam_b200", "cxx"),
                           os.path.join(ROOT, "tools", "flopcount", "count_flops.cpp"), "-o", exe, "-L", libdir, "-lawfl_b200",
                           "-Wl,-rpath," + libdir])
    return json.loads(subprocess.check_output([exe]).decode())


def main():
    sys.path.insert(0, ROOT)
    from pam_b200 import roofline
    out = {}
    with tempfile.TemporaryDirectory() as td:
        for o in (3, 5, 7, 9):
            c = count(o, td)
            hand = roofline.recon_flops_hand_formula(o)
            c["hand_formula"] = {"flops": hand[0], "div": hand[1], "delta_rel": (c["flops"] - hand[0]) / hand[0]}
            c["frozen_in_roofline_py"] = list(roofline.recon_flops(o))
            out[str(o)] = c
    txt = json.dumps(out, indent=1)
    print(txt)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(txt + "\n")


if __name__ == "__main__":
    main()
