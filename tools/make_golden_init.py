#!/usr/bin/env python3
"""Generate tests/golden/init_*.npz: the idealized initial states produced by the compiled reference (oracle/_ref:
Dycore::init_supercell resp. the thermal branch of Dycore::init, then convert_dynamics_to_coupler) for the cases of
tests/test_init_oracle.py, plus the reference's 9-point GLL tables. Runs only where /root/reference exists.

usage: python tools/make_golden_init.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import test_init_oracle as T              # noqa: E402
from oracle import reflib                 # noqa: E402


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    r = reflib.RefDycore(8, 1, 8, 1, 16000., 2000., np.arange(9) * 100.)
    pts, wts = r.gll9()
    r.close()
    for name, kw in sorted(T.CASES.items()):
        case = T.build(kw)
        ref = T.reference_init(case, kw["kind"])
        ref["gll_pts"], ref["gll_wts"] = pts, wts
        path = os.path.join(out_dir, "init_%s.npz" % name)
        np.savez_compressed(path, **ref)
        print("%-28s %7.1f kB" % (os.path.basename(path), os.path.getsize(path) / 1e3))


if __name__ == "__main__":
    main()
