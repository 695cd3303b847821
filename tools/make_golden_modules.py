#!/usr/bin/env python3
"""Generate tests/golden/modules_*.npz: outputs of the reference's own coupler modules (sponge_layer, compute/apply GCM
forcing with hole filling, perturb_temperature; compiled in place into oracle/_ref, serial backend) on the seeded cases of
tests/test_modules_oracle.py. Runs only where /root/reference exists; the fixtures travel with the repository.

usage: python tools/make_golden_modules.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers as H                       # noqa: E402
import test_modules_oracle as T           # noqa: E402


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    for name, kw in sorted(T.CASES.items()):
        case, dm, opts = H.make_modules_case(**kw)
        ids = np.arange(case["nens"], dtype=np.int32) * 3 + 1
        ref = T.run_reference_sequence(case, dm, opts, ids)
        ref["ids"] = ids
        path = os.path.join(out_dir, "modules_%s.npz" % name)
        np.savez_compressed(path, **ref)
        print("%-24s %7.1f kB" % (os.path.basename(path), os.path.getsize(path) / 1e3))


if __name__ == "__main__":
    main()
