import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def rel_l2(a, b, floor=0.0):
    """||a-b||_2 / max(||b||_2, floor)"""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    d = np.linalg.norm(a - b)
    n = max(np.linalg.norm(b), floor)
    return 0.0 if d == 0.0 else d / n


def field_floors(ref_fields):
    """Per-field floors for the relative L2 norm: winds are normalised by at least 1e-3*||uvel||
    (vvel is identically 0 in 2-D, wvel ~ 0 in balanced states); tracers by 1e-30 (may be identically 0)."""
    un = float(np.linalg.norm(ref_fields["uvel"]))
    fl = {k: 1e-300 for k in ref_fields}
    for k in ("uvel", "vvel", "wvel"):
        fl[k] = 1e-3 * un
    return fl
