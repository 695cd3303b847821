"""The entry point that allocates nothing: awfl_compute_mass_dynamics_noalloc (include/awfl_b200.h) against the existing
awfl_compute_mass_dynamics, through the C ABI, on two calls in a row with different inputs (the static workspace keeps the
first call's contents: the second result must not see them). No counterpart in the reference; the file name makes pytest
collect it last."""
import ctypes as C

import numpy as np
import pytest

import helpers as H
from pam_b200 import _lib, synthetic as syn

pytestmark = pytest.mark.gpu
TOL = 1e-13          # the tolerance of the haloed compute_mass checks (tests/test_cxx_shim.py); same kernels -> in fact equal bits


def _haloed(lib, d, c, nz, ny, nx, nens, nt, hs, torch):
    """convert_coupler_to_dynamics through the C ABI: haloed (state, tracers) device arrays of the coupler's fields."""
    shape = (nz + 2 * hs, ny + 2 * hs, nx + 2 * hs, nens)
    st = torch.zeros((5,) + shape, dtype=torch.float64, device=c.device)
    tr = torch.zeros((nt,) + shape, dtype=torch.float64, device=c.device)
    f, keep = d._fields(c)
    assert lib.awfl_convert_coupler_to_dynamics(d.h, C.byref(f), st.data_ptr(), tr.data_ptr()) == 0, lib.awfl_last_error(d.h)
    return st, tr


@pytest.mark.parametrize("ny,nt", [(1, 1), (4, 10)])
def test_no_alloc_mass_matches_the_handle_entry_point_on_two_calls_in_a_row(ny, nt):
    import torch
    lib = _lib.load()
    ord, nz, nx, nens = 5, 12, 10, 40
    hs = (ord + 1) // 2
    tracers = syn.TRACERS_NT1 if nt == 1 else syn.TRACERS_P3_SHOC
    case = syn.make_case(nz, ny, nx, nens, tracers=tracers, grid="tanh", per_ens_grid=True)
    c, d = H.make_gpu(case, ord, crm_dt=2.0)
    zint = np.asarray(case["zint"], dtype=np.float64)
    dz = torch.from_numpy(np.ascontiguousarray(np.diff(zint, axis=0))).to(c.device)          # (nz, nens)
    assert dz.shape == (nz, nens)
    assert lib.awfl_noalloc_setup() == 0, lib.awfl_last_error(None)
    n_out = (nt + 2) * nens
    results = []
    for call in range(2):
        if call == 1:
            d.declare_current_profile_as_hydrostatic(c)
            d.timeStep(c)                                   # different fields for the second call
        st, tr = _haloed(lib, d, c, nz, ny, nx, nens, nt, hs, torch)
        want = torch.full((n_out,), float("nan"), dtype=torch.float64, device=c.device)
        got = torch.full((n_out,), float("nan"), dtype=torch.float64, device=c.device)
        assert lib.awfl_compute_mass_dynamics(d.h, st.data_ptr(), tr.data_ptr(), want.data_ptr()) == 0, lib.awfl_last_error(d.h)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.awfl_compute_mass_dynamics_noalloc(nz, ny, nx, nens, nt, ord, st.data_ptr(), tr.data_ptr(), dz.data_ptr(),
                                                    got.data_ptr(), None)
        assert rc == 0, lib.awfl_last_error(None)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before                    # nothing was allocated on the device
        w, g = want.cpu().numpy(), got.cpu().numpy()
        assert np.all(np.isfinite(w)) and np.all(np.isfinite(g))
        assert np.max(np.abs(g - w) / (np.abs(w) + 1e-300)) <= TOL
        results.append(w)
    assert not np.array_equal(results[0], results[1])                         # the two calls really had different inputs


def test_no_alloc_rejects_sizes_beyond_the_static_workspace():
    import torch
    lib = _lib.load()
    cap = int(lib.awfl_noalloc_max_doubles())
    assert cap == 1 << 21
    x = torch.zeros(8, dtype=torch.float64, device="cuda:0")      # never touched: the call must fail before any launch
    nz, ny, nx, nens, nt = 64, 1, 128, 64, 1                       # 6 * 524288 doubles > 2^21
    assert (5 + nt) * nz * ny * nx * nens > cap
    rc = lib.awfl_compute_mass_dynamics_noalloc(nz, ny, nx, nens, nt, 5, x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), None)
    assert rc != 0
    msg = lib.awfl_last_error(None).decode()
    assert "exceeds the static workspace" in msg and str(cap) in msg
    rc = lib.awfl_compute_mass_dynamics_noalloc(0, 1, 8, 8, 1, 5, x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), None)
    assert rc != 0 and "bad dimensions" in lib.awfl_last_error(None).decode()
    torch.cuda.synchronize()
