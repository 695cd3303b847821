"""CPU tests of the host-side mirror of the reference interface (PamCoupler / DataManager / Options), the ensemble
sharding helper, the synthetic-input generator and the frozen roofline arithmetic."""
import os

import numpy as np
import pytest

from pam_b200 import PamCoupler, roofline, synthetic as syn
from pam_b200.parallel import shard_range


def _coupler():
    c = PamCoupler("cpu")
    c.allocate_coupler_state(6, 1, 4, 3)
    c.set_grid(8000.0, 2000.0, np.arange(7) * 250.0)
    return c


def test_coupler_registers_the_reference_field_names():
    c = _coupler()
    dm = c.get_data_manager_device_readonly()
    for n in ("density_dry", "uvel", "vvel", "wvel", "temp"):                 # pam_coupler.h:259-263
        assert tuple(dm.get(n, 4, readonly=True).shape) == (6, 1, 4, 3)
    assert tuple(dm.get("vertical_cell_dz", readonly=True).shape) == (6, 3)   # pam_coupler.h:265
    assert tuple(dm.get("vertical_interface_height", readonly=True).shape) == (7, 3)
    assert np.allclose(dm.get("vertical_cell_dz", readonly=True).numpy(), 250.0)
    assert np.allclose(dm.get("vertical_midpoint_height", readonly=True).numpy()[:, 0], 125.0 + 250.0 * np.arange(6))
    assert c.get_dx() == 2000.0 and c.get_dy() == 2000.0 and c.get_nens() == 3


def test_datamanager_errors_like_the_reference():
    c = _coupler()
    dm = c.get_data_manager_device_readwrite()
    with pytest.raises(RuntimeError, match="Could not find entry"):           # DataManager.h:553-560
        dm.get("nope")
    with pytest.raises(RuntimeError, match="dimensions"):                     # DataManager.h:569-575
        dm.get("uvel", 3)
    with pytest.raises(RuntimeError, match="Duplicate"):                      # DataManager.h:104-108
        dm.register_and_allocate("uvel", "", (1,))
    assert not dm.dirty["temp"]
    dm.get("temp")
    assert dm.dirty["temp"]                                                   # DataManager.h:306
    with pytest.raises(RuntimeError, match="option key"):                     # Options.h:84-87
        c.get_option("crm_dt")
    c.set_option("crm_dt", 2.0)
    assert c.get_option("crm_dt") == 2.0 and c.option_exists("crm_dt")


def test_tracer_registry():
    c = _coupler()
    c.add_tracer("cloud_water", "cw", True, True)
    c.add_tracer("water_vapor", "wv", True, True)
    c.add_tracer("tke", "", True, False)
    assert c.get_tracer_names() == ["cloud_water", "water_vapor", "tke"]
    assert c.get_tracer_info("tke") == ("", True, True, False)
    assert c.get_tracer_info("absent")[1] is False
    assert c.get_num_tracers() == 3


@pytest.mark.parametrize("nens,world", [(4096, 8), (1024, 3), (7, 4), (5, 8)])
def test_shard_range_partitions_the_ensemble(nens, world):
    spans = [shard_range(nens, r, world) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == nens
    for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
        assert a1 == b0 and a1 >= a0
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(nens, world, world)


def test_synthetic_inputs_are_deterministic_and_shardable():
    a = syn.make_case(12, 1, 6, 6, grid="tanh")
    b = syn.make_case(12, 1, 6, 6, grid="tanh")
    for k in a["fields"]:
        assert np.array_equal(a["fields"][k], b["fields"][k])
    # a shard generated with ens_offset equals the slice of the global case (what bench.py relies on)
    s = syn.make_case(12, 1, 6, 2, grid="tanh", ens_offset=3)
    for k in a["fields"]:
        assert np.array_equal(s["fields"][k], a["fields"][k][..., 3:5]), k
    nt10 = syn.make_case(12, 1, 12, 2, tracers=syn.TRACERS_P3_SHOC)
    assert [t[0] for t in nt10["tracers"]].index("water_vapor") == 8          # idWV = 8
    zero_frac = np.mean(nt10["fields"]["rain"] == 0.0)
    assert zero_frac >= 0.3                                                   # FCT must have something to limit
    for k, v in nt10["fields"].items():
        assert np.isfinite(v).all(), k


def test_vertical_grids():
    z = syn.vertical_grid(58, kind="equal")
    assert z[0] == 0.0 and abs(z[-1] - 20000.0) < 1e-9 and np.allclose(np.diff(z), 20000.0 / 58)
    zt = syn.vertical_grid(58, kind="tanh")
    dz = np.diff(zt)
    assert zt[0] == 0.0 and abs(zt[-1] - 20000.0) < 1e-6 and (dz > 0).all() and dz[-1] > 3 * dz[0]


def test_roofline_counts_are_frozen():
    # (flops, of which divides) of one Dycore::reconstruct: the INSTRUMENTED count (SURVEY.md 8d asks for it; its hand count
    # was 143 / 287 / 840 +-10 %). tools/count_flops.py runs the reference's algorithm on a counting scalar type.
    assert roofline.recon_flops(3) == (149, 16)
    assert roofline.recon_flops(5) == (295, 22)
    assert roofline.recon_flops(7) == (525, 28)
    assert roofline.recon_flops(9) == (845, 34)
    for o in (3, 5, 7, 9):     # the hand formula it replaced stays within 2.1 %
        assert abs(roofline.recon_flops_hand_formula(o)[0] / roofline.recon_flops(o)[0] - 1.0) < 0.021
    assert roofline.bytes_per_cell_update(1) == 384 and roofline.bytes_per_cell_update(10) == 960
    assert roofline.state_sweep_flops_per_cell(5) == 8 * 295 + 10 + 12 + 8 + 15
    # 2-D: the reconstruction and flux of v, which the reference computes and discards (Dycore.h:565), is not algorithmic work
    assert roofline.state_sweep_flops_per_cell(5, sim2d=True) == 7 * 295 + 10 + 12 + 6 + 12
    # ... and a 2-D march that carries the first tracer counts that tracer's share of the tracer march
    assert roofline.state_sweep_flops_per_cell(5, sim2d=True, fused_tracer=True) == 7 * 295 + 10 + 12 + 6 + 12 + 295 + 4
    assert roofline.state_sweep_bytes_per_cell(last=True, fused_tracer=True) == 8 * 21 + 16
    assert roofline.flops_per_cell_stage(5, 1, 3) == 3 * (9 * 295 + 10 + 12 + 10 + 18) + 5 + 4 + 1 + 30 + 30
    assert roofline.flops_per_cell_stage(5, 1, 2) == 2 * (8 * 295 + 10 + 12 + 8 + 15) + 5 + 4 + 1 + 30 + 30


def test_instrumented_flop_count_reproduces_the_frozen_numbers(tmp_path):
    """Re-count: build the counting program (g++, no GPU) for every order and compare with roofline.RECON_OPS."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import count_flops
    for o in (3, 5, 7, 9):
        c = count_flops.count(o, str(tmp_path))
        assert (c["add"], c["mul"], c["div"]) == roofline.RECON_OPS[o], o
