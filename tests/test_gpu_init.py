"""GPU parity tests of the dycore's idealized initial states (awfl_init_idealized: Dycore::init's thermal / supercell
branches) against the CPU oracle (oracle/init_oracle.c, pinned bit for bit to the compiled
reference) and the reference's golden vectors. Tolerance: relative L2 <= 1e-13 per field (CUDA pow/exp/cos vs glibc)."""
import os

import numpy as np
import pytest
import torch
import yaml

from test_init_oracle import CASES, GOLDEN, build, oracle_init

pytestmark = pytest.mark.gpu
TOL = 1e-13


def make_coupler(case):
    from pam_b200 import PamCoupler
    c = PamCoupler("cuda:0")
    c.set_option("crm_dt", 1.0)
    for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav"):
        c.set_option(k, case["consts"][k])
    c.allocate_coupler_state(case["nz"], case["ny"], case["nx"], case["nens"])
    c.set_grid(case["xlen"], case["ylen"], case["zint"])
    for name, pos, adds in case["tracers"]:
        c.add_tracer(name, name, pos, adds)
    return c


@pytest.mark.parametrize("name", sorted(CASES))
def test_idealized_init_matches_oracle_and_golden(name, tmp_path):
    from pam_b200 import Dycore
    kw = CASES[name]
    case = build(kw)
    c = make_coupler(case)
    # the way the standalone driver asks for it: option "standalone_input_file" -> YAML key initData (Dycore.h:989-1003)
    cfg = tmp_path / "input.yaml"
    cfg.write_text(yaml.safe_dump({"initData": kw["kind"]}))
    c.set_option("standalone_input_file", str(cfg))
    d = Dycore(5)
    d.init(c)
    ref = oracle_init(case, kw["kind"])
    g = np.load(os.path.join(GOLDEN, "init_%s.npz" % name))
    dm = c.get_data_manager_device_readonly()
    un = float(np.linalg.norm(ref["temp"])) * 1e-3
    for k, v in ref.items():
        got = dm.get(k, readonly=True).cpu().numpy()
        floor = un if k in ("uvel", "vvel", "wvel") else 1e-300
        e = float(np.linalg.norm(got - v) / max(np.linalg.norm(v), floor))
        assert e <= TOL, (k, e)
        assert float(np.linalg.norm(got - g[k]) / max(np.linalg.norm(g[k]), floor)) <= TOL, (k, "golden")
    # and the state steps: C1 is "standalone driver, supercell sounding" -- one dycore step from the initial state
    if name == "c1_supercell":
        d.declare_current_profile_as_hydrostatic(c)
        assert d.timeStep(c) >= 1
        assert torch.isfinite(dm.get("temp", readonly=True)).all()


def test_invalid_init_data_is_an_error(tmp_path):
    from pam_b200 import Dycore
    case = build(CASES["thermal_2d"])
    c = make_coupler(case)
    cfg = tmp_path / "input.yaml"
    cfg.write_text(yaml.safe_dump({"initData": "bubble"}))
    c.set_option("standalone_input_file", str(cfg))
    with pytest.raises(RuntimeError):
        Dycore(5).init(c)
