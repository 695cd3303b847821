"""Synthetic idealized-sounding inputs for the AWFL dycore, shared by tests, bench and the
CPU baselines so that every implementation sees bit-identical FP64 inputs.

Restates, in numpy, the *shape* of the reference's test inputs:
  * vertical grids          utils/generate_vertical_levels.py:102-138  ("equal" and "tanh")
  * supercell base column   standalone/mmf_simplified/supercell_init.h:7-135,
                            pam_core/idealized_profiles.h:55-110
  * column broadcast        pam_core/modules/broadcast_initial_gcm_column.h:32-39
  * temperature noise       pam_core/modules/perturb_temperature.h:44-53 (own counter-based RNG instead of
                            yakl::Random so CPU and GPU runs agree bit-for-bit)
  * P3 + SHOC tracer set    physics/micro/p3/Microphysics.h:119-127, physics/sgs/shoc/SGS.h:103
Host-side only (numpy); nothing here touches the GPU or the oracle.
"""
import numpy as np

P3_CONSTS = dict(R_d=287.042, cp_d=1004.64, R_v=461.505, cp_v=1859.0, p0=1.0e5, grav=9.80616)
DEFAULT_CONSTS = dict(R_d=287.0, cp_d=1003.0, R_v=461.0, cp_v=1859.0, p0=1.0e5, grav=9.81)

# (name, positive, adds_mass) in registration order
TRACERS_NT1 = (("water_vapor", True, True),)
TRACERS_P3_SHOC = (
    ("cloud_water", True, True), ("cloud_water_num", True, False), ("rain", True, True),
    ("rain_num", True, False), ("ice", True, True), ("ice_num", True, False),
    ("ice_rime", True, False), ("ice_rime_vol", True, False), ("water_vapor", True, True),
    ("tke", True, False),
)

# BASELINE.json configs (C1..C5)
CONFIGS = {
    "C1": dict(nz=50, ny=1, nx=64, nens=1, ord=5, tracers=TRACERS_NT1, consts=DEFAULT_CONSTS, grid="equal"),
    "C2": dict(nz=58, ny=1, nx=64, nens=4096, ord=5, tracers=TRACERS_NT1, consts=P3_CONSTS, grid="equal"),
    "C3": dict(nz=58, ny=32, nx=32, nens=1024, ord=5, tracers=TRACERS_NT1, consts=P3_CONSTS, grid="equal"),
    "C4": dict(nz=58, ny=1, nx=128, nens=2048, ord=5, tracers=TRACERS_NT1, consts=P3_CONSTS, grid="equal"),
    "C5": dict(nz=58, ny=1, nx=64, nens=4096, ord=5, tracers=TRACERS_P3_SHOC, consts=P3_CONSTS, grid="equal"),
}


def derived_consts(c):
    """gamma_d, kappa_d, C0 the way Dycore::init derives them (dynamics/awfl/Dycore.h:871-891)."""
    d = dict(c)
    d.setdefault("cv_d", d["cp_d"] - d["R_d"])
    d.setdefault("gamma_d", d["cp_d"] / d["cv_d"])
    d.setdefault("kappa_d", d["R_d"] / d["cp_d"])
    d.setdefault("cv_v", d["R_v"] - d["cp_v"])
    d.setdefault("C0", float(np.power(d["R_d"] * np.power(d["p0"], -d["kappa_d"]), d["gamma_d"])))
    return d


def vertical_grid(nz, ztop=20000.0, kind="equal", inflect=2000.0, steep=8.0, scale=10.0):
    """Interface heights zint[nz+1]. 'equal': k*ztop/nz. 'tanh': dz follows a tanh template, fixed-point
    iterated until dz stops changing (same recipe as utils/generate_vertical_levels.py)."""
    if kind == "equal":
        return np.arange(nz + 1, dtype=np.float64) * (ztop / nz)
    if kind != "tanh":
        raise ValueError(kind)
    infl = inflect / ztop

    def tmpl(z):
        return ((np.tanh((z / ztop - infl) * steep) + 1.0) / 2.0 * (scale - 1.0)) + 1.0

    dz = np.full(nz, ztop / nz)
    for _ in range(200):
        old = dz.copy()
        zmid = np.cumsum(dz) - 0.5 * dz
        t = tmpl(zmid)
        dz = dz[0] * np.concatenate(([1.0], np.cumprod(t[1:] / t[:-1])))
        dz = dz * ztop / dz.sum()
        if np.abs(dz - old).sum() / dz.sum() < 1e-15:
            break
    return np.concatenate(([0.0], np.cumsum(dz)))


_GLL5_PTS = np.array([-0.5, -0.32732683535398857, 0.0, 0.32732683535398857, 0.5])
_GLL5_WTS = np.array([0.05, 0.2722222222222222, 0.35555555555555557, 0.2722222222222222, 0.05])


def _sc_temperature(z, ztop):
    z_trop, T0, Tt = 12000.0, 300.0, 213.0
    lapse = -(Tt - T0) / z_trop
    return np.where(z <= z_trop, T0 - lapse * z, Tt)


def _sc_pressure_dry(z, ztop, R_d, grav):
    z_trop, T0, Tt, p0 = 12000.0, 300.0, 213.0, 1.0e5
    lapse = -(Tt - T0) / z_trop
    T = _sc_temperature(z, ztop)
    p_low = p0 * np.power(T / T0, grav / (R_d * lapse))
    p_trop = p0 * np.power(Tt / T0, grav / (R_d * lapse))
    p_hi = p_trop * np.exp(-grav * (z - z_trop) / (R_d * Tt))
    return np.where(z <= z_trop, p_low, p_hi)


def _sc_qv(z, ztop, R_d, grav):
    T = _sc_temperature(z, ztop)
    pd = _sc_pressure_dry(z, ztop, R_d, grav)
    qvs = 380.0 / pd * np.exp(17.27 * (T - 273.0) / (T - 36.0))
    rh = np.where(z <= 12000.0, 1.0 - 0.75 * np.power(np.clip(z, 0.0, 12000.0) / 12000.0, 1.25), 0.25)   # z can round to -1e-13 at the surface
    return np.minimum(0.014, qvs * rh), T


def supercell_column(zint, R_d, R_v, grav):
    """Cell averages of (rho_d, u, v, w, T, rho_v) for the supercell sounding: hydrostatic integration of
    ln p with 5-point GLL quadrature between GLL points, then GLL cell averages."""
    zint = np.asarray(zint, dtype=np.float64)
    nz = zint.size - 1
    ztop = zint[-1]
    dz = np.diff(zint)
    zm = 0.5 * (zint[:-1] + zint[1:])
    pts, wts = _GLL5_PTS, _GLL5_WTS
    p_gll = np.empty((nz, 5))
    p = 1.0e5
    for k in range(nz):
        p_gll[k, 0] = p
        for kk in range(4):
            b = zm[k] + pts[kk] * dz[k]
            t = zm[k] + pts[kk + 1] * dz[k]
            sub = dz[k] * (pts[kk + 1] - pts[kk])
            zq = 0.5 * (b + t) + sub * pts
            qv, T = _sc_qv(zq, ztop, R_d, grav)
            integ = np.sum(-(1.0 + qv) * grav / (R_d + qv * R_v) / T * wts) * sub
            p = p * np.exp(integ)
            p_gll[k, kk + 1] = p
    zq = zm[:, None] + pts[None, :] * dz[:, None]
    qv, T = _sc_qv(zq, ztop, R_d, grav)
    rho_d = p_gll / (R_d + qv * R_v) / T
    u = np.where(zq < 5000.0, 30.0 * (zq / 5000.0) - 15.0, 15.0)
    col = dict(density_dry=(rho_d * wts).sum(1), uvel=(u * wts).sum(1), vvel=np.zeros(nz), wvel=np.zeros(nz),
               temp=(T * wts).sum(1), water_vapor=(qv * rho_d * wts).sum(1))
    return col


def _mix64(x):
    """splitmix64 finaliser on uint64 arrays (wrap-around arithmetic)."""
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def counter_uniform(seed, iens, k, j, i, stream=0):
    """Uniform [0,1) doubles from a counter-based hash keyed on (seed, stream, iens, k, j, i) (broadcast)."""
    with np.errstate(over="ignore"):
        h = _mix64(np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * np.uint64(stream + 1))
        for c in (iens, k, j, i):
            h = _mix64(h ^ (np.asarray(c).astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)))
    return (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def synthetic_fields(nz, ny, nx, nens, zint, tracers=TRACERS_NT1, consts=P3_CONSTS, seed=12345,
                     ens_offset=0, wind_v=0.0):
    """Coupler fields {name: (nz,ny,nx,nens) float64} for local ensemble members
    [ens_offset, ens_offset+nens). Per-member variation: u += 0.5*(iens mod 17 - 8) m/s and a temperature
    perturbation uniform in +-0.1 K * (1 - k/(nz/4)) below nz/4. For ny>1 a weak v shear is added so the
    y-direction is exercised with both flux signs."""
    zint = np.asarray(zint, dtype=np.float64)
    z1 = zint if zint.ndim == 1 else zint[:, 0]
    col = supercell_column(z1, consts["R_d"], consts["R_v"], consts["grav"])
    shape = (nz, ny, nx, nens)
    K = np.arange(nz).reshape(nz, 1, 1, 1)
    J = np.arange(ny).reshape(1, ny, 1, 1)
    I = np.arange(nx).reshape(1, 1, nx, 1)
    E = (np.arange(nens) + ens_offset).reshape(1, 1, 1, nens)
    f = {}
    for name in ("density_dry", "uvel", "vvel", "wvel", "temp"):
        f[name] = np.broadcast_to(col[name].reshape(nz, 1, 1, 1), shape).copy()
    f["uvel"] += 0.5 * ((E % 17) - 8.0)
    if ny > 1:
        zm = 0.5 * (z1[:-1] + z1[1:]).reshape(nz, 1, 1, 1)
        f["vvel"] += 6.0 * np.sin(2.0 * np.pi * zm / z1[-1]) + 0.25 * ((E % 5) - 2.0) + wind_v
    nlev = max(nz // 4, 1)
    r = counter_uniform(seed, E, K, J, I, stream=0) * 2.0 - 1.0
    scal = np.maximum(0.0, 1.0 - K / float(nlev))
    f["temp"] += 0.1 * r * scal
    # a smooth warm bubble so w, theta and tracers develop gradients in every direction
    xc = (I + 0.5) / nx - 0.5
    yc = (J + 0.5) / ny - 0.5 if ny > 1 else np.zeros((1, 1, 1, 1))
    zc = (K + 0.5) / nz - 0.15
    d2 = (xc / 0.18) ** 2 + (yc / 0.18) ** 2 + (zc / 0.1) ** 2
    f["temp"] += 1.5 * np.exp(-d2) * (1.0 + 0.05 * ((E % 7) - 3.0))
    names = [t[0] for t in tracers]
    for it, name in enumerate(names):
        if name == "water_vapor":
            f[name] = np.broadcast_to(col["water_vapor"].reshape(nz, 1, 1, 1), shape).copy()
            f[name] *= 1.0 + 0.02 * (counter_uniform(seed, E, K, J, I, stream=1) - 0.5)
            continue
        # smooth non-negative blob with exact zeros outside (>= 30 % of cells), different place per tracer
        cx = -0.3 + 0.07 * it
        cz = 0.2 + 0.06 * it
        xr = ((I + 0.5) / nx - 0.5 - cx + 0.5) % 1.0 - 0.5
        yr = ((J + 0.5) / ny - 0.5) if ny > 1 else np.zeros((1, 1, 1, 1))
        zr = (K + 0.5) / nz - cz
        rad = np.sqrt((xr / 0.22) ** 2 + (yr / 0.3) ** 2 + (zr / 0.16) ** 2)
        blob = np.maximum(0.0, np.cos(0.5 * np.pi * np.minimum(rad, 1.0))) ** 2
        blob = np.where(rad < 1.0, blob, 0.0)
        amp = {"cloud_water": 1e-3, "rain": 5e-4, "ice": 3e-4, "ice_rime": 1e-4, "tke": 0.5,
               "cloud_water_num": 1e8, "rain_num": 1e4, "ice_num": 1e5, "ice_rime_vol": 1e-7}.get(name, 1e-3)
        f[name] = amp * blob * (1.0 + 0.1 * ((E % 3) - 1.0)) * np.ones(shape)
    for k_ in f:
        f[k_] = np.ascontiguousarray(f[k_], dtype=np.float64)
    return f


def make_case(nz, ny, nx, nens, tracers=TRACERS_NT1, consts=P3_CONSTS, grid="equal", ztop=20000.0,
              dx=2000.0, seed=12345, ens_offset=0, per_ens_grid=False):
    """Everything needed to set up a coupler: dims, xlen/ylen, zint, tracer table, constants, fields."""
    zint = vertical_grid(nz, ztop, grid)
    if per_ens_grid:
        # member-dependent stretching (API allows per-iens dz, dynamics/awfl/Dycore.h:904-937)
        e = (np.arange(nens) + ens_offset)
        s = 1.0 + 0.02 * ((e % 3) - 1.0)
        zint = zint[:, None] * s[None, :]
    fields = synthetic_fields(nz, ny, nx, nens, zint, tracers, consts, seed, ens_offset)
    return dict(nz=nz, ny=ny, nx=nx, nens=nens, xlen=nx * dx, ylen=max(ny, 1) * dx, zint=zint,
                tracers=tuple(tracers), consts=dict(consts), fields=fields)
