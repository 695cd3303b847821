// The reference's HALOED dynamics arrays at the boundary. Included only by awfl_b200.cu.
//
// Dycore::timeStep keeps its state in two haloed copies -- state (5, nz+2hs, ny+2hs, nx+2hs, nens) and tracers
// (nt, ...), hs = (ord+1)/2 -- and every other public method of the class takes them as arguments
// (dynamics/awfl/Dycore.h:37, :262-267, :608-611, :1096, :1281-1283, :1336-1338). The time step of this library never
// materialises them; these kernels exist so that a caller who holds such arrays -- an MMF driver, a unit
// test written against the reference class -- finds the same methods with the same effects on the same memory:
// conversion to and from the coupler fields, the halo fill with its wall boundary conditions, one semi-discrete
// right-hand side including the in-place divide / multiply-back of the prognostic variables, and the mass diagnostic.
// One thread per interior cell on the 32-bit row map (lane = member: coalesced); init- and test-time work, HBM-bound.
#pragma once

namespace awfl {

struct HaloDims {
  int hs, nyh, nxh;
  long nh;                    // elements of one haloed variable: (nz+2hs)(ny+2hs)(nx+2hs) nens
};
static HaloDims halo_dims(const Geom &g, int hs) {
  HaloDims d;
  d.hs = hs; d.nyh = g.ny + 2 * hs; d.nxh = g.nx + 2 * hs;
  d.nh = (long)(g.nz + 2 * hs) * d.nyh * d.nxh * g.nens;
  return d;
}
// haloed index of array cell (kh, jh, ih) -- halo included -- of member e
__device__ __forceinline__ long hidx(const Geom &g, const HaloDims &d, int kh, int jh, int ih, int e) {
  return (((long)kh * d.nyh + jh) * d.nxh + ih) * g.nens + e;
}

// interior of (state | tracers) -> the library's un-haloed N-field buffer
__global__ void __launch_bounds__(256)
k_h_gather(const Geom g, const RowMap rm, const HaloDims d, const double *__restrict__ st, const double *__restrict__ tr, double *__restrict__ q) {
  long cell; int k, j, i, e;
  if (!row_coords(g, rm, cell, k, j, i, e)) return;
  const long hi = hidx(g, d, d.hs + k, d.hs + j, d.hs + i, e);
  for (int l = 0; l < NS; l++) q[l * g.ncell + cell] = st[l * d.nh + hi];
  for (int t = 0; t < g.nt; t++) q[(NS + t) * g.ncell + cell] = tr[t * d.nh + hi];
}

// Dycore::convert_coupler_to_dynamics (Dycore.h:1336-1388): interior cells only, no tracer clip (that is timeStep's, :130-134)
__global__ void __launch_bounds__(256)
k_h_convert_in(const Geom g, const Phys ph, const FieldPtrs fp, int idWV, const unsigned char *adds_mass, const RowMap rm, const HaloDims d,
               double *st, double *tr) {
  long cell; int k, j, i, e;
  if (!row_coords(g, rm, cell, k, j, i, e)) return;
  const long hi = hidx(g, d, d.hs + k, d.hs + j, d.hs + i, e);
  const double rho_d = fp.f[0][cell], u = fp.f[1][cell], v = fp.f[2][cell], w = fp.f[3][cell], temp = fp.f[4][cell];
  const double rho_v = fp.tr[idWV][cell];
  const double press = rho_d * ph.R_d * temp + rho_v * ph.R_v * temp;
  double rho = rho_d;
  for (int t = 0; t < g.nt; t++) {
    const double x = fp.tr[t][cell];
    if (adds_mass[t]) rho += x;
    tr[t * d.nh + hi] = x;
  }
  const double theta = pow(press / ph.C0, 1.0 / ph.gamma) / rho;
  st[idR * d.nh + hi] = rho;
  st[idU * d.nh + hi] = rho * u;
  st[idV * d.nh + hi] = rho * v;
  st[idW * d.nh + hi] = rho * w;
  st[idT * d.nh + hi] = rho * theta;
}

// Dycore::convert_dynamics_to_coupler (Dycore.h:1281-1331)
__global__ void __launch_bounds__(256)
k_h_convert_out(const Geom g, const Phys ph, const FieldPtrs fp, int idWV, const unsigned char *adds_mass, const RowMap rm, const HaloDims d,
                const double *__restrict__ st, const double *__restrict__ tr) {
  long cell; int k, j, i, e;
  if (!row_coords(g, rm, cell, k, j, i, e)) return;
  const long hi = hidx(g, d, d.hs + k, d.hs + j, d.hs + i, e);
  const double rho = st[idR * d.nh + hi];
  const double u = st[idU * d.nh + hi] / rho, v = st[idV * d.nh + hi] / rho, w = st[idW * d.nh + hi] / rho;
  const double theta = st[idT * d.nh + hi] / rho;
  const double press = ph.C0 * pow(rho * theta, ph.gamma);
  const double rho_v = tr[idWV * d.nh + hi];
  double rho_d = rho;
  for (int t = 0; t < g.nt; t++) {
    const double x = tr[t * d.nh + hi];
    if (adds_mass[t]) rho_d -= x;
    fp.tr[t][cell] = x;
  }
  const double temp = press / (rho_d * ph.R_d + rho_v * ph.R_v);
  fp.f[0][cell] = rho_d; fp.f[1][cell] = u; fp.f[2][cell] = v; fp.f[3][cell] = w; fp.f[4][cell] = temp;
}

// Dycore.h:310-321: pressure of the interior cells and the in-place division by density (a true division: these are the
// values the caller sees afterwards). mode 0: divide + pressure, mode 1: multiply back (Dycore.h:527-532).
__global__ void __launch_bounds__(256)
k_h_divide_multiply(const Geom g, const Phys ph, const RowMap rm, const HaloDims d, double *st, double *tr, double *pr, int mode) {
  long cell; int k, j, i, e;
  if (!row_coords(g, rm, cell, k, j, i, e)) return;
  const long hi = hidx(g, d, d.hs + k, d.hs + j, d.hs + i, e);
  const double rho = st[idR * d.nh + hi];
  if (mode == 0) {
    pr[hi] = ph.C0 * pow(st[idT * d.nh + hi], ph.gamma);
    for (int l = idU; l <= idT; l++) st[l * d.nh + hi] /= rho;
    for (int t = 0; t < g.nt; t++) tr[t * d.nh + hi] /= rho;
  } else {
    for (int l = idU; l <= idT; l++) st[l * d.nh + hi] *= rho;
    for (int t = 0; t < g.nt; t++) tr[t * d.nh + hi] *= rho;
  }
}

// Dycore::halo_exchange, periodic part (Dycore.h:629-657): the x ghosts of the interior rows, then (3-D) the y ghosts of
// the interior columns; corner ghosts are never touched, as in the reference. One thread per (ghost pair, member):
// blockIdx.y = packed variable (state, tracers, pressure), dir 0: (k, j, ii < hs), dir 1: (k, jj < hs, i).
__global__ void __launch_bounds__(256)
k_h_periodic(const Geom g, const HaloDims d, int dir, double *st, double *tr, double *pr) {
  const int v = blockIdx.y;
  double *a = v < NS ? st + v * d.nh : (v < NS + g.nt ? tr + (long)(v - NS) * d.nh : pr);
  const long id = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int n1 = dir == 0 ? g.ny : d.hs, n2 = dir == 0 ? d.hs : g.nx;
  if (id >= (long)g.nz * n1 * n2 * g.nens) return;
  const int e = (int)(id % g.nens);
  long r = id / g.nens;
  const int c2 = (int)(r % n2); r /= n2;
  const int c1 = (int)(r % n1);
  const int k = (int)(r / n1);
  const int hs = d.hs;
  if (dir == 0) {
    const int j = c1, ii = c2;
    a[hidx(g, d, hs + k, hs + j, g.nx + hs + ii, e)] = a[hidx(g, d, hs + k, hs + j, hs + ii, e)];
    a[hidx(g, d, hs + k, hs + j, ii, e)] = a[hidx(g, d, hs + k, hs + j, g.nx + ii, e)];
  } else {
    const int jj = c1, i = c2;
    a[hidx(g, d, hs + k, g.ny + hs + jj, hs + i, e)] = a[hidx(g, d, hs + k, hs + jj, hs + i, e)];
    a[hidx(g, d, hs + k, jj, hs + i, e)] = a[hidx(g, d, hs + k, g.ny + jj, hs + i, e)];
  }
}

// Dycore::halo_exchange, wall part (Dycore.h:662-710) with the intended semantics of SURVEY.md F2: every ghost copy of
// a column is made first (w = 0, the others zero-gradient), then the ghost density by hydrostatic extrapolation at
// constant theta and the ghost pressure from it. One thread per interior column (j, i, member).
__global__ void __launch_bounds__(128)
k_h_wall(const Geom g, const Phys ph, const HaloDims d, const double *__restrict__ dz, double *st, double *tr, double *pr) {
  const long id = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= (long)g.ny * g.nx * g.nens) return;
  const int e = (int)(id % g.nens);
  const int i = (int)((id / g.nens) % g.nx), j = (int)(id / ((long)g.nens * g.nx));
  const int hs = d.hs, nz = g.nz;
  const long sk = (long)d.nyh * d.nxh * g.nens;                 // one haloed level
  const long b0 = hidx(g, d, hs, hs + j, hs + i, e), b1 = hidx(g, d, hs + nz - 1, hs + j, hs + i, e);
  for (int kk = 0; kk < hs; kk++) {
    const long lo = hidx(g, d, kk, hs + j, hs + i, e), hi = hidx(g, d, hs + nz + kk, hs + j, hs + i, e);
    for (int l = 1; l < NS; l++) {
      st[l * d.nh + lo] = l == idW ? 0.0 : st[l * d.nh + b0];
      st[l * d.nh + hi] = l == idW ? 0.0 : st[l * d.nh + b1];
    }
    for (int t = 0; t < g.nt; t++) {
      tr[t * d.nh + lo] = tr[t * d.nh + b0];
      tr[t * d.nh + hi] = tr[t * d.nh + b1];
    }
  }
  for (int side = 0; side < 2; side++) {
    const long b = side == 0 ? b0 : b1;
    const double rho0 = st[idR * d.nh + b], theta0 = st[idT * d.nh + b];
    const double rho0_gm1 = pow(rho0, ph.gamma - 1), theta0_g = pow(theta0, ph.gamma);
    const double dz0 = dz[(long)(side == 0 ? 0 : nz - 1) * g.nens + e];
    for (int kk = 0; kk < hs; kk++) {
      const long c = side == 0 ? b - (long)(kk + 1) * sk : b + (long)(kk + 1) * sk;
      const double inc = ph.grav * (ph.gamma - 1) * dz0 * (kk + 1) / (ph.gamma * ph.C0 * theta0_g);
      const double rg = pow(side == 0 ? rho0_gm1 + inc : rho0_gm1 - inc, 1.0 / (ph.gamma - 1));
      st[idR * d.nh + c] = rg;
      pr[c] = ph.C0 * pow(rg * st[idT * d.nh + c], ph.gamma);
    }
  }
}

}  // namespace awfl
