// Order-templated device code of the AWFL time step: the marching sweep kernels, the hydrostatic interface-pressure
// kernel and the reconstruct test hook. Included only by awfl_ord.cu (one instantiation per translation unit).
#pragma once
#include <math.h>
#include "awfl_types.h"
#include "awfl_weno.cuh"

namespace awfl {





// ---------------------------------------------------------------------------------------------------------------
// Cell loading: conserved -> primitives (Dycore.h:310-321), pressure p = C0 (rho theta)^gamma, and the wall ghost
// cells of Dycore.h:662-710 (intended semantics, SURVEY.md F2: ghost theta = theta of the boundary cell).
// ---------------------------------------------------------------------------------------------------------------
struct Prim {
  double u, v, w, t, m, p;  // velocities, theta, m = rho * normal velocity, pressure
};

template <int DIR>
__device__ __forceinline__ Prim load_prim(const double *__restrict__ q, long idx, long nc, const Phys &ph, double &rho) {
  const double r = __ldg(q + idx);
  const double ru = __ldg(q + nc + idx), rv = __ldg(q + 2 * nc + idx), rw = __ldg(q + 3 * nc + idx);
  const double rt = __ldg(q + 4 * nc + idx);
  Prim c;
  c.p = ph.C0 * pow(rt, ph.gamma);
  c.u = ru / r;
  c.v = rv / r;
  c.w = rw / r;
  c.t = rt / r;
  c.m = r * (DIR == 0 ? c.u : (DIR == 1 ? c.v : c.w));
  rho = r;
  return c;
}

// Hydrostatic, constant-theta extrapolation of density below / above a wall (Dycore.h:683-694, :697-708).
struct WallCtx {
  Prim b;            // boundary cell primitives
  double rho_gm1;    // rho0^(gamma-1)
  double coef;       // grav*(gamma-1)*dz0 / (gamma*C0*theta0^gamma), per unit of (kk+1)
};
__device__ __forceinline__ WallCtx make_wall(const Prim &b, double rho0, double dz0, const Phys &ph) {
  WallCtx w;
  w.b = b;
  w.rho_gm1 = pow(rho0, ph.gamma - 1.0);
  w.coef = ph.grav * (ph.gamma - 1.0) * dz0 / (ph.gamma * ph.C0 * pow(b.t, ph.gamma));
  return w;
}
__device__ __forceinline__ Prim ghost_prim(const WallCtx &w, int dist, double sign, const Phys &ph) {
  const double rg = pow(w.rho_gm1 + sign * (w.coef * dist), 1.0 / (ph.gamma - 1.0));
  Prim g = w.b;
  g.w = 0.0;
  g.m = 0.0;
  g.p = ph.C0 * pow(rg * w.b.t, ph.gamma);
  return g;
}

__device__ __forceinline__ long stride_of(int dir, const Geom &g) {
  return dir == 0 ? (long)g.nens : (dir == 1 ? (long)g.nx * g.nens : (long)g.ny * g.nx * g.nens);
}

// Work-item decode shared by the sweep kernels: item -> (ensemble tile, segment, line); returns false if out of range.
struct Item {
  int e;          // ensemble member of this lane
  long base;      // linear index of cell 0 of the line (this lane)
  int c0, c1;     // cell range [c0,c1) along the sweep direction
  int la, lb;     // the two indices of the line: x: (k,j)  y: (k,i)  z: (j,i)
};
template <int DIR>
__device__ __forceinline__ bool decode_item(long item, const Geom &g, int nseg, Item &it) {
  const int netile = (g.nens + 31) >> 5;
  const int n = DIR == 0 ? g.nx : (DIR == 1 ? g.ny : g.nz);
  const long nlines = DIR == 0 ? (long)g.nz * g.ny : (DIR == 1 ? (long)g.nz * g.nx : (long)g.ny * g.nx);
  if (item >= (long)netile * nseg * nlines) return false;
  const int et = (int)(item % netile);
  long r = item / netile;
  const int seg = (int)(r % nseg);
  const long line = r / nseg;
  it.e = et * 32 + (threadIdx.x & 31);
  if (it.e >= g.nens) return false;
  it.c0 = (int)((long)seg * n / nseg);
  it.c1 = (int)((long)(seg + 1) * n / nseg);
  if (DIR == 0) {        // line = k*ny + j
    it.la = (int)(line / g.ny); it.lb = (int)(line % g.ny);
    it.base = (((long)it.la * g.ny + it.lb) * g.nx) * g.nens + it.e;
  } else if (DIR == 1) { // line = k*nx + i
    it.la = (int)(line / g.nx); it.lb = (int)(line % g.nx);
    it.base = (((long)it.la * g.ny) * g.nx + it.lb) * g.nens + it.e;
  } else {               // line = j*nx + i
    it.la = (int)(line / g.nx); it.lb = (int)(line % g.nx);
    it.base = ((long)it.la * g.nx + it.lb) * g.nens + it.e;
  }
  return true;
}

// ---------------------------------------------------------------------------------------------------------------
// State sweep: Dycore.h:334-519 (flux kernel, one direction) + :553-566 (flux divergence, gravity) and, on the last
// sweep of a stage, the SSPRK3 stage combination Dycore.h:162-165 / :182-187 / :206-211.
//   MODE 0: write (first sweep) or accumulate this direction's flux divergence into tend[5]
//   MODE 1: last sweep: finish the tendency and write qout = a*qbase + b*roundtrip(qin) + c*tend
//   MODE 2: last sweep: finish the tendency and write it to tend_out (tendency-level parity entry point)
// ---------------------------------------------------------------------------------------------------------------


template <int ORD, int DIR, int MODE, bool SHARED>
__global__ void __launch_bounds__(32 * WPB)
k_sweep_state(const Geom g, const Phys ph, const SweepArgs a) {
  constexpr int HS = (ORD + 1) / 2;
  constexpr double cs = 350.0;   // Dycore.h:335
  Item it;
  if (!decode_item<DIR>((long)blockIdx.x * WPB + (threadIdx.x >> 5), g, a.nseg, it)) return;
  const int n = DIR == 0 ? g.nx : (DIR == 1 ? g.ny : g.nz);
  const long sd = stride_of(DIR, g);
  const long nc = g.ncell;
  const double *__restrict__ q = a.qin;

  // wall contexts (z only)
  WallCtx wb, wt;
  if (DIR == 2) {
    if (it.c0 - HS < 0) {
      double r0; Prim b = load_prim<DIR>(q, it.base, nc, ph, r0);
      wb = make_wall(b, r0, __ldg(a.dz + it.e), ph);
    }
    if (it.c1 + HS - 1 >= n) {
      double r0; Prim b = load_prim<DIR>(q, it.base + (long)(n - 1) * sd, nc, ph, r0);
      wt = make_wall(b, r0, __ldg(a.dz + (long)(n - 1) * g.nens + it.e), ph);
    }
  }
  auto get_cell = [&](int c) -> Prim {
    double rr;
    if (DIR == 2) {
      if (c < 0) return ghost_prim(wb, -c, 1.0, ph);
      if (c >= n) return ghost_prim(wt, c - n + 1, -1.0, ph);
      return load_prim<DIR>(q, it.base + (long)c * sd, nc, ph, rr);
    } else {
      int cw = c < 0 ? c + n : (c >= n ? c - n : c);
      return load_prim<DIR>(q, it.base + (long)cw * sd, nc, ph, rr);
    }
  };

  // window of ORD+1 cells: W[s] = cell f-HS+s for the current interface f
  double Wu[ORD + 1], Wv[ORD + 1], Ww[ORD + 1], Wt[ORD + 1], Wm[ORD + 1], Wp[ORD + 1];
#pragma unroll
  for (int s = 0; s < ORD + 1; s++) {
    Prim c = get_cell(it.c0 - HS + s);
    Wu[s] = c.u; Wv[s] = c.v; Ww[s] = c.w; Wt[s] = c.t; Wm[s] = c.m; Wp[s] = c.p;
  }
  const long vst = SHARED ? 1 : g.nens;
  auto vmat = [&](int level) -> VMat<ORD> {
    VMat<ORD> m;
    m.p = a.vtab + (long)level * VMat<ORD>::NEL * vst + (SHARED ? 0 : it.e);
    m.st = vst;
    return m;
  };
  HMat<ORD> hm;

  // acoustic polynomial of cell c0-1 -> left states of interface c0
  double mL, pL;
  {
    double sm[ORD], sp[ORD], E, O;
#pragma unroll
    for (int s = 0; s < ORD; s++) { sm[s] = Wm[s]; sp[s] = Wp[s]; }
    if (DIR == 2) { VMat<ORD> vm = vmat(it.c0); weno_EO<ORD>(sm, vm, E, O); mL = E + O; weno_EO<ORD>(sp, vm, E, O); pL = E + O; }
    else          { weno_EO<ORD>(sm, hm, E, O); mL = E + O; weno_EO<ORD>(sp, hm, E, O); pL = E + O; }
  }
  double Fp[NS];
#pragma unroll
  for (int l = 0; l < NS; l++) Fp[l] = 0.0;

  for (int f = it.c0; f <= it.c1; f++) {
    // acoustic polynomial of cell f -> right states of interface f and left states of interface f+1
    double mR, pR, mLn, pLn;
    {
      double sm[ORD], sp[ORD], E, O;
#pragma unroll
      for (int s = 0; s < ORD; s++) { sm[s] = Wm[s + 1]; sp[s] = Wp[s + 1]; }
      if (DIR == 2) { VMat<ORD> vm = vmat(f + 1); weno_EO<ORD>(sm, vm, E, O); mR = E - O; mLn = E + O; weno_EO<ORD>(sp, vm, E, O); pR = E - O; pLn = E + O; }
      else          { weno_EO<ORD>(sm, hm, E, O); mR = E - O; mLn = E + O; weno_EO<ORD>(sp, hm, E, O); pR = E - O; pLn = E + O; }
    }
    const bool wall = (DIR == 2) && (f == 0 || f == n);   // Dycore.h:477,482,496
    if (wall) { mL = 0.0; mR = 0.0; }
    // characteristic upwinding, Dycore.h:361-365
    const double w1 = 0.5 * (pR - cs * mR);
    const double w2 = 0.5 * (pL + cs * mL);
    const double pp = w1 + w2;
    double rn = (w2 - w1) / cs;
    if (wall) rn = 0.0;
    const int upw = rn > 0.0 ? 0 : 1;   // Dycore.h:368
    double F[NS];
    F[idR] = rn;
    {
      double su[ORD], sv[ORD], sw[ORD], st[ORD], E, O;
#pragma unroll
      for (int s = 0; s < ORD; s++) {
        su[s] = upw ? Wu[s + 1] : Wu[s];
        sv[s] = upw ? Wv[s + 1] : Wv[s];
        sw[s] = upw ? Ww[s + 1] : Ww[s];
        st[s] = upw ? Wt[s + 1] : Wt[s];
      }
      const double sg = upw ? -1.0 : 1.0;
      if (DIR == 2) {
        VMat<ORD> vm = vmat(f + upw);
        weno_EO<ORD>(su, vm, E, O); F[idU] = rn * fma(sg, O, E);
        weno_EO<ORD>(sv, vm, E, O); F[idV] = rn * fma(sg, O, E);
        weno_EO<ORD>(sw, vm, E, O); F[idW] = rn * fma(sg, O, E);
        weno_EO<ORD>(st, vm, E, O); F[idT] = rn * fma(sg, O, E);
      } else {
        weno_EO<ORD>(su, hm, E, O); F[idU] = rn * fma(sg, O, E);
        weno_EO<ORD>(sv, hm, E, O); F[idV] = rn * fma(sg, O, E);
        weno_EO<ORD>(sw, hm, E, O); F[idW] = rn * fma(sg, O, E);
        weno_EO<ORD>(st, hm, E, O); F[idT] = rn * fma(sg, O, E);
      }
    }
    F[DIR == 0 ? idU : (DIR == 1 ? idV : idW)] += pp;

    // interface mass flux for the tracer sweep (x,y: n entries, seam stored at 0; z: n+1 entries)
    if (f < it.c1 || (DIR == 2 && f == n)) a.mflux[it.base + (long)f * sd] = rn;

    if (f > it.c0) {
      const int c = f - 1;
      const long idx = it.base + (long)c * sd;
      const int k = DIR == 2 ? c : it.la;
      const double dd = DIR == 0 ? g.dx : (DIR == 1 ? g.dy : __ldg(a.dz + (long)k * g.nens + it.e));
      double T[NS];
#pragma unroll
      for (int l = 0; l < NS; l++) {
        const double div = (F[l] - Fp[l]) / dd;
        T[l] = a.first ? -div : a.tend[l * nc + idx] - div;
      }
      if (MODE == 0) {
#pragma unroll
        for (int l = 0; l < NS; l++) a.tend[l * nc + idx] = T[l];
      } else {
        const double rho = __ldg(q + idx);
        T[idW] += -__ldg(a.grav_var + (long)k * g.nens + it.e) * rho;   // Dycore.h:560
        if (g.sim2d) T[idV] = 0.0;                                       // Dycore.h:565
        if (MODE == 2) {
#pragma unroll
          for (int l = 0; l < NS; l++) a.qout[l * nc + idx] = T[l];
        } else {
          // round-tripped stage input (x/rho)*rho, Dycore.h:316-319 + :527-530
          double rt[NS];
          rt[idR] = rho; rt[idU] = Wu[HS - 1] * rho; rt[idV] = Wv[HS - 1] * rho; rt[idW] = Ww[HS - 1] * rho; rt[idT] = Wt[HS - 1] * rho;
#pragma unroll
          for (int l = 0; l < NS; l++) {
            double o = fma(a.c, T[l], a.b * rt[l]);
            if (a.a != 0.0) o = fma(a.a, a.qbase[l * nc + idx], a.b * rt[l]) + a.c * T[l];   // plain load: qout may alias qbase
            a.qout[l * nc + idx] = o;
          }
        }
      }
    }
#pragma unroll
    for (int l = 0; l < NS; l++) Fp[l] = F[l];
    mL = mLn; pL = pLn;
    if (f < it.c1) {
#pragma unroll
      for (int s = 0; s < ORD; s++) { Wu[s] = Wu[s + 1]; Wv[s] = Wv[s + 1]; Ww[s] = Ww[s + 1]; Wt[s] = Wt[s + 1]; Wm[s] = Wm[s + 1]; Wp[s] = Wp[s + 1]; }
      Prim c = get_cell(f + HS);
      Wu[ORD] = c.u; Wv[ORD] = c.v; Ww[ORD] = c.w; Wt[ORD] = c.t; Wm[ORD] = c.m; Wp[ORD] = c.p;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Tracer sweep: the tracer part of Dycore.h:381-385 / :434-438 / :513-517. flux = massflux * upwind WENO edge value
// of (tracer/rho). Work item = (tracer, line, segment, ensemble tile).
// ---------------------------------------------------------------------------------------------------------------


template <int ORD, int DIR, bool SHARED>
__global__ void __launch_bounds__(32 * WPB)
k_sweep_tracer(const Geom g, const TracerSweepArgs a) {
  constexpr int HS = (ORD + 1) / 2;
  const long gitem = (long)blockIdx.x * WPB + (threadIdx.x >> 5);
  const int tr = (int)(gitem / a.items_per_tracer);
  if (tr >= g.nt) return;
  Item it;
  if (!decode_item<DIR>(gitem % a.items_per_tracer, g, a.nseg, it)) return;
  const int n = DIR == 0 ? g.nx : (DIR == 1 ? g.ny : g.nz);
  const long sd = stride_of(DIR, g);
  const long nc = g.ncell;
  const double *__restrict__ rho = a.qin;
  const double *__restrict__ tq = a.qin + (long)(NS + tr) * nc;
  double *__restrict__ fl = a.tflux + (long)tr * (DIR == 2 ? nc + nc / g.nz : nc);
  auto get_q = [&](int c) -> double {
    int cw;
    if (DIR == 2) cw = c < 0 ? 0 : (c >= n ? n - 1 : c);       // wall: zero-gradient mixing ratio, Dycore.h:674-677
    else cw = c < 0 ? c + n : (c >= n ? c - n : c);            // periodic, Dycore.h:629-657
    const long idx = it.base + (long)cw * sd;
    return __ldg(tq + idx) / __ldg(rho + idx);                 // Dycore.h:320
  };
  double W[ORD + 1];
#pragma unroll
  for (int s = 0; s < ORD + 1; s++) W[s] = get_q(it.c0 - HS + s);
  const long vst = SHARED ? 1 : g.nens;
  HMat<ORD> hm;
  for (int f = it.c0; f < it.c1; f++) {
    const double rn = __ldg(a.mflux + it.base + (long)f * sd);
    double flux = 0.0;
    if (!(DIR == 2 && f == 0)) {
      const int upw = rn > 0.0 ? 0 : 1;
      double st[ORD], E, O;
#pragma unroll
      for (int s = 0; s < ORD; s++) st[s] = upw ? W[s + 1] : W[s];
      if (DIR == 2) {
        VMat<ORD> vm;
        vm.p = a.vtab + (long)(f + upw) * VMat<ORD>::NEL * vst + (SHARED ? 0 : it.e);
        vm.st = vst;
        weno_EO<ORD>(st, vm, E, O);
      } else {
        weno_EO<ORD>(st, hm, E, O);
      }
      flux = rn * (upw ? E - O : E + O);
    }
    fl[it.base + (long)f * sd] = flux;
    if (f + 1 < it.c1) {
#pragma unroll
      for (int s = 0; s < ORD; s++) W[s] = W[s + 1];
      W[ORD] = get_q(f + HS);
    }
  }
  if (DIR == 2 && it.c1 == n) fl[it.base + (long)n * sd] = 0.0;   // top wall: rw = 0 (Dycore.h:496)
}

// ---------------------------------------------------------------------------------------------------------------
// declare_current_profile_as_hydrostatic (Dycore.h:1392-1504): interface pressure exactly as the solver reconstructs
// it, then the column mean of -(dp)/(rho dz) -- one thread per (k,iens) summing in (j,i) order (deterministic;
// the reference uses atomicAdd, Dycore.h:1487).
// ---------------------------------------------------------------------------------------------------------------


template <int ORD, bool SHARED>
__global__ void __launch_bounds__(128) k_hydro_pint(const Geom g, const Phys ph, const HydroArgs a) {
  constexpr int HS = (ORD + 1) / 2;
  const long col = (long)blockIdx.x * blockDim.x + threadIdx.x;   // (j,i,e)
  const long ncol = (long)g.ny * g.nx * g.nens;
  if (col >= ncol) return;
  const int e = (int)(col % g.nens);
  const long sz = ncol;
  const int n = g.nz;
  auto rho_rt = [&](int k, double &rho, double &rt) {
    if (a.q) { rho = a.q[(long)k * sz + col]; rt = a.q[4 * g.ncell + (long)k * sz + col]; }
    else {
      const long c = (long)k * g.nens + e;
      rho = a.g_rd[c] + a.g_rv[c] + a.g_rc[c] + a.g_ri[c];
      const double p = (a.g_rd[c] * ph.R_d + a.g_rv[c] * ph.R_v) * a.g_t[c];
      rt = pow(p / ph.C0, 1.0 / ph.gamma);
    }
  };
  auto wall_ctx = [&](int k) -> WallCtx {
    double rho, rt; rho_rt(k, rho, rt);
    Prim b; b.u = b.v = b.w = b.m = 0.0; b.t = rt / rho; b.p = 0.0;
    return make_wall(b, rho, a.dz[(long)k * g.nens + e], ph);
  };
  const WallCtx wb = wall_ctx(0), wt = wall_ctx(n - 1);
  auto get_p = [&](int c) -> double {
    if (c < 0) return ghost_prim(wb, -c, 1.0, ph).p;
    if (c >= n) return ghost_prim(wt, c - n + 1, -1.0, ph).p;
    double rho, rt; rho_rt(c, rho, rt);
    return ph.C0 * pow(rt, ph.gamma);
  };
  double W[ORD + 1];
#pragma unroll
  for (int s = 0; s < ORD + 1; s++) W[s] = get_p(-HS + s);
  const long vst = SHARED ? 1 : g.nens;
  double pL;
  {
    double sp[ORD], E, O;
#pragma unroll
    for (int s = 0; s < ORD; s++) sp[s] = W[s];
    VMat<ORD> vm; vm.p = a.vtab + (SHARED ? 0 : e); vm.st = vst;
    weno_EO<ORD>(sp, vm, E, O); pL = E + O;
  }
  for (int f = 0; f <= n; f++) {
    double sp[ORD], E, O;
#pragma unroll
    for (int s = 0; s < ORD; s++) sp[s] = W[s + 1];
    VMat<ORD> vm; vm.p = a.vtab + (long)(f + 1) * VMat<ORD>::NEL * vst + (SHARED ? 0 : e); vm.st = vst;
    weno_EO<ORD>(sp, vm, E, O);
    a.pint[(long)f * sz + col] = 0.5 * (pL + (E - O));
    pL = E + O;
    if (f < n) {
#pragma unroll
      for (int s = 0; s < ORD; s++) W[s] = W[s + 1];
      W[ORD] = get_p(f + HS);
    }
  }
}

// Test hook: Dycore::reconstruct on arbitrary stencils.
template <int ORD, bool SHARED>
__global__ void k_reconstruct(const double *st, int n, int ind, int level, int iens, int nens, const double *vtab, double *out) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n) return;
  double u[ORD], E, O;
#pragma unroll
  for (int s = 0; s < ORD; s++) u[s] = st[(long)q * ORD + s];
  if (level < 0) { HMat<ORD> hm; weno_EO<ORD>(u, hm, E, O); }
  else {
    const long vst = SHARED ? 1 : nens;
    VMat<ORD> vm; vm.p = vtab + (long)level * VMat<ORD>::NEL * vst + (SHARED ? 0 : iens); vm.st = vst;
    weno_EO<ORD>(u, vm, E, O);
  }
  out[q] = ind ? E + O : E - O;
}

}  // namespace awfl
