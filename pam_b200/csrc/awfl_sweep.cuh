// Order-templated device code shared by the AWFL kernels: primitive recovery and wall ghost cells, the work-item decode
// of the marching kernels (awfl_march.cuh), the hydrostatic interface-pressure kernel and the reconstruct test hook.
// Included only by awfl_ord.cu (one instantiation per translation unit).
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

// Work-item decode of the marching kernels. A work item (one warp = one block) is (member tile, line, segment[, tracer]);
// the grid carries it without any runtime division -- blockIdx.x = member tile, blockIdx.y = second line index,
// blockIdx.z = (tracer << (labits+sbits)) | (first line index << sbits) | segment, with 2^sbits segments per line --
// because an integer division by a runtime value goes through the floating-point reciprocal and ptxas then no longer
// treats the march position as warp-uniform (every per-level matrix element becomes a vector load).
struct Item {
  int e;          // ensemble member of this lane
  long base;      // linear index of cell 0 of the line (this lane)
  int c0, c1;     // cell range [c0,c1) along the sweep direction
  int la, lb;     // the two indices of the line: x: (k,j)  y: (k,i)  z: (j,i)
  int tr;         // tracer index (tracer march)
};
// LA_LINES / LB_LINES: lines per work item along the first / second line index (the tracer march takes two adjacent
// lines per warp); it.la / it.lb are then the indices of the item's FIRST line.
template <int DIR, int LA_LINES = 1, int LB_LINES = 1>
__device__ __forceinline__ bool decode_item(const Geom &g, int sbits, int labits, Item &it, int seg_only = -1) {
  const int n = DIR == 0 ? g.nx : (DIR == 1 ? g.ny : g.nz);
  const int la_count = DIR == 2 ? g.ny : g.nz;
  const unsigned z = blockIdx.z;
  const int seg = (int)(z & ((1u << sbits) - 1u));
  if (seg_only >= 0 && seg != seg_only) return false;      // block-uniform: the launch works on one page of levels
  it.la = (int)((z >> sbits) & ((1u << labits) - 1u)) * LA_LINES;
  it.tr = (int)(z >> (sbits + labits));
  it.lb = (int)blockIdx.y * LB_LINES;
  if (it.la >= la_count) return false;
  // lanes past the last member of a ragged slab repeat that member (same values written twice) instead of leaving: with
  // every lane alive the warp stays provably convergent and ptxas keeps loop control and table loads on the uniform path
  it.e = g.e0 + (int)blockIdx.x * 32 + (int)(threadIdx.x & 31);
  if (it.e > g.e0 + g.ne - 1) it.e = g.e0 + g.ne - 1;
  it.c0 = (seg * n) >> sbits;
  it.c1 = ((seg + 1) * n) >> sbits;
  if (DIR == 0) it.base = (((long)it.la * g.ny + it.lb) * g.nx) * g.nens + it.e;          // (k,j)
  else if (DIR == 1) it.base = (((long)it.la * g.ny) * g.nx + it.lb) * g.nens + it.e;     // (k,i)
  else it.base = ((long)it.la * g.nx + it.lb) * g.nens + it.e;                             // (j,i)
  return true;
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
