// libawfl_b200.so -- hand-written sm_100a CUDA implementation of the AWFL dycore time step behind the C ABI of
// include/awfl_b200.h. Reference behaviour: E3SM-Project/PAM dynamics/awfl/Dycore.h (cited per kernel below).
//
// Design: the reference materialises two haloed state copies, a pressure array and six flux arrays
// per Runge-Kutta stage and launches ~10 generic one-thread-per-index kernels. Here the dycore state lives UN-haloed
// in three persistent buffers; per stage and per direction ONE marching kernel (awfl_march.cuh) walks each warp along a
// line with the stencil window in a shared-memory ring (32 lanes = 32 consecutive ensemble members, so every global access
// is a coalesced 256-byte line), fusing primitive recovery, pressure, periodic wrap / wall boundary conditions, WENO
// reconstruction, the characteristic Riemann solve, the upwind advective fluxes and the flux divergence; the last
// march of a stage also applies gravity and the SSPRK3 stage combination. Tracers ride on the stored interface
// mass flux in a second, lighter march; FCT positivity + the tracer update are one strip-marching kernel in 2-D, two
// cell-local kernels in 3-D. This file holds those tracer kernels, the conversions, the diagnostics, the idealized
// initial states, the handle and the C ABI; awfl_haloed.cuh the entry points on the reference's haloed arrays.
//
// There is NO CPU fallback in this file: every entry point launches kernels or fails.
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/awfl_b200.h"
#include "awfl_tables.h"
#include "awfl_types.h"
#ifndef AWFL_FUSED_TRACER_Z_MAX_ORD
#define AWFL_FUSED_TRACER_Z_MAX_ORD 7
#endif

#ifndef AWFL_FCT_STRIP
#define AWFL_FCT_STRIP 4
#endif

namespace awfl {

// ---------------------------------------------------------------------------------------------------------------
// FCT positivity (Dycore.h:525-550), cell-local part: the factor by which a cell scales its OUTGOING tracer fluxes.
// start mass = a*T0 + b*Ts (the value the reference stashes in tracers_tend, Dycore.h:158,173,197).
// ---------------------------------------------------------------------------------------------------------------
struct FctArgs {
  const double *q0;       // step-start state (N fields)
  const double *qs;       // stage input state (N fields)
  const double *tfx, *tfy, *tfz;
  const double *fct_start;  // explicit start mass (tendency entry point) or NULL
  double *mult;           // nt fields
  const double *dz, *rdz;
  const unsigned char *positive;
  double sa, sb;          // start = sa*T0 + sb*Ts
  double dt;
  double rdx, rdy;
};

// Work decomposition of the two cell-local tracer kernels: blockIdx.x = (k*ny + j)*rblocks + rb, where the rblocks
// blocks of a row cover its nx*nens (i, member) pairs; blockIdx.y = tracer. Every index is 32-bit and the row part is
// block-uniform (the first version decoded a 64-bit linear cell index with three runtime divisions per thread and was
// issue-bound at a third of the HBM rate: profiles/r01d, 2.7 ms per launch at NT=10).
struct RowMap { int rblocks, rowlen; };   // rowlen = nx * (members of the slab), rblocks = ceil(rowlen / blockDim.x)
static RowMap row_map(const Geom &g, int threads) {
  RowMap m;
  m.rowlen = g.nx * g.ne; m.rblocks = (m.rowlen + threads - 1) / threads;
  return m;
}
static unsigned row_blocks(const Geom &g, const RowMap &m) { return (unsigned)m.rblocks * (unsigned)(g.nz * g.ny); }
__device__ __forceinline__ bool row_coords(const Geom &g, const RowMap m, long &cell, int &k, int &j, int &i, int &e) {
  const unsigned kj = blockIdx.x / (unsigned)m.rblocks, rb = blockIdx.x - kj * (unsigned)m.rblocks;
  const int r = (int)(rb * blockDim.x + threadIdx.x);
  k = (int)(kj / (unsigned)g.ny); j = (int)(kj - (unsigned)k * (unsigned)g.ny);
  i = (int)((unsigned)r / (unsigned)g.ne); e = g.e0 + (r - i * g.ne);
  cell = ((long)kj * g.nx + i) * g.nens + e;
  return r < m.rowlen;
}

__device__ __forceinline__ double rcp_pos_b(double x) {   // same as awfl_weno.cuh rcp_pos
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  const double e = fma(-x, r, 1.0);
  return fma(r, fma(e, e, e), r);
}

// max(x, 0) and min(x, 0) of the FCT mass budget as arithmetic: x + |x| = 2 max(x, 0) and x - |x| = 2 min(x, 0) EXACTLY,
// and the factor 2 is folded into the constant that multiplies the difference anyway (a power of two: the same bits as
// (max - min) * c). fmax / fmin of doubles compile to 7-instruction compare / select / move sequences (their NaN rule);
// in the fused FCT kernel they were 42 % of all non-FP64 instructions (profiles/r02s_fct_line_mix.txt), and that kernel
// is issue-bound (68 % issue-active at 51 % of the HBM rate).
__device__ __forceinline__ double pos2(double x) { return x + fabs(x); }
__device__ __forceinline__ double neg2(double x) { return x - fabs(x); }
// avail / out for out > avail >= 0 (else 1): reciprocal-multiply (MUFU seed + one cubic step, <= 1 ulp, no divergent slow
// path); + 1e-290 keeps the seed in the normal range (a no-op for out > 1e-274; below it avail < out and the product < 1)
__device__ __forceinline__ double fct_multiplier(double avail, double out) {
  const double m = avail * rcp_pos_b(out + 1.0e-290);
  return out > avail ? m : 1.0;
}

__global__ void __launch_bounds__(256) k_fct_mult(const Geom g, const FctArgs a, const RowMap rm) {
  long cell; int k, j, i, e;
  if (!row_coords(g, rm, cell, k, j, i, e)) return;
  const int tr = blockIdx.y;
  const long nc = g.ncell;
  double mult = 1.0;
  if (a.positive[tr]) {
    const long sx = g.nens, sy = (long)g.nx * g.nens, sz = (long)g.ny * g.nx * g.nens;
    const double dzk = __ldg(a.dz + (long)k * g.nens + e), rdzk = __ldg(a.rdz + (long)k * g.nens + e);
    double start;
    if (a.fct_start) start = __ldg(a.fct_start + tr * nc + cell);
    else {
      start = __ldg(a.q0 + (NS + tr) * nc + cell);
      if (a.sb != 0.0) start = a.sa * start + a.sb * __ldg(a.qs + (NS + tr) * nc + cell);
    }
    const double vol = g.dx * g.dy * dzk;
    const double avail = pos2(start) * (0.5 * vol);
    const double *fx = a.tfx + tr * nc, *fz = a.tfz + tr * (nc + sz);
    const double fx0 = __ldg(fx + cell), fx1 = __ldg(fx + (i == g.nx - 1 ? cell - (long)(g.nx - 1) * sx : cell + sx));
    const double ox = (pos2(fx1) - neg2(fx0)) * (0.5 * a.rdx);
    double oy = 0.0;
    if (!g.sim2d) {
      const double *fy = a.tfy + tr * nc;
      const double fy0 = __ldg(fy + cell), fy1 = __ldg(fy + (j == g.ny - 1 ? cell - (long)(g.ny - 1) * sy : cell + sy));
      oy = (pos2(fy1) - neg2(fy0)) * (0.5 * a.rdy);
    }
    const double fz0 = __ldg(fz + cell), fz1 = __ldg(fz + cell + sz);
    const double oz = (pos2(fz1) - neg2(fz0)) * (0.5 * rdzk);
    const double out = (ox + oy + oz) * a.dt * vol;
    mult = fct_multiplier(avail, out);
  }
  a.mult[tr * nc + cell] = mult;
}

// Tracer flux divergence with the FCT-scaled fluxes (Dycore.h:567-583, incl. the periodic-seam min() quirk)
// and the SSPRK3 stage combination + positivity clip (Dycore.h:166-175 / :188-199 / :212-220).
struct TracerUpdArgs {
  const double *q0, *qs;  // step-start and stage-input state (N fields)
  double *qout;           // stage output state (N fields) or tracer tendencies (nt fields) when tend_only
  const double *tfx, *tfy, *tfz, *mult, *rdz;
  const unsigned char *positive;
  double a, b, c;
  double rdx, rdy;
  int tend_only;
};

// scaled flux through an interface: F>0 leaves the low-side cell, F<0 leaves the high-side cell
__device__ __forceinline__ double fct_scaled(double F, double mult_lo, double mult_hi, bool seam) {
  // F == 0 takes either factor (0 * m = 0); periodic seam with F < 0: min() of the two edge copies selects the unscaled one
  // (Dycore.h:574-579). A select and a product, no branch.
  const double m = F > 0.0 ? mult_lo : (seam ? 1.0 : mult_hi);
  return F * m;
}

__global__ void __launch_bounds__(256) k_tracer_update(const Geom g, const TracerUpdArgs a, const RowMap rm) {
  long cell; int k, j, i, e;
  if (!row_coords(g, rm, cell, k, j, i, e)) return;
  const int tr = blockIdx.y;
  const long nc = g.ncell;
  const long sx = g.nens, sy = (long)g.nx * g.nens, sz = (long)g.ny * g.nx * g.nens;
  const double *fx = a.tfx + tr * nc, *fz = a.tfz + tr * (nc + sz), *mu = a.mult + tr * nc;
  const double rdzk = __ldg(a.rdz + (long)k * g.nens + e);
  const double m_c = __ldg(mu + cell);
  // x faces
  const long cxm = i == 0 ? cell + (long)(g.nx - 1) * sx : cell - sx;
  const long cxp = i == g.nx - 1 ? cell - (long)(g.nx - 1) * sx : cell + sx;
  const double f0 = fct_scaled(__ldg(fx + cell), __ldg(mu + cxm), m_c, i == 0);
  const double f1 = fct_scaled(__ldg(fx + cxp), m_c, __ldg(mu + cxp), i == g.nx - 1);
  double tend = -(f1 - f0) * a.rdx;
  // y faces
  if (!g.sim2d) {
    const double *fy = a.tfy + tr * nc;
    const long cym = j == 0 ? cell + (long)(g.ny - 1) * sy : cell - sy;
    const long cyp = j == g.ny - 1 ? cell - (long)(g.ny - 1) * sy : cell + sy;
    const double g0 = fct_scaled(__ldg(fy + cell), __ldg(mu + cym), m_c, j == 0);
    const double g1 = fct_scaled(__ldg(fy + cyp), m_c, __ldg(mu + cyp), j == g.ny - 1);
    tend -= (g1 - g0) * a.rdy;
  }
  // z faces (walls carry zero flux)
  const double h0 = fct_scaled(__ldg(fz + cell), k > 0 ? __ldg(mu + cell - sz) : 1.0, m_c, false);
  const double h1 = fct_scaled(__ldg(fz + cell + sz), m_c, k < g.nz - 1 ? __ldg(mu + cell + sz) : 1.0, false);
  tend -= (h1 - h0) * rdzk;
  if (a.tend_only) { a.qout[tr * nc + cell] = tend; return; }
  const double ts = __ldg(a.qs + (NS + tr) * nc + cell);
  double v = a.b * ts;
  if (a.a != 0.0) v = fma(a.a, a.q0[(NS + tr) * nc + cell], v);   // plain load: qout may alias q0
  v += a.c * tend;
  if (a.positive[tr]) v = 0.5 * pos2(v);
  a.qout[(NS + tr) * nc + cell] = v;
}

// ---------------------------------------------------------------------------------------------------------------
// 2-D fused form of the two kernels above (k_fct_mult + k_tracer_update): one warp = 32 members x a strip of W columns,
// marching up a segment of levels. A level's FCT multipliers are computed for the strip and its two halo columns from
// loads that stay in registers, used at once for the x faces and the lower z face of that level, and carried one level
// for its upper z face -- the multiplier field is never written, the tracer fluxes and the two tracer states are read
// once (plus the halo columns, which the neighbouring strip reads anyway: L2). 5 doubles per (cell, tracer) of HBM
// traffic instead of 12. Same formulas as the two kernels (which remain the 3-D path and the tendency entry point).
// ---------------------------------------------------------------------------------------------------------------
struct FctUpdArgs {
  const double *q0, *qs;   // step-start and stage-input state (N fields)
  double *qout;            // stage output state (N fields)
  const double *tfx, *tfz, *dz, *rdz;
  const unsigned char *positive;
  double a, b, c;          // stage combination: out = a*q0 + b*qs + c*tend; FCT start mass = a*q0 + b*qs
  double dt;               // stage time step of the FCT mass budget
  double rdx;
  int ksegs;               // level segments per column
};

template <int W>
__global__ void __launch_bounds__(32, W <= 4 ? 16 : 10) k_fct_update_2d(const Geom g, const FctUpdArgs a) {
  const int e = g.e0 + (int)blockIdx.x * 32 + (int)threadIdx.x;
  if (e >= g.e0 + g.ne) return;
  const int i0 = (int)blockIdx.y * W;
  const int tr = (int)(blockIdx.z / (unsigned)a.ksegs), seg = (int)(blockIdx.z - (unsigned)tr * (unsigned)a.ksegs);
  const int k0 = (seg * g.nz) / a.ksegs, k1 = ((seg + 1) * g.nz) / a.ksegs;
  const long nc = g.ncell, sz = (long)g.nx * g.nens;
  const double *__restrict__ fx = a.tfx + tr * nc;
  const double *__restrict__ fz = a.tfz + tr * (nc + sz);
  const double *__restrict__ q0 = a.q0 + (NS + tr) * nc;
  const double *__restrict__ qs = a.qs + (NS + tr) * nc;
  double *__restrict__ qo = a.qout + (NS + tr) * nc;
  const bool positive = a.positive[tr] != 0;
  // column offsets of the strip, its halo columns (-1, W) and the interface beyond (W+1), periodic in x
  // (32-bit: nx*nens < 2^31 is checked at create; added to a per-level base pointer, one IMAD.WIDE per load instead of
  // a 64-bit index computation: the kernel spent 30 % of its instructions there)
  unsigned off[W + 3];
#pragma unroll
  for (int c = 0; c < W + 3; c++) {
    const int i = ((i0 + c - 1) % g.nx + g.nx) % g.nx;      // once per thread; a ragged last strip can run past the seam
    off[c] = (unsigned)(i * g.nens + e);
  }
  const int kstart = k0 > 0 ? k0 - 1 : 0;
  double A[W + 2], Mp[W], P[W];       // carried: z fluxes of the interface below the level, multipliers and partial result of the level below
  double crdz_prev = 0.0;
#pragma unroll
  for (int c = 0; c < W + 2; c++) A[c] = __ldg(fz + (long)kstart * sz + off[c]);
  const double *__restrict__ fxk = fx + (long)kstart * sz, *__restrict__ fzk = fz + (long)(kstart + 1) * sz;
  const double *__restrict__ qsk = qs + (long)kstart * sz, *q0k = q0 + (long)kstart * sz;
  double *qok = qo + (long)(kstart - 1) * sz;           // level below the current one
#pragma unroll
  for (int c = 0; c < W; c++) { Mp[c] = 1.0; P[c] = 0.0; }
#pragma unroll 1
  for (int kk = kstart; kk <= k1; kk++, fxk += sz, fzk += sz, qsk += sz, q0k += sz, qok += sz) {
    const bool inside = kk < g.nz;
    double M[W + 2], X[W + 3], B[W + 2], V[W + 2];
    double dzk = 1.0, rdzk = 0.0;
    if (inside) {
      dzk = __ldg(a.dz + (long)kk * g.nens + e); rdzk = __ldg(a.rdz + (long)kk * g.nens + e);
#pragma unroll
      for (int c = 0; c < W + 3; c++) X[c] = __ldg(fxk + off[c]);
#pragma unroll
      for (int c = 0; c < W + 2; c++) B[c] = __ldg(fzk + off[c]);
#pragma unroll
      for (int c = 0; c < W + 2; c++) V[c] = a.b * __ldg(qsk + off[c]);
      if (a.a != 0.0) {
#pragma unroll
        for (int c = 0; c < W + 2; c++) V[c] = fma(a.a, q0k[off[c]], V[c]);   // plain load: qout may alias q0
      }
      const double vol = g.dx * g.dy * dzk, hrdx = 0.5 * a.rdx;
#pragma unroll
      for (int c = 0; c < W + 2; c++) {
        double m = 1.0;
        if (positive) {
          const double avail = pos2(V[c]) * (0.5 * vol);
          const double ox = (pos2(X[c + 1]) - neg2(X[c])) * hrdx;
          const double oz = (pos2(B[c]) - neg2(A[c])) * (0.5 * rdzk);
          const double out = (ox + oz) * a.dt * vol;
          m = fct_multiplier(avail, out);
        }
        M[c] = m;
      }
    } else {
#pragma unroll
      for (int c = 0; c < W + 2; c++) { M[c] = 1.0; B[c] = 0.0; V[c] = 0.0; }
#pragma unroll
      for (int c = 0; c < W + 3; c++) X[c] = 0.0;
    }
    const bool finish = kk - 1 >= k0, start = kk >= k0 && kk < k1;
    const double crdz = a.c * rdzk, crdx = a.c * a.rdx;
#pragma unroll
    for (int c = 0; c < W; c++) {
      const int i = i0 + c;
      if (i >= g.nx) continue;
      // interface kk: upper face of level kk-1, lower face of level kk
      const double hA = fct_scaled(A[c + 1], kk > 0 ? Mp[c] : 1.0, M[c + 1], false);
      if (finish) {
        double v = fma(-crdz_prev, hA, P[c]);
        if (positive) v = 0.5 * pos2(v);
        qok[off[c + 1]] = v;
      }
      if (start) {
        const double f0 = fct_scaled(X[c + 1], M[c], M[c + 1], i == 0);
        const double f1 = fct_scaled(X[c + 2], M[c + 1], M[c + 2], i == g.nx - 1);
        P[c] = fma(crdz, hA, fma(-crdx, f1 - f0, V[c + 1]));
      }
      Mp[c] = M[c + 1];
    }
#pragma unroll
    for (int c = 0; c < W + 2; c++) A[c] = B[c];
    crdz_prev = crdz;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Coupler <-> dycore conversion (Dycore.h:1336-1388, :1281-1331), tracer clip (:130-134) and CFL time step (:65-102)
// ---------------------------------------------------------------------------------------------------------------
struct FieldPtrs {
  double *f[5];
  double *const *tr;   // device array of nt device pointers
};

}  // namespace awfl
#include "awfl_haloed.cuh"
namespace awfl {

__device__ __forceinline__ double atomic_min_pos(double *addr, double v) {
  // positive doubles order like their bit patterns
  unsigned long long old = atomicMin((unsigned long long *)addr, (unsigned long long)__double_as_longlong(v));
  return __longlong_as_double((long long)old);
}

__global__ void __launch_bounds__(256)
k_coupler_to_dycore(const Geom g, const Phys ph, const FieldPtrs fp, int idWV, const unsigned char *adds_mass,
                    const unsigned char *positive, double *q, double cfl, const double *dz, double *dt_min, const RowMap rm) {
  long cell; int k, j, i, e;
  const bool in = row_coords(g, rm, cell, k, j, i, e);
  double dtc = INFINITY;
  if (in) {
    const long nc = g.ncell;
    const double rho_d = fp.f[0][cell], u = fp.f[1][cell], v = fp.f[2][cell], w = fp.f[3][cell], temp = fp.f[4][cell];
    const double rho_v = fp.tr[idWV][cell];
    const double press = rho_d * ph.R_d * temp + rho_v * ph.R_v * temp;
    double rho = rho_d;
    for (int t = 0; t < g.nt; t++) {
      const double x = fp.tr[t][cell];
      if (adds_mass[t]) rho += x;
      q[(NS + t) * nc + cell] = positive[t] ? fmax(0.0, x) : x;
    }
    const double theta = pow(press / ph.C0, 1.0 / ph.gamma) / rho;
    q[cell] = rho;
    q[nc + cell] = rho * u;
    q[2 * nc + cell] = rho * v;
    q[3 * nc + cell] = rho * w;
    q[4 * nc + cell] = rho * theta;
    if (dt_min) {
      const double rho2 = rho_d + rho_v;
      const double p = (rho_d * ph.R_d + rho_v * ph.R_v) * temp;
      const double c = sqrt(ph.gamma * p / rho2);
      const double dtx = cfl * g.dx / (fabs(u) + c), dty = cfl * g.dy / (fabs(v) + c);
      const double dtz = cfl * dz[(long)k * g.nens + e] / (fabs(w) + c);
      dtc = fmin(fmin(dtx, dty), dtz);
    }
  }
  if (dt_min) {
    for (int o = 16; o > 0; o >>= 1) dtc = fmin(dtc, __shfl_xor_sync(0xffffffffu, dtc, o));
    __shared__ double sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = dtc;
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w2 = 1; w2 < (int)(blockDim.x >> 5); w2++) dtc = fmin(dtc, sm[w2]);
      if (dtc < INFINITY) atomic_min_pos(dt_min, dtc);
    }
  }
}

__global__ void __launch_bounds__(256)
k_cfl_only(const Geom g, const Phys ph, const FieldPtrs fp, int idWV, double cfl, const double *dz, double *dt_min, const RowMap rm) {
  long cell; int k, j, i, e;
  const bool in = row_coords(g, rm, cell, k, j, i, e);
  double dtc = INFINITY;
  if (in) {
    const double rho_d = fp.f[0][cell], u = fp.f[1][cell], v = fp.f[2][cell], w = fp.f[3][cell], temp = fp.f[4][cell];
    const double rho_v = fp.tr[idWV][cell];
    const double rho = rho_d + rho_v;
    const double p = (rho_d * ph.R_d + rho_v * ph.R_v) * temp;
    const double c = sqrt(ph.gamma * p / rho);
    const double dtx = cfl * g.dx / (fabs(u) + c), dty = cfl * g.dy / (fabs(v) + c);
    const double dtz = cfl * dz[(long)k * g.nens + e] / (fabs(w) + c);
    dtc = fmin(fmin(dtx, dty), dtz);
  }
  for (int o = 16; o > 0; o >>= 1) dtc = fmin(dtc, __shfl_xor_sync(0xffffffffu, dtc, o));
  __shared__ double sm[8];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = dtc;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w2 = 1; w2 < (int)(blockDim.x >> 5); w2++) dtc = fmin(dtc, sm[w2]);
    if (dtc < INFINITY) atomic_min_pos(dt_min, dtc);
  }
}

__global__ void __launch_bounds__(256)
k_dycore_to_coupler(const Geom g, const Phys ph, const FieldPtrs fp, int idWV, const unsigned char *adds_mass, const double *q,
                    const RowMap rm) {
  long cell; int k, j, i, e;
  if (!row_coords(g, rm, cell, k, j, i, e)) return;
  const long nc = g.ncell;
  const double rho = q[cell];
  const double u = q[nc + cell] / rho, v = q[2 * nc + cell] / rho, w = q[3 * nc + cell] / rho, theta = q[4 * nc + cell] / rho;
  const double press = ph.C0 * pow(rho * theta, ph.gamma);
  const double rho_v = q[(NS + idWV) * nc + cell];
  double rho_d = rho;
  for (int t = 0; t < g.nt; t++) {
    const double x = q[(NS + t) * nc + cell];
    if (adds_mass[t]) rho_d -= x;
    fp.tr[t][cell] = x;
  }
  const double temp = press / (rho_d * ph.R_d + rho_v * ph.R_v);
  fp.f[0][cell] = rho_d; fp.f[1][cell] = u; fp.f[2][cell] = v; fp.f[3][cell] = w; fp.f[4][cell] = temp;
}

__global__ void __launch_bounds__(128) k_hydro_mean(const Geom g, const HydroArgs a) {
  const long id = (long)blockIdx.x * blockDim.x + threadIdx.x;   // (k,e)
  if (id >= (long)g.nz * g.nens) return;
  const int e = (int)(id % g.nens), k = (int)(id / g.nens);
  const long sz = (long)g.ny * g.nx * g.nens;
  const double r_nx_ny = 1.0 / (g.nx * g.ny);
  const double dzk = a.dz[id];
  double acc = 0.0;
  for (int ji = 0; ji < g.ny * g.nx; ji++) {
    const long col = (long)ji * g.nens + e;
    double rho;
    if (a.q) rho = a.q[(long)k * sz + col];
    else { const long c = id; rho = a.g_rd[c] + a.g_rv[c] + a.g_rc[c] + a.g_ri[c]; }
    acc += -(a.pint[(long)(k + 1) * sz + col] - a.pint[(long)k * sz + col]) / (rho * dzk) * r_nx_ny;
  }
  a.grav_var[id] = acc;
}

// Dycore::compute_mass (Dycore.h:37-60): one thread per (variable, iens), cells summed in (k,j,i) order.
__global__ void __launch_bounds__(128) k_mass(const Geom g, const double *q, const double *dz, double *out) {
  const long id = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= (long)(g.nt + 2) * g.nens) return;
  const int e = (int)(id % g.nens), v = (int)(id / g.nens);
  const int field = v < g.nt ? NS + v : (v == g.nt ? idR : idT);
  const double *src = q + (long)field * g.ncell;
  double s = 0.0;
  for (int k = 0; k < g.nz; k++) {
    const double dzk = dz[(long)k * g.nens + e];
    for (int ji = 0; ji < g.ny * g.nx; ji++) s += src[((long)k * g.ny * g.nx + ji) * g.nens + e] * dzk;
  }
  out[id] = s / ((double)g.nz * g.ny * g.nx);
}

// FP64 FMA-pipe peak: 8 independent dependent-chains per thread.
__global__ void __launch_bounds__(256) k_fp64_peak(int iters, double seed, double *sink) {
  double x0 = seed + threadIdx.x * 1e-9, x1 = x0 + 1e-3, x2 = x0 + 2e-3, x3 = x0 + 3e-3, x4 = x0 + 4e-3, x5 = x0 + 5e-3,
         x6 = x0 + 6e-3, x7 = x0 + 7e-3;
  const double a = 0.9999999, b = 1e-7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
      x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
      x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
  }
  const double s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 12345.6789) sink[0] = s;
}

// ---------------------------------------------------------------------------------------------------------------
// Idealized initial states of Dycore::init (Dycore.h:985-1091): supercell sounding (init_supercell, :1096-1276, point
// functions :774-830) and rising thermal (:1017-1086, point functions :742-769), each followed by
// convert_dynamics_to_coupler (:1281-1331). Init-time kernels: a column kernel (hydrostatic integration up the column:
// one thread per member resp. per (level, member)) and a cell kernel (9x9x9 Gauss-Legendre-Lobatto cell averages, in
// the reference's summation order). Every member integrates its OWN column (the reference's cell-average kernel
// overwrites all members from each thread, Dycore.h:1228-1231 -- the same thing whenever members share a grid).
// ---------------------------------------------------------------------------------------------------------------
__constant__ double c_gll9_pts[9] = AWFL_GLL9_PTS_INIT;
__constant__ double c_gll9_wts[9] = AWFL_GLL9_WTS_INIT;

struct InitArgs {
  const double *zint;            // (nz+1,nens)
  double *pg, *vapg;             // (nz,9,nens): pressure / vapour density at the GLL points of every cell
  double *hy_dens, *hy_pres;     // (nz,nens)
  double xlen, ylen, cp_d, p0;
};

__device__ __forceinline__ double sc_temperature(double z, double z_top) {
  const double z_0 = 0, z_trop = 12000, T_0 = 300, T_trop = 213, T_top = 213;
  if (z <= z_trop) { const double lapse = -(T_trop - T_0) / (z_trop - z_0); return T_0 - lapse * (z - z_0); }
  const double lapse = -(T_top - T_trop) / (z_top - z_trop);
  return T_trop - lapse * (z - z_trop);
}
__device__ __forceinline__ double sc_pressure_dry(double z, double z_top, double R_d, double grav) {
  const double z_0 = 0, z_trop = 12000, T_0 = 300, T_trop = 213, p_0 = 100000;
  const double lapse = -(T_trop - T_0) / (z_trop - z_0);
  if (z <= z_trop) return p_0 * pow(sc_temperature(z, z_top) / T_0, grav / (R_d * lapse));
  const double p_trop = p_0 * pow(T_trop / T_0, grav / (R_d * lapse));
  return p_trop * exp(-grav * (z - z_trop) / (R_d * T_trop));      // T_top == T_trop: the isothermal branch (:805-807)
}
__device__ __forceinline__ double sc_qv(double z, double z_top, double R_d, double grav, double &temp) {
  temp = sc_temperature(z, z_top);
  const double qvs = 380 / sc_pressure_dry(z, z_top, R_d, grav) * exp(17.27 * (temp - 273) / (temp - 36));
  double relhum = z <= 12000.0 ? 1.0 - 0.75 * pow(z / 12000.0, 1.25) : 0.25;
  if (relhum * qvs > 0.014) relhum = 0.014 / qvs;
  return fmin(0.014, qvs * relhum);       // like std::min(0.014, x): a NaN x (height rounded below zero) gives 0.014
}

__global__ void __launch_bounds__(64) k_init_supercell_column(const Geom g, const Phys ph, const InitArgs a) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= g.nens) return;
  const double z_top = a.zint[(long)g.nz * g.nens + e];
  double p = 100000.0;
  for (int k = 0; k < g.nz; k++) {
    const double z0 = a.zint[(long)k * g.nens + e], z1 = a.zint[(long)(k + 1) * g.nens + e];
    const double dz = z1 - z0, zmid = 0.5 * (z0 + z1);
    a.pg[((long)k * 9) * g.nens + e] = p;
    for (int kk = 0; kk < 8; kk++) {
      const double b = zmid + c_gll9_pts[kk] * dz, t = zmid + c_gll9_pts[kk + 1] * dz, m = 0.5 * (b + t);
      const double sub = dz * (c_gll9_pts[kk + 1] - c_gll9_pts[kk]);
      double tot = 0;
      for (int q = 0; q < 9; q++) {
        double temp;
        const double qv = sc_qv(m + sub * c_gll9_pts[q], z_top, ph.R_d, ph.grav, temp);
        tot += (-(1 + qv) * ph.grav / (ph.R_d + qv * ph.R_v) / temp) * c_gll9_wts[q];
      }
      tot *= sub;
      p = p * exp(tot);
      a.pg[((long)k * 9 + kk + 1) * g.nens + e] = p;
    }
    double press_tot = 0, dens_tot = 0;
    for (int kk = 0; kk < 9; kk++) {
      double temp;
      const double qv = sc_qv(zmid + c_gll9_pts[kk] * dz, z_top, ph.R_d, ph.grav, temp);
      const double press = a.pg[((long)k * 9 + kk) * g.nens + e];
      const double dens_dry = press / (ph.R_d + qv * ph.R_v) / temp, dens_vap = qv * dens_dry;
      a.vapg[((long)k * 9 + kk) * g.nens + e] = dens_vap;
      press_tot += press * c_gll9_wts[kk];
      dens_tot += (dens_dry + dens_vap) * c_gll9_wts[kk];
    }
    a.hy_dens[(long)k * g.nens + e] = dens_tot;
    a.hy_pres[(long)k * g.nens + e] = press_tot;
  }
}

__global__ void __launch_bounds__(128) k_init_thermal_column(const Geom g, const Phys ph, const InitArgs a) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
  if (e >= g.nens) return;
  const double z0 = a.zint[(long)k * g.nens + e], z1 = a.zint[(long)(k + 1) * g.nens + e];
  const double dz = z1 - z0, zmid = 0.5 * (z0 + z1);
  double hd = 0, hp = 0;
  for (int kk = 0; kk < 9; kk++) {
    const double z = zmid + c_gll9_pts[kk] * dz;
    const double exner = 1.0 - ph.grav * z / (a.cp_d * 300.0);               // hydro_const_theta, Dycore.h:742-752
    const double p = a.p0 * pow(exner, a.cp_d / ph.R_d);
    const double rt = pow(p / ph.C0, 1.0 / ph.gamma);
    const double hr = rt / 300.0;
    hd += hr * c_gll9_wts[kk];
    hp += ph.C0 * pow(hr * 300.0, ph.gamma) * c_gll9_wts[kk];
  }
  a.hy_dens[(long)k * g.nens + e] = hd;
  a.hy_pres[(long)k * g.nens + e] = hp;
}

// spec 0: thermal, 1: supercell. One thread per cell; writes the coupler fields.
__global__ void __launch_bounds__(128)
k_init_cells(const Geom g, const Phys ph, const InitArgs a, const FieldPtrs fp, int idWV, const unsigned char *adds_mass, int spec,
             const RowMap rm, double *dyn_st, double *dyn_tr, const HaloDims hdim) {
  long cell; int k, j, i, e;
  if (!row_coords(g, rm, cell, k, j, i, e)) return;
  const double z0 = a.zint[(long)k * g.nens + e], z1 = a.zint[(long)(k + 1) * g.nens + e];
  const double dz = z1 - z0, zmid = 0.5 * (z0 + z1);
  const double hd = a.hy_dens[(long)k * g.nens + e], hp = a.hy_pres[(long)k * g.nens + e];
  double rho, ru = 0, rv = 0, rw = 0, rt, wv = 0;
  if (spec == 1) {
    rho = hd;
    rt = pow(hp / ph.C0, 1.0 / ph.gamma);
    for (int kk = 0; kk < 9; kk++) {
      const double zloc = zmid + c_gll9_pts[kk] * dz;
      const double uvel = zloc < 5000.0 ? 30.0 * (zloc / 5000.0) - 15.0 : 30.0 - 15.0;
      const double dens_vap = a.vapg[((long)k * 9 + kk) * g.nens + e];
      for (int jj = 0; jj < 9; jj++)
        for (int ii = 0; ii < 9; ii++) {
          const double factor = c_gll9_wts[ii] * c_gll9_wts[jj] * c_gll9_wts[kk];
          ru += rho * uvel * factor;
          wv += dens_vap * factor;
        }
    }
  } else {
    const double dxx = g.dx, dyy = g.dy;
    const double ht = pow(hp / ph.C0, 1.0 / ph.gamma) / hd;
    rho = 0; rt = 0;
    for (int kk = 0; kk < 9; kk++)
      for (int jj = 0; jj < 9; jj++)
        for (int ii = 0; ii < 9; ii++) {
          const double x = (i + 0.5) * dxx + c_gll9_pts[ii] * dxx;
          double y = (j + 0.5) * dyy + c_gll9_pts[jj] * dyy;
          if (g.sim2d) y = a.ylen / 2;
          const double z = zmid + c_gll9_pts[kk] * dz;
          // sample_ellipse_cosine(2, x,y,z, xlen/2, ylen/2, 2000, 2000,2000,2000), Dycore.h:757-769
          const double xr = (x - a.xlen / 2) / 2000., yr = (y - a.ylen / 2) / 2000., zr = (z - 2000.) / 2000.;
          const double dist = sqrt(xr * xr + yr * yr + zr * zr) * M_PI / 2.;
          double bump = 0.;
          if (dist <= M_PI / 2.) { const double cd = cos(dist); bump = 2.0 * (cd * cd); }
          const double theta = ht + bump;
          const double wt = c_gll9_wts[ii] * c_gll9_wts[jj] * c_gll9_wts[kk];
          rho += hd * wt;
          rt += hd * theta * wt;
        }
  }
  if (dyn_st) {
    // Dycore::init_supercell (Dycore.h:1096-1276) proper: the haloed dynamics arrays, interior cells
    const long hi = hidx(g, hdim, hdim.hs + k, hdim.hs + j, hdim.hs + i, e);
    dyn_st[idR * hdim.nh + hi] = rho; dyn_st[idU * hdim.nh + hi] = ru; dyn_st[idV * hdim.nh + hi] = rv;
    dyn_st[idW * hdim.nh + hi] = rw; dyn_st[idT * hdim.nh + hi] = rt;
    for (int t = 0; t < g.nt; t++) dyn_tr[t * hdim.nh + hi] = t == idWV ? wv : 0.0;
    return;
  }
  // convert_dynamics_to_coupler, Dycore.h:1313-1330
  const double theta = rt / rho;
  const double press = ph.C0 * pow(rho * theta, ph.gamma);
  double rho_d = rho;
  for (int t = 0; t < g.nt; t++) {
    const double x = t == idWV ? wv : 0.0;
    if (adds_mass[t]) rho_d -= x;
    fp.tr[t][cell] = x;
  }
  fp.f[0][cell] = rho_d; fp.f[1][cell] = ru / rho; fp.f[2][cell] = rv / rho; fp.f[3][cell] = rw / rho;
  fp.f[4][cell] = press / (rho_d * ph.R_d + wv * ph.R_v);
}

}  // namespace awfl

// =================================================================================================================
// Host side: handle, workspace, launch sequencing, C ABI
// =================================================================================================================
using namespace awfl;

struct awfl_handle {
  awfl_config cfg;
  Geom g;
  Phys ph;
  int ord, hs;
  double cv_d, kappa;
  bool shared_grid;
  int seg_len;
  cudaStream_t stream;
  bool own_stream;
  std::string err;
  long launches;
  // device memory
  double *Q[3];          // dycore state buffers, (5+nt) fields each
  double *tend;          // 5 fields
  double *mflux[3];      // x,y (ncell), z (ncell + one level)
  double *tflux[3];      // nt fields each
  double *mult;          // nt fields
  double *dz, *rdz, *grav_var, *vtab, *vs2c, *vwrl, *pint;
  bool use_cvtab;               // shared vertical grid whose table fits the constant bank
  // ... or does not fit, but the levels ONE SEGMENT of the vertical marches needs do: the z marches then run segment by
  // segment, each launch preceded by the upload of its page (pinned host copy -> constant bank, on the launching stream)
  bool use_cvpage;
  bool cvpage_ok;               // false while the host-buffer entry point runs slabs on several streams (they would overwrite each other's pages)
  int cvpage_sbits;             // 2^sbits segments = pages per line
  double *vtab_pin;             // pinned host copy of the whole kernel table
  int *vlvl_pin;                // pinned: per page, level -> offset in the page
  double *dt_min_d, *dt_min_h;   // device scalar + pinned host scalar
  double *mass_d;
  unsigned char *positive_d, *adds_mass_d;
  std::vector<unsigned char> positive_h, adds_mass_h;
  double **trptr_d;      // device array of tracer field pointers
  double **trptr_out_d;  // the same for the result staging of the host-buffer entry point
  double *stage_out;     // result staging of the host-buffer entry point (the staged inputs stay intact beside it)
  // optional per-kernel-class CUDA-event timing (awfl_profile_begin/end)
  bool prof;
  std::vector<cudaEvent_t> ev_pool;
  std::vector<int> ev_class;     // class of event pair i (events 2i, 2i+1)
  size_t ev_used;
  double *stage_fields;  // staging buffer for the host-buffer entry point (5+nt fields)
  int host_last_ncycles;      // sub-cycle count derived by the previous awfl_time_step_host (0 = none yet): the prediction
  long host_mispredictions;   // calls whose prediction was wrong and were redone un-overlapped
  bool tracer_pairs;     // tracer march: two tracers per warp on one line instead of two lines of one tracer (AWFL_TRACER_PAIRS)
  int fused_tracer;      // 2-D: the first tracer is reconstructed by the state march of x (bit 0) / z (bit 1) instead of the tracer march (AWFL_FUSED_TRACER)
  bool general_2d;       // 2-D on the general state marches (AWFL_GENERAL_2D=1; see SweepArgs::general_2d)
  bool fused_fct;        // 2-D: fused FCT + tracer update kernel (AWFL_FUSED_FCT=0 selects the two-kernel path)
  int host_slabs;        // member slabs of the host-buffer entry point (copy/compute overlap); 1 = no pipelining, 0 = auto
  cudaStream_t s_in, s_out;              // copy streams of the host-buffer entry point
  cudaStream_t s_cmp[4];                 // its compute streams: consecutive slabs alternate, so the tail of one slab's
                                         // kernels (a partial wave of one-warp blocks) overlaps the head of the next slab's
  int host_cstreams;                     // 1 = all slabs on the handle's stream, 2..4 = round robin over s_cmp
  std::vector<cudaEvent_t> slab_ev;      // 2 events per slab: inputs on the device / results ready on the device
  double *tmp_state;     // scratch for awfl_compute_tendencies
  double *pcache;        // pressure of the current stage input (SweepArgs::pcache)
  double *halo_pressure; // haloed pressure scratch of awfl_compute_tendencies_haloed (allocated on first use)
};

// Process-wide state is limited to these two objects. The message of a failed awfl_create is per host thread (one
// coupler per host thread in the reference, pam_core/pam_interface/pam_interface.h:16-32). The constant-bank copy of the
// vertical matrices (c_vtab, one per device and reconstruction order) is CLAIMED at awfl_create and kept until
// awfl_destroy: handles of one (device, order) whose tables are bit-identical share it (reference count); a handle that
// finds it taken by a different grid reads its own global-memory table instead (MatSel VT 3 / 2). Nothing is uploaded
// or synchronised inside a time step, so handles on different host threads or devices never race on it.
static thread_local std::string g_create_err;
struct CvtabSlot { int refs = 0; bool exclusive = false; std::vector<double> table; std::vector<int> levels; };
static std::mutex g_cvtab_mu;
static std::map<std::pair<int, int>, CvtabSlot> g_cvtab;   // (device, ord) -> owner table

// Every entry point runs on the handle's device whatever the caller's current device is, and leaves the caller's
// current device as it found it.
struct DeviceGuard {
  int prev = -1, want;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int dev) : want(dev) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != want) err = cudaSetDevice(want);
  }
  ~DeviceGuard() { if (prev >= 0 && prev != want) cudaSetDevice(prev); }
};
#define ENTER(h)                                                                                        \
  DeviceGuard dev_guard_((h)->cfg.device);                                                              \
  if (dev_guard_.err != cudaSuccess) {                                                                  \
    (h)->err = std::string("cudaSetDevice: ") + cudaGetErrorString(dev_guard_.err);                     \
    return 1;                                                                                           \
  }

#define CK(call)                                                                                        \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) {                                                                            \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                      \
      return 1;                                                                                         \
    }                                                                                                   \
  } while (0)

static int fail(awfl_handle *h, const std::string &m) { h->err = m; return 1; }

// kernel classes for awfl_profile_end (keep in sync with include/awfl_b200.h)
enum { KC_IN = 0, KC_OUT, KC_STATE_X, KC_STATE_Y, KC_STATE_Z, KC_TRACER_X, KC_TRACER_Y, KC_TRACER_Z, KC_FCT, KC_TRACER_UPD, KC_OTHER, KC_COUNT };

struct ProfScope {
  awfl_handle *h;
  bool on;
  cudaStream_t st;
  ProfScope(awfl_handle *h_, int cls, cudaStream_t st_ = nullptr);
  ~ProfScope();
};

// ---- host-side vertical matrices: Dycore.h:897-939 + TransformMatrices_variable.h:10-69 + YAKL matinv_ge ----
static void matinv_ge(int n, const double *a, double *inv) {
  std::vector<double> sc(a, a + n * n);
  for (int r = 0; r < n; r++) for (int c = 0; c < n; c++) inv[r * n + c] = r == c ? 1.0 : 0.0;
  for (int d = 0; d < n; d++) {
    const double f = 1.0 / sc[d * n + d];
    for (int c = d; c < n; c++) sc[d * n + c] *= f;
    for (int c = 0; c < n; c++) inv[d * n + c] *= f;
    for (int r = d + 1; r < n; r++) {
      const double gq = sc[r * n + d];
      for (int c = d; c < n; c++) sc[r * n + c] -= gq * sc[d * n + c];
      for (int c = 0; c < n; c++) inv[r * n + c] -= gq * inv[d * n + c];
    }
  }
  for (int d = n - 1; d >= 1; d--)
    for (int r = 0; r < d; r++) {
      const double gq = sc[r * n + d];
      for (int c = r + 1; c < n; c++) sc[r * n + c] -= gq * sc[d * n + c];
      for (int c = 0; c < n; c++) inv[r * n + c] -= gq * inv[d * n + c];
    }
}
static void sten_to_coefs_var(int n, const double *locs, double *out) {
  std::vector<double> pw(locs, locs + n + 1), c2s(n * n);
  for (int j = 0; j < n; j++) c2s[j] = 1.0;
  for (int i = 1; i < n; i++) {
    for (int j = 0; j < n + 1; j++) pw[j] *= locs[j];
    for (int j = 0; j < n; j++) c2s[i * n + j] = 1. / (i + 1.) * (pw[j] - pw[j + 1]) / (locs[j] - locs[j + 1]);
  }
  matinv_ge(n, c2s.data(), out);
}
static void ideal_weights(int ord, double *idl) {
  const int m = (ord + 1) / 2;
  const double *raw = ord == 3 ? AWFL_IDL_RAW_3 : ord == 5 ? AWFL_IDL_RAW_5 : ord == 7 ? AWFL_IDL_RAW_7 : AWFL_IDL_RAW_9;
  double sum = 0.0;
  for (int i = 0; i < m + 1; i++) sum += raw[i];
  for (int i = 0; i < m + 1; i++) idl[i] = raw[i] / (sum + 1.0e-20);
}

// Coefficient-space maps of the "row" form (tools/gen_tables.py tv_rows): glo (hs x hs), ghi (ord x ord), row-major.
static void tv_row_maps(int ord, std::vector<double> &glo, std::vector<double> &ghi) {
  static const double g3l[2][2] = AWFL_GLO_3_INIT, g3h[3][3] = AWFL_GHI_3_INIT;
  static const double g5l[3][3] = AWFL_GLO_5_INIT, g5h[5][5] = AWFL_GHI_5_INIT;
  static const double g7l[4][4] = AWFL_GLO_7_INIT, g7h[7][7] = AWFL_GHI_7_INIT;
  static const double g9l[5][5] = AWFL_GLO_9_INIT, g9h[9][9] = AWFL_GHI_9_INIT;
  const int hs = (ord + 1) / 2;
  const double *l = ord == 3 ? &g3l[0][0] : ord == 5 ? &g5l[0][0] : ord == 7 ? &g7l[0][0] : &g9l[0][0];
  const double *hh = ord == 3 ? &g3h[0][0] : ord == 5 ? &g5h[0][0] : ord == 7 ? &g7h[0][0] : &g9h[0][0];
  glo.assign(l, l + hs * hs);
  ghi.assign(hh, hh + ord * ord);
}

// Builds vs2c (nz+2,ord,ord,nens), vwrl (nz+2,hs,hs,hs,nens) exactly as the reference does, and the kernel table
// vtab: per level [rlo(hs^3) | rhi(ord^2)] in row form, element-major, ensemble fastest (or no ensemble axis if shared).
static void build_vertical(int ord, int nz, int nens, const double *dz, bool shared, std::vector<double> &vs2c,
                           std::vector<double> &vwrl, std::vector<double> &vtab) {
  const int hs = (ord + 1) / 2, nel = hs * hs * hs + ord * ord;
  vs2c.assign((size_t)(nz + 2) * ord * ord * nens, 0.0);
  vwrl.assign((size_t)(nz + 2) * hs * hs * hs * nens, 0.0);
  vtab.assign((size_t)(nz + 2) * nel * (shared ? 1 : nens), 0.0);
  std::vector<double> idl(hs + 1);
  ideal_weights(ord, idl.data());
  std::vector<double> glo, ghi;
  tv_row_maps(ord, glo, ghi);
  std::vector<double> dzl(ord), locs(ord + 1), hi(ord * ord), lo(hs * hs), ll(hs + 1), los(hs * hs * hs);
  for (int k = 0; k < nz + 2; k++)
    for (int e = 0; e < nens; e++) {
      if (shared && e > 0) {
        for (int q = 0; q < ord * ord; q++) vs2c[((size_t)k * ord * ord + q) * nens + e] = vs2c[((size_t)k * ord * ord + q) * nens];
        for (int q = 0; q < hs * hs * hs; q++) vwrl[((size_t)k * hs * hs * hs + q) * nens + e] = vwrl[((size_t)k * hs * hs * hs + q) * nens];
        continue;
      }
      for (int kk = 0; kk < ord; kk++) {
        int i1 = -1 + k + kk, i2 = -1 + k;
        i1 = i1 < 0 ? 0 : (i1 > nz - 1 ? nz - 1 : i1);
        i2 = i2 < 0 ? 0 : (i2 > nz - 1 ? nz - 1 : i2);
        dzl[kk] = dz[(size_t)i1 * nens + e] / dz[(size_t)i2 * nens + e];
      }
      locs[0] = 0;
      for (int kk = 1; kk < ord + 1; kk++) locs[kk] = locs[kk - 1] + dzl[kk - 1];
      const double mid = (locs[(ord - 1) / 2] + locs[(ord + 1) / 2]) / 2;
      for (int kk = 0; kk < ord + 1; kk++) locs[kk] = locs[kk] - mid;
      sten_to_coefs_var(ord, locs.data(), hi.data());
      for (int i = 0; i < hs; i++) {
        for (int ii = 0; ii < hs + 1; ii++) ll[ii] = locs[i + ii];
        sten_to_coefs_var(hs, ll.data(), lo.data());
        for (int q = 0; q < hs * hs; q++) los[i * hs * hs + q] = lo[q];
      }
      for (int q = 0; q < ord * ord; q++) vs2c[((size_t)k * ord * ord + q) * nens + e] = hi[q];
      for (int q = 0; q < hs * hs * hs; q++) vwrl[((size_t)k * hs * hs * hs + q) * nens + e] = los[q];
      // bridge matrix of this level: (hi - sum_i idl_i * embed(lo_i)) / idl_m   (WenoLimiter.h:131-138)
      std::vector<double> br(hi);
      for (int i = 0; i < hs; i++)
        for (int s = 0; s < hs; s++)
          for (int ii = 0; ii < hs; ii++) br[(i + s) * ord + ii] -= idl[i] * los[(i * hs + s) * hs + ii];
      for (int q = 0; q < ord * ord; q++) br[q] /= idl[hs];
      const size_t st = shared ? 1 : nens, off = shared ? 0 : e;
      // kernel table in "row" form (awfl_weno.cuh): rlo[i][r][s] = sum_ii glo[r][ii]*lo_i[s][ii], rhi[r][s] = sum_ii ghi[r][ii]*br[s][ii]
      for (int i = 0; i < hs; i++)
        for (int r = 0; r < hs; r++)
          for (int s2 = 0; s2 < hs; s2++) {
            double acc = 0.0;
            for (int ii = 0; ii < hs; ii++) acc += glo[r * hs + ii] * los[(i * hs + s2) * hs + ii];
            vtab[((size_t)k * nel + (i * hs + r) * hs + s2) * st + off] = acc;
          }
      for (int r = 0; r < ord; r++)
        for (int s2 = 0; s2 < ord; s2++) {
          double acc = 0.0;
          for (int ii = 0; ii < ord; ii++) acc += ghi[r * ord + ii] * br[s2 * ord + ii];
          vtab[((size_t)k * nel + hs * hs * hs + r * ord + s2) * st + off] = acc;
        }
    }
}

// NVTX range names of the kernel classes: every launch sits in one (the reference wraps each of its parallel_for launches
// the same way under YAKL_PROFILE, externals/YAKL/src/YAKL_parallel_for_common.h:516-518,555-557). Header-only NVTX v3:
// without a profiler attached a range is a null function-pointer test.
static const char *const kc_names[] = {"awfl:coupler_to_dycore", "awfl:dycore_to_coupler", "awfl:state_march_x", "awfl:state_march_y",
                                       "awfl:state_march_z", "awfl:tracer_march_x", "awfl:tracer_march_y", "awfl:tracer_march_z",
                                       "awfl:fct_multipliers", "awfl:tracer_update", "awfl:other"};

ProfScope::ProfScope(awfl_handle *h_, int cls, cudaStream_t st_) : h(h_), on(h_->prof), st(st_ ? st_ : h_->stream) {
  h->launches++;
  nvtxRangePushA(kc_names[cls]);
  if (!on) return;
  if (h->ev_used + 2 > h->ev_pool.size()) {
    for (int i = 0; i < 2; i++) { cudaEvent_t e; cudaEventCreate(&e); h->ev_pool.push_back(e); }
  }
  h->ev_class.push_back(cls);
  cudaEventRecord(h->ev_pool[h->ev_used], st);
}
ProfScope::~ProfScope() {
  nvtxRangePop();
  if (!on) return;
  cudaEventRecord(h->ev_pool[h->ev_used + 1], st);
  h->ev_used += 2;
}

// Segments per line: a power of two (the grid encodes the segment in the low bits of blockIdx.z, see decode_item),
// chosen so that a segment is about seg_len cells and never shorter than 4.
static int sbits_for(const awfl_handle *h, int n, int seg_len = 0) {
  if (seg_len <= 0) seg_len = h->seg_len;
  int sb = 0;
  while (sb < 4 && (n >> (sb + 1)) >= 4 && (double)n / (1 << sb) > 1.5 * seg_len) sb++;
  return sb;
}
// State march: a segment re-reads and re-converts (x^gamma) its halo cells, so whole lines are cheaper as long as there
// are enough of them to fill the GPU (C2: 6.13 -> 6.04 ms in x, 7.12 -> 7.03 ms in z); a member slab of the host-buffer
// entry point brings few lines and keeps the shorter segments (e2e 1.17 vs 1.27 G with whole lines). The tracer march has
// no such conversion and prefers the shorter segments either way (1.18 vs 1.08 ms).
static long march_lines(const Geom &g, int dir) {
  const long lines = dir == 0 ? (long)g.nz * g.ny : (dir == 1 ? (long)g.nz * g.nx : (long)g.ny * g.nx);
  return lines * ((g.ne + 31) / 32);
}
// A launch that brings too few one-warp blocks to occupy the GPU even once (a standalone CRM: BASELINE.json configs[0] has 50
// lines) is bound by the LATENCY of one warp walking its line: shorter segments, down to 8 cells, until the blocks fill the
// 148 x 12 slots of one wave. The halo cells a segment recomputes cost nothing there. (C1: 1.09 -> see DESIGN.md 5.)
static int sbits_fill(int sb, int n, long warps) {
  while (sb < 4 && (n >> (sb + 1)) >= 8 && (warps << (sb + 1)) <= 148L * 12) sb++;
  return sb;
}
static int sbits_state(const awfl_handle *h, const Geom &g, int dir, int n) {
  const long warps = march_lines(g, dir);
  return sbits_fill(sbits_for(h, n, warps >= 4096 ? 2 * h->seg_len : h->seg_len), n, warps);
}
static int labits_for(const awfl_handle *h, int dir) {
  const int la = dir == 2 ? h->g.ny : h->g.nz;
  int b = 0;
  while ((1 << b) < la) b++;
  return b;
}

#define AWFL_ORD_DISPATCH(CALL)                  \
  switch (h->ord) {                              \
    case 3: OrdLaunch<3>::CALL; break;           \
    case 5: OrdLaunch<5>::CALL; break;           \
    case 7: OrdLaunch<7>::CALL; break;           \
    default: OrdLaunch<9>::CALL; break;          \
  }

// Claim the constant-bank table of (device, order) for this handle (see g_cvtab). Called from awfl_create with the
// handle's device current. Returns true when the handle may read c_vtab.
static bool claim_cvtab(awfl_handle *h, const std::vector<double> &vtab, const std::vector<int> &levels) {
  std::lock_guard<std::mutex> lk(g_cvtab_mu);
  CvtabSlot &slot = g_cvtab[std::make_pair(h->cfg.device, h->ord)];
  if (slot.refs > 0) {
    if (slot.exclusive || slot.table != vtab || slot.levels != levels) return false;
    slot.refs++;
    return true;
  }
  int rc = 0;
  switch (h->ord) {
    case 3: rc = OrdLaunch<3>::upload_cvtab(vtab.data(), vtab.size(), levels.data(), levels.size()); break;
    case 5: rc = OrdLaunch<5>::upload_cvtab(vtab.data(), vtab.size(), levels.data(), levels.size()); break;
    case 7: rc = OrdLaunch<7>::upload_cvtab(vtab.data(), vtab.size(), levels.data(), levels.size()); break;
    default: rc = OrdLaunch<9>::upload_cvtab(vtab.data(), vtab.size(), levels.data(), levels.size()); break;
  }
  if (rc != 0) return false;
  slot.table = vtab;
  slot.levels = levels;
  slot.refs = 1;
  return true;
}
// Paged use rewrites the bank inside every step: the handle must be its only user.
static bool claim_cvtab_exclusive(awfl_handle *h) {
  std::lock_guard<std::mutex> lk(g_cvtab_mu);
  CvtabSlot &slot = g_cvtab[std::make_pair(h->cfg.device, h->ord)];
  if (slot.refs > 0) return false;
  slot.refs = 1; slot.exclusive = true; slot.table.clear(); slot.levels.clear();
  return true;
}
static void release_cvtab(awfl_handle *h) {
  if (!h->use_cvtab && !h->use_cvpage) return;
  std::lock_guard<std::mutex> lk(g_cvtab_mu);
  auto it = g_cvtab.find(std::make_pair(h->cfg.device, h->ord));
  if (it != g_cvtab.end() && --it->second.refs <= 0) {
    // the last reader is gone once its kernels have drained; only then may a later handle overwrite the table
    cudaDeviceSynchronize();
    g_cvtab.erase(it);
  }
  h->use_cvtab = false; h->use_cvpage = false;
}

// home of the vertical matrices a z march reads: 1 constant bank (whole table, or the current page), 3 / 2 global memory
static int vtable_home(const awfl_handle *h, int mode, bool tracer) {
  if (h->use_cvtab || (h->use_cvpage && h->cvpage_ok)) return 1;
  return h->shared_grid && (tracer || mode == 1) ? 3 : 2;
}
static void launch_state_sweep(awfl_handle *h, const Geom &g, int dir, int mode, const SweepArgs &a) {
  ProfScope ps(h, KC_STATE_X + dir);
  AWFL_ORD_DISPATCH(march_state(h->stream, g, h->ph, a, dir, mode, dir == 2 ? vtable_home(h, mode, false) : 0));
}

static void launch_tracer_sweep(awfl_handle *h, const Geom &g, int dir, TracerSweepArgs a, int first_tracer = 0) {
  if (g.nt <= first_tracer) return;
  // the grid carries (tracer, line index, segment) in blockIdx.z, which CUDA limits to 65535: tall grids with many tracers
  // (nz = 128, nx >= 768: 2048 work items per tracer) go in chunks of tracers
  static const int zmax = getenv("AWFL_GRID_Z_MAX") && atoi(getenv("AWFL_GRID_Z_MAX")) > 0 ? atoi(getenv("AWFL_GRID_Z_MAX")) : 65535;   // tests
  int per_launch = zmax >> (a.labits + a.sbits);
  if (per_launch < 1) per_launch = 1;
  const int vt = dir == 2 ? vtable_home(h, 1, true) : 0;
  // Two tracers per warp on one line share the line's density, its reciprocal and the mass flux (h->tracer_pairs); an odd
  // tracer is left for a launch of the two-lines-per-warp form.
  int t0 = first_tracer;
  if (h->tracer_pairs) {
    const int npair_total = (g.nt - first_tracer) / 2;
    for (int p0 = 0; p0 < npair_total; p0 += per_launch) {
      const int np = npair_total - p0 < per_launch ? npair_total - p0 : per_launch;
      a.tr0 = first_tracer + 2 * p0; a.ntr = 2 * np;
      ProfScope ps(h, KC_TRACER_X + dir);
      AWFL_ORD_DISPATCH(march_tracer(h->stream, g, a, dir, vt, true));
    }
    t0 = first_tracer + 2 * npair_total;
  }
  for (; t0 < g.nt; t0 += per_launch) {
    a.tr0 = t0; a.ntr = g.nt - t0 < per_launch ? g.nt - t0 : per_launch;
    ProfScope ps(h, KC_TRACER_X + dir);
    AWFL_ORD_DISPATCH(march_tracer(h->stream, g, a, dir, vt, false));
  }
}

// Page p of the vertical table: the levels segment p of the z marches reads (cells c0-1 .. c1 -> levels c0 .. c1+1).
static int upload_cvpage(awfl_handle *h, int p) {
  const int n = h->g.nz, sb = h->cvpage_sbits;
  const int c0 = (p * n) >> sb, c1 = ((p + 1) * n) >> sb;
  const size_t nel = (size_t)h->hs * h->hs * h->hs + (size_t)h->ord * h->ord;
  const double *src = h->vtab_pin + (size_t)c0 * nel;
  const size_t cnt = (size_t)(c1 + 2 - c0) * nel;
  const int *lv = h->vlvl_pin + (size_t)p * (n + 2);
  int rc = 0;
  switch (h->ord) {
    case 3: rc = OrdLaunch<3>::upload_cvpage(h->stream, src, cnt, lv, n + 2); break;
    case 5: rc = OrdLaunch<5>::upload_cvpage(h->stream, src, cnt, lv, n + 2); break;
    case 7: rc = OrdLaunch<7>::upload_cvpage(h->stream, src, cnt, lv, n + 2); break;
    default: rc = OrdLaunch<9>::upload_cvpage(h->stream, src, cnt, lv, n + 2); break;
  }
  return rc;
}

// One semi-discrete right-hand side + stage combination.
//   qin: stage input; q0: step-start state; qout: stage output.  mode 1: RK update, mode 2: tendencies to (st_out,tr_out)
static void run_stage(awfl_handle *h, const Geom &g, const double *qin, const double *q0, double *qout, double a, double b, double dt_stage,
                      double dt_full, double sa, double sb, int mode, const double *fct_start, double *st_out, double *tr_out) {
  const int ndir = g.sim2d ? 2 : 3;
  const int dirs3[3] = {0, 1, 2}, dirs2[2] = {0, 2};
  const int *dirs = g.sim2d ? dirs2 : dirs3;
  const int nn[3] = {g.nx, g.ny, g.nz};
  for (int d = 0; d < ndir; d++) {
    const int dir = dirs[d];
    const bool last = d == ndir - 1;
    SweepArgs s;
    s.qin = qin; s.qbase = q0; s.qout = mode == 2 ? st_out : qout; s.tend = h->tend; s.mflux = h->mflux[dir];
    s.dz = h->dz; s.grav_var = h->grav_var; s.vtab = h->vtab; s.rdz = h->rdz; s.pcache = h->pcache; s.vst = h->shared_grid ? 1 : g.nens;
    s.a = a; s.b = b; s.c = b * dt_full; s.first = d == 0; s.sbits = sbits_state(h, g, dir, nn[dir]); s.labits = labits_for(h, dir);
    s.seg_only = -1; s.general_2d = h->general_2d;
    // 2-D: the first tracer rides in the state march, in the slot the reconstruction of v leaves free (awfl_march.cuh RingF)
    const int fused = g.sim2d && !h->general_2d && g.nt > 0 && (h->fused_tracer >> (dir == 2 ? 1 : 0) & 1) ? 1 : 0;
    s.fused_tracer = fused; s.tflux = h->tflux[dir];
    TracerSweepArgs t;
    t.qin = qin; t.mflux = h->mflux[dir]; t.tflux = h->tflux[dir]; t.vtab = h->vtab; t.vst = s.vst; t.sbits = sbits_fill(sbits_for(h, nn[dir]), nn[dir], (march_lines(g, dir) + 1) / 2 * g.nt); t.labits = s.labits;
    t.seg_only = -1;
    if (dir == 2 && h->use_cvpage && h->cvpage_ok) {
      // paged vertical table: segment by segment, the state and the tracer march of a page behind its upload (stream order
      // keeps a page in place until the launches that read it are done)
      s.sbits = t.sbits = h->cvpage_sbits;
      for (int p = 0; p < (1 << h->cvpage_sbits); p++) {
        if (upload_cvpage(h, p) != 0) { h->err = "upload of a vertical-table page failed"; return; }
        s.seg_only = t.seg_only = p;
        launch_state_sweep(h, g, dir, last ? mode : 0, s);
        launch_tracer_sweep(h, g, dir, t, fused);
      }
      continue;
    }
    launch_state_sweep(h, g, dir, last ? mode : 0, s);
    launch_tracer_sweep(h, g, dir, t, fused);
  }
  if (g.nt > 0 && g.sim2d && mode == 1 && !fct_start && h->fused_fct) {
    // fused FCT + update (2-D): strips of W columns, enough level segments for a few waves of one-warp blocks
    // W = 4, or 8 where the strips divide the line (6 halo-inclusive multipliers per 4 results against 10 per 8: 7.21 -> 7.01 ms
    // at C5; ragged strips lose it again, profiles/r02u_fct_variants.txt)
    // (with few tracers the narrower strips' larger grid wins: C2, one tracer, 0.90 against 0.93 ms)
    const int W = (g.nx % 8 == 0 && g.nx >= 16 && g.nt >= 4) ? 8 : AWFL_FCT_STRIP;
    FctUpdArgs u;
    u.q0 = q0; u.qs = qin; u.qout = qout; u.tfx = h->tflux[0]; u.tfz = h->tflux[2]; u.dz = h->dz; u.rdz = h->rdz;
    u.positive = h->positive_d; u.a = a; u.b = b; u.c = b * dt_full; u.dt = dt_stage; u.rdx = 1.0 / g.dx;
    const unsigned tiles = (unsigned)(g.ne + 31) / 32, strips = (unsigned)(g.nx + W - 1) / W;
    const long per_seg = (long)tiles * strips * g.nt;
    long ks = (16000 + per_seg - 1) / per_seg;
    if (ks > g.nz / 8) ks = g.nz / 8;
    if (ks < 1) ks = 1;
    if (const char *ev = getenv("AWFL_FCT_KSEGS")) { const int v = atoi(ev); if (v >= 1 && v <= g.nz) ks = v; }   // experiments
    u.ksegs = (int)ks;
    ProfScope ps(h, KC_TRACER_UPD);
    if (W == 8) k_fct_update_2d<8><<<dim3(tiles, strips, (unsigned)(g.nt * ks)), 32, 0, h->stream>>>(g, u);
    else k_fct_update_2d<AWFL_FCT_STRIP><<<dim3(tiles, strips, (unsigned)(g.nt * ks)), 32, 0, h->stream>>>(g, u);
  } else if (g.nt > 0) {
    const RowMap rm = row_map(g, 256);
    const unsigned blocks = row_blocks(g, rm);
    FctArgs f;
    f.q0 = q0; f.qs = qin; f.tfx = h->tflux[0]; f.tfy = h->tflux[1]; f.tfz = h->tflux[2]; f.fct_start = fct_start;
    f.mult = h->mult; f.dz = h->dz; f.rdz = h->rdz; f.rdx = 1.0 / g.dx; f.rdy = 1.0 / g.dy; f.positive = h->positive_d; f.sa = sa; f.sb = sb; f.dt = dt_stage;
    { ProfScope ps(h, KC_FCT); k_fct_mult<<<dim3(blocks, g.nt), 256, 0, h->stream>>>(g, f, rm); }
    TracerUpdArgs u;
    u.q0 = q0; u.qs = qin; u.qout = mode == 2 ? tr_out : qout; u.tfx = h->tflux[0]; u.tfy = h->tflux[1]; u.tfz = h->tflux[2];
    u.mult = h->mult; u.rdz = h->rdz; u.rdx = f.rdx; u.rdy = f.rdy; u.positive = h->positive_d; u.a = a; u.b = b; u.c = b * dt_full; u.tend_only = mode == 2;
    { ProfScope ps(h, KC_TRACER_UPD); k_tracer_update<<<dim3(blocks, g.nt), 256, 0, h->stream>>>(g, u, rm); }
  }
}

static int make_field_ptrs(awfl_handle *h, const awfl_fields *f, FieldPtrs &fp) {
  fp.f[0] = f->density_dry; fp.f[1] = f->uvel; fp.f[2] = f->vvel; fp.f[3] = f->wvel; fp.f[4] = f->temp;
  fp.tr = h->trptr_d;
  if (h->g.nt > 0) CK(cudaMemcpyAsync(h->trptr_d, f->tracers, sizeof(double *) * h->g.nt, cudaMemcpyHostToDevice, h->stream));
  return 0;
}

extern "C" {

const char *awfl_dycore_name(void) { return "SSPRK3+WENO+FV A-grid"; }

const char *awfl_last_error(const awfl_handle *h) { return h ? h->err.c_str() : g_create_err.c_str(); }

long awfl_launch_count(const awfl_handle *h) { return h->launches; }

int awfl_fused_tracer(const awfl_handle *h) { return h->g.sim2d && !h->general_2d && h->g.nt > 0 ? h->fused_tracer : 0; }

int awfl_table_home(const awfl_handle *h) { return h->use_cvtab ? 1 : (h->use_cvpage ? 4 : (h->shared_grid ? 3 : 2)); }

int awfl_create(const awfl_config *cfg, awfl_handle **out) {
  *out = nullptr;
  if (!cfg || cfg->nz < 1 || cfg->ny < 1 || cfg->nx < 1 || cfg->nens < 1 || cfg->num_tracers < 0) { g_create_err = "awfl_create: bad dimensions"; return 1; }
  if (cfg->ord != 3 && cfg->ord != 5 && cfg->ord != 7 && cfg->ord != 9) { g_create_err = "awfl_create: ord must be 3, 5, 7 or 9"; return 1; }
  const int hs0 = (cfg->ord + 1) / 2;
  if ((long)cfg->nx * cfg->nens >= (1L << 31) || ((long)cfg->nx * cfg->nens + 255) / 256 * cfg->nz * cfg->ny >= (1L << 31)) {
    g_create_err = "awfl_create: nx*nens and the block count of one field must fit 31 bits"; return 1;
  }
  if ((double)cfg->nz * cfg->ny * cfg->nx * cfg->nens >= 4294967296.0) { g_create_err = "awfl_create: one field must hold fewer than 2^32 cells (34 GB)"; return 1; }
  {
    int lab = 0;
    while ((1 << lab) < (cfg->nz > cfg->ny ? cfg->nz : cfg->ny)) lab++;
    if ((1L << (lab + 4)) > 65535) { g_create_err = "awfl_create: nz and ny must not exceed 2048 (line index + segment bits of a march must fit gridDim.z)"; return 1; }
  }
  if (cfg->nx < hs0 || (cfg->ny > 1 && cfg->ny < hs0) || cfg->nz < hs0) { g_create_err = "awfl_create: each active dimension needs at least (ord+1)/2 cells"; return 1; }
  if (!cfg->dz_host) { g_create_err = "awfl_create: dz_host is NULL"; return 1; }
  if (cfg->num_tracers > 0 && (!cfg->tracer_positive || !cfg->tracer_adds_mass)) { g_create_err = "awfl_create: tracer flag arrays are NULL"; return 1; }
  if (cfg->num_tracers > 0 && (cfg->idWV < 0 || cfg->idWV >= cfg->num_tracers)) { g_create_err = "awfl_create: idWV out of range (a water_vapor tracer is required, Dycore.h:83)"; return 1; }
  if (cfg->num_tracers == 0) { g_create_err = "awfl_create: the dycore needs the water_vapor tracer (Dycore.h:83,1376)"; return 1; }
  awfl_handle *h = new awfl_handle();
  h->cfg = *cfg;
  h->launches = 0;
  h->use_cvpage = false; h->cvpage_ok = false; h->cvpage_sbits = 0; h->vtab_pin = nullptr; h->vlvl_pin = nullptr;
  h->prof = false; h->ev_used = 0;
  h->ord = cfg->ord; h->hs = hs0;
  Geom &g = h->g;
  g.nz = cfg->nz; g.ny = cfg->ny; g.nx = cfg->nx; g.nens = cfg->nens; g.nt = cfg->num_tracers;
  g.sim2d = cfg->ny == 1;
  g.ncell = (long)g.nz * g.ny * g.nx * g.nens;
  g.e0 = 0; g.ne = g.nens;
  g.dx = cfg->xlen / g.nx; g.dy = cfg->ylen / g.ny;
  // Dycore.h:871-891
  const double R_d = cfg->R_d, cp_d = cfg->cp_d;
  h->cv_d = cp_d - R_d;
  const double gamma = cfg->gamma_d != 0.0 ? cfg->gamma_d : cp_d / h->cv_d;
  const double kappa = cfg->kappa_d != 0.0 ? cfg->kappa_d : R_d / cp_d;
  const double C0 = cfg->C0 != 0.0 ? cfg->C0 : pow(R_d * pow(cfg->p0, -kappa), gamma);
  h->kappa = kappa;
  h->ph.R_d = R_d; h->ph.R_v = cfg->R_v; h->ph.C0 = C0; h->ph.gamma = gamma; h->ph.grav = cfg->grav;
  h->seg_len = 32;
  if (const char *s = getenv("AWFL_SEG_LEN")) { int v = atoi(s); if (v >= 4) h->seg_len = v; }
  auto bail = [&](const std::string &m) { g_create_err = m; awfl_destroy(h); return 1; };
  DeviceGuard dev_guard_(cfg->device);
  if (dev_guard_.err != cudaSuccess) {
    g_create_err = std::string("cudaSetDevice: ") + cudaGetErrorString(dev_guard_.err) + " (no CPU fallback exists)";
    delete h;
    return 1;
  }
  // NULL = the legacy default stream, which is where the reference (YAKL) enqueues everything and which orders
  // with the caller's own default-stream work (copies, other modules)
  h->own_stream = false;
  h->stream = (cudaStream_t)cfg->stream;
  const int nt = g.nt, N = NS + nt;
  h->positive_h.assign(cfg->tracer_positive, cfg->tracer_positive + nt);
  h->adds_mass_h.assign(cfg->tracer_adds_mass, cfg->tracer_adds_mass + nt);
  // shared vertical grid?
  h->shared_grid = true;
  for (int k = 0; k < g.nz && h->shared_grid; k++)
    for (int e = 1; e < g.nens; e++)
      if (cfg->dz_host[(size_t)k * g.nens + e] != cfg->dz_host[(size_t)k * g.nens]) { h->shared_grid = false; break; }
  if (getenv("AWFL_FORCE_PER_ENS_GRID")) h->shared_grid = false;
  std::vector<double> vs2c, vwrl, vtab;
  build_vertical(h->ord, g.nz, g.nens, cfg->dz_host, h->shared_grid, vs2c, vwrl, vtab);
  // the marching kernels opt in to their shared-memory rings on THIS device (a per-device, per-kernel attribute)
  switch (h->ord) {
    case 3: OrdLaunch<3>::configure(); break;
    case 5: OrdLaunch<5>::configure(); break;
    case 7: OrdLaunch<7>::configure(); break;
    default: OrdLaunch<9>::configure(); break;
  }
  if (h->shared_grid && !getenv("AWFL_NO_CVTAB")) {
    // constant-bank copy: levels with bit-identical tables share one (an equally spaced grid has a single table at any order)
    const size_t nel = (size_t)h->hs * h->hs * h->hs + (size_t)h->ord * h->ord;
    std::vector<double> ctab;
    std::vector<int> lvl(g.nz + 2);
    for (int k = 0; k < g.nz + 2; k++) {
      const double *tk = vtab.data() + (size_t)k * nel;
      int found = -1;
      for (size_t u = 0; u + nel <= ctab.size() && found < 0; u += nel)
        if (memcmp(ctab.data() + u, tk, nel * sizeof(double)) == 0) found = (int)u;
      if (found < 0) { found = (int)ctab.size(); ctab.insert(ctab.end(), tk, tk + nel); }
      lvl[k] = found;
    }
    h->use_cvtab = ctab.size() <= OrdLaunch<5>::cvtab_capacity() && lvl.size() <= OrdLaunch<5>::cvtab_max_levels() && claim_cvtab(h, ctab, lvl);
    if (!h->use_cvtab && !getenv("AWFL_NO_CVPAGE") && lvl.size() <= OrdLaunch<5>::cvtab_max_levels()) {
      // too large as a whole (ord 9 at 58 levels: 60 x 206 doubles = 97 KB): page it by segments of the vertical marches
      int sb = 1;
      auto fits = [&](int b) {
        for (int p = 0; p < (1 << b); p++) {
          const int c0 = (p * g.nz) >> b, c1 = ((p + 1) * g.nz) >> b;
          if ((size_t)(c1 + 2 - c0) * nel > OrdLaunch<5>::cvtab_capacity()) return false;
        }
        return true;
      };
      while (sb < 4 && !fits(sb)) sb++;
      if (fits(sb) && (g.nz >> sb) >= 4 && claim_cvtab_exclusive(h)) {
        h->use_cvpage = true; h->cvpage_ok = true; h->cvpage_sbits = sb;
        const int P = 1 << sb;
        if (cudaMallocHost((void **)&h->vtab_pin, sizeof(double) * vtab.size()) != cudaSuccess ||
            cudaMallocHost((void **)&h->vlvl_pin, sizeof(int) * (size_t)P * (g.nz + 2)) != cudaSuccess) return bail("cudaMallocHost failed");
        memcpy(h->vtab_pin, vtab.data(), sizeof(double) * vtab.size());
        for (int p = 0; p < P; p++) {
          const int c0 = (p * g.nz) >> sb, c1 = ((p + 1) * g.nz) >> sb;
          for (int k = 0; k < g.nz + 2; k++) h->vlvl_pin[(size_t)p * (g.nz + 2) + k] = (k >= c0 && k <= c1 + 1) ? (int)((k - c0) * nel) : 0;
        }
      }
    }
  }
  const size_t fb = sizeof(double) * (size_t)g.ncell;
  const size_t lvl = sizeof(double) * (size_t)g.ny * g.nx * g.nens;
#define ALLOC(ptr, bytes) do { if (cudaMalloc((void **)&(ptr), (bytes)) != cudaSuccess) return bail(std::string("cudaMalloc failed for ") + #ptr); } while (0)
  for (int b = 0; b < 3; b++) ALLOC(h->Q[b], fb * N);
  ALLOC(h->tend, fb * NS);
  ALLOC(h->pcache, fb);
  ALLOC(h->mflux[0], fb); ALLOC(h->mflux[1], fb); ALLOC(h->mflux[2], fb + lvl);
  ALLOC(h->tflux[0], fb * nt); ALLOC(h->tflux[1], fb * nt); ALLOC(h->tflux[2], (fb + lvl) * nt);
  ALLOC(h->mult, fb * nt);
  ALLOC(h->dz, sizeof(double) * g.nz * g.nens);
  ALLOC(h->rdz, sizeof(double) * g.nz * g.nens);
  ALLOC(h->grav_var, sizeof(double) * g.nz * g.nens);
  ALLOC(h->vtab, sizeof(double) * vtab.size());
  ALLOC(h->vs2c, sizeof(double) * vs2c.size());
  ALLOC(h->vwrl, sizeof(double) * vwrl.size());
  ALLOC(h->pint, fb + lvl);
  ALLOC(h->dt_min_d, sizeof(double));
  ALLOC(h->mass_d, sizeof(double) * (nt + 2) * g.nens);
  ALLOC(h->positive_d, nt); ALLOC(h->adds_mass_d, nt);
  ALLOC(h->trptr_d, sizeof(double *) * nt);
  ALLOC(h->trptr_out_d, sizeof(double *) * nt);
  h->stage_fields = nullptr; h->stage_out = nullptr; h->tmp_state = nullptr; h->halo_pressure = nullptr;
  { const char *ev = getenv("AWFL_FUSED_FCT"); h->fused_fct = !(ev && atoi(ev) == 0); }
  { const char *ev = getenv("AWFL_GENERAL_2D"); h->general_2d = ev && atoi(ev) == 1; }
  // measured (profiles/r02G_fused_tracer_ab.txt): both marches up to ord 7; at ord 9 the vertical march, already at the
  // register cap with its dense per-level tables, loses more than the tracer march costs
  h->fused_tracer = cfg->ord <= AWFL_FUSED_TRACER_Z_MAX_ORD ? 3 : 1;
  { const char *ev = getenv("AWFL_FUSED_TRACER"); if (ev) h->fused_tracer = atoi(ev) & 3; }
  { const char *ev = getenv("AWFL_TRACER_PAIRS"); h->tracer_pairs = !(ev && atoi(ev) == 0); }
  h->s_in = nullptr; h->s_out = nullptr; h->host_last_ncycles = 0; h->host_mispredictions = 0;
  {
    // member slabs of awfl_time_step_host: chosen from the grid by default (host_slab_count); AWFL_HOST_SLABS overrides
    const char *ev = getenv("AWFL_HOST_SLABS");
    h->host_slabs = ev && atoi(ev) > 0 ? atoi(ev) : 0;
    const char *ec = getenv("AWFL_HOST_CSTREAMS");
    h->host_cstreams = ec && atoi(ec) > 0 ? atoi(ec) : 3;
  }
  if (cudaMallocHost((void **)&h->dt_min_h, sizeof(double)) != cudaSuccess) return bail("cudaMallocHost failed");
  cudaMemcpy(h->dz, cfg->dz_host, sizeof(double) * g.nz * g.nens, cudaMemcpyHostToDevice);
  {
    std::vector<double> rdz((size_t)g.nz * g.nens);
    for (size_t i = 0; i < rdz.size(); i++) rdz[i] = 1.0 / cfg->dz_host[i];
    cudaMemcpy(h->rdz, rdz.data(), sizeof(double) * rdz.size(), cudaMemcpyHostToDevice);
  }
  cudaMemset(h->grav_var, 0, sizeof(double) * g.nz * g.nens);   // the reference leaves it uninitialised
  cudaMemcpy(h->vtab, vtab.data(), sizeof(double) * vtab.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(h->vs2c, vs2c.data(), sizeof(double) * vs2c.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(h->vwrl, vwrl.data(), sizeof(double) * vwrl.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(h->positive_d, h->positive_h.data(), nt, cudaMemcpyHostToDevice);
  cudaMemcpy(h->adds_mass_d, h->adds_mass_h.data(), nt, cudaMemcpyHostToDevice);
  if (cudaDeviceSynchronize() != cudaSuccess) return bail(std::string("awfl_create: ") + cudaGetErrorString(cudaGetLastError()));
  *out = h;
  return 0;
}

void awfl_destroy(awfl_handle *h) {
  if (!h) return;
  DeviceGuard dev_guard_(h->cfg.device);
  for (int b = 0; b < 3; b++) cudaFree(h->Q[b]);
  cudaFree(h->tend);
  for (int d = 0; d < 3; d++) { cudaFree(h->mflux[d]); cudaFree(h->tflux[d]); }
  release_cvtab(h);
  cudaFree(h->mult); cudaFree(h->dz); cudaFree(h->rdz); cudaFree(h->grav_var); cudaFree(h->vtab); cudaFree(h->vs2c); cudaFree(h->vwrl);
  cudaFree(h->pint); cudaFree(h->dt_min_d); cudaFree(h->mass_d); cudaFree(h->positive_d); cudaFree(h->adds_mass_d);
  cudaFree(h->pcache); cudaFree(h->halo_pressure); cudaFree(h->trptr_d); cudaFree(h->trptr_out_d); cudaFree(h->stage_fields); cudaFree(h->stage_out); cudaFree(h->tmp_state);
  if (h->s_in) { cudaStreamDestroy(h->s_in); cudaStreamDestroy(h->s_out); for (int c = 0; c < 4; c++) cudaStreamDestroy(h->s_cmp[c]); }
  for (cudaEvent_t e : h->slab_ev) cudaEventDestroy(e);
  if (h->dt_min_h) cudaFreeHost(h->dt_min_h);
  if (h->vtab_pin) cudaFreeHost(h->vtab_pin);
  if (h->vlvl_pin) cudaFreeHost(h->vlvl_pin);
  for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
  if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

double awfl_get_option(const awfl_handle *h, const char *name) {
  const std::string n(name);
  if (n == "R_d") return h->ph.R_d;
  if (n == "cp_d") return h->cfg.cp_d;
  if (n == "R_v") return h->ph.R_v;
  if (n == "cp_v") return h->cfg.cp_v;
  if (n == "p0") return h->cfg.p0;
  if (n == "grav") return h->ph.grav;
  if (n == "cv_d") return h->cv_d;
  if (n == "gamma_d") return h->ph.gamma;
  if (n == "kappa_d") return h->kappa;
  if (n == "C0") return h->ph.C0;
  return NAN;
}

double *awfl_variable_gravity(awfl_handle *h) { return h->grav_var; }
const double *awfl_vert_sten_to_coefs(awfl_handle *h) { return h->vs2c; }
const double *awfl_vert_weno_recon_lower(awfl_handle *h) { return h->vwrl; }

int awfl_copy_to_host(void *dst_host, const void *src_device, unsigned long bytes) {
  return cudaMemcpy(dst_host, src_device, bytes, cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : 1;
}
int awfl_copy_to_device(void *dst_device, const void *src_host, unsigned long bytes) {
  return cudaMemcpy(dst_device, src_host, bytes, cudaMemcpyHostToDevice) == cudaSuccess ? 0 : 1;
}
int awfl_device_alloc(void **ptr, unsigned long bytes) {
  if (cudaMalloc(ptr, bytes) != cudaSuccess) { *ptr = nullptr; return 1; }
  return cudaMemset(*ptr, 0, bytes) == cudaSuccess ? 0 : 1;
}
void awfl_device_free(void *ptr) { if (ptr) cudaFree(ptr); }
int awfl_current_device(void) {
  int d = 0;
  return cudaGetDevice(&d) == cudaSuccess ? d : 0;
}

int awfl_set_variable_gravity(awfl_handle *h, const double *variable_gravity) {
  ENTER(h);
  if (!variable_gravity) return fail(h, "awfl_set_variable_gravity: NULL pointer");
  if (variable_gravity == h->grav_var) return 0;
  CK(cudaMemcpyAsync(h->grav_var, variable_gravity, sizeof(double) * h->g.nz * h->g.nens, cudaMemcpyDeviceToDevice, h->stream));
  return 0;
}

int awfl_export_tables(awfl_handle *h, double *variable_gravity, double *vert_sten_to_coefs, double *vert_weno_recon_lower) {
  ENTER(h);
  const Geom &g = h->g;
  const size_t o = h->ord, hs = h->hs, lv = (size_t)(g.nz + 2) * g.nens;
  if (variable_gravity) CK(cudaMemcpyAsync(variable_gravity, h->grav_var, sizeof(double) * g.nz * g.nens, cudaMemcpyDeviceToDevice, h->stream));
  if (vert_sten_to_coefs) CK(cudaMemcpyAsync(vert_sten_to_coefs, h->vs2c, sizeof(double) * lv * o * o, cudaMemcpyDeviceToDevice, h->stream));
  if (vert_weno_recon_lower) CK(cudaMemcpyAsync(vert_weno_recon_lower, h->vwrl, sizeof(double) * lv * hs * hs * hs, cudaMemcpyDeviceToDevice, h->stream));
  return 0;
}

int awfl_compute_time_step(awfl_handle *h, const awfl_fields *f, double cfl, double *dt_host) {
  ENTER(h);
  FieldPtrs fp;
  if (make_field_ptrs(h, f, fp)) return 1;
  const double inf = INFINITY;
  CK(cudaMemcpyAsync(h->dt_min_d, &inf, sizeof(double), cudaMemcpyHostToDevice, h->stream));
  const RowMap rm = row_map(h->g, 256);
  k_cfl_only<<<row_blocks(h->g, rm), 256, 0, h->stream>>>(h->g, h->ph, fp, h->cfg.idWV, cfl, h->dz, h->dt_min_d, rm);
  h->launches++;
  CK(cudaMemcpyAsync(h->dt_min_h, h->dt_min_d, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  *dt_host = *h->dt_min_h;
  return 0;
}

// The three pieces of Dycore::timeStep (Dycore.h:107-255) on the member slab `g` describes (the whole ensemble, or one
// slab of the host-buffer entry point); all work goes to h->stream.
// (1) convert_coupler_to_dynamics + tracer clip + CFL minimum (Dycore.h:128-134,141)
static void step_convert_in(awfl_handle *h, const Geom &g, const FieldPtrs &fp, bool need_dt, cudaStream_t st = nullptr) {
  if (!st) st = h->stream;
  ProfScope ps(h, KC_IN, st);
  const RowMap rm = row_map(g, 256);
  k_coupler_to_dycore<<<row_blocks(g, rm), 256, 0, st>>>(g, h->ph, fp, h->cfg.idWV, h->adds_mass_d, h->positive_d, h->Q[0], 0.8,
                                                                h->dz, need_dt ? h->dt_min_d : nullptr, rm);
}
// (2) read the CFL minimum back and derive the sub-cycle count (Dycore.h:144): the only host synchronisation of a step
static int step_ncycles(awfl_handle *h, double crm_dt, int *ncycles, cudaStream_t st = nullptr) {
  if (!st) st = h->stream;
  CK(cudaMemcpyAsync(h->dt_min_h, h->dt_min_d, sizeof(double), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const double dt_dyn = *h->dt_min_h;
  if (!(dt_dyn > 0.0) || !isfinite(dt_dyn)) return fail(h, "awfl_time_step: non-finite CFL time step (NaN or non-positive state)");
  *ncycles = (int)ceil(crm_dt / dt_dyn);
  return 0;
}
// (3) ncycles SSPRK3 sub-cycles (Dycore.h:147-222). The step state enters in Q[0]; the three buffers rotate -- stage 1:
// C -> S1, stage 2: S1 (+C) -> S2, stage 3: S2 (+C) -> S1, which becomes the next C -- so that no stage writes the buffer
// another block of the same launch still reads (the fused FCT kernel reads halo columns of the step-start tracers).
// Returns the index of the buffer that holds the result.
static int step_advance(awfl_handle *h, const Geom &g, int ncycles, double dt) {
  int c = 0;
  for (int n = 0; n < ncycles; n++) {
    const int s1 = (c + 1) % 3, s2 = (c + 2) % 3;
    run_stage(h, g, h->Q[c], h->Q[c], h->Q[s1], 0.0, 1.0, dt, dt, 1.0, 0.0, 1, nullptr, nullptr, nullptr);
    run_stage(h, g, h->Q[s1], h->Q[c], h->Q[s2], 3.0 / 4.0, 1.0 / 4.0, (1.0 / 4.0) * dt, dt, 3.0 / 4.0, 1.0 / 4.0, 1, nullptr, nullptr, nullptr);
    run_stage(h, g, h->Q[s2], h->Q[c], h->Q[s1], 1.0 / 3.0, 2.0 / 3.0, (2.0 / 3.0) * dt, dt, 1.0 / 3.0, 2.0 / 3.0, 1, nullptr, nullptr, nullptr);
    c = s1;
  }
  return c;
}
// (4) convert_dynamics_to_coupler (Dycore.h:1281-1331)
static void step_convert_out(awfl_handle *h, const Geom &g, const FieldPtrs &fp, int qi, cudaStream_t st = nullptr) {
  if (!st) st = h->stream;
  ProfScope ps(h, KC_OUT, st);
  const RowMap rm = row_map(g, 256);
  k_dycore_to_coupler<<<row_blocks(g, rm), 256, 0, st>>>(g, h->ph, fp, h->cfg.idWV, h->adds_mass_d, h->Q[qi], rm);
}

// module-level NVTX range of a public entry point (the reference's per-module timers: pam_core/pam_coupler.h:144-150)
struct NvtxRange {
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

int awfl_time_step(awfl_handle *h, const awfl_fields *f, double crm_dt, int ncycles_override, int *ncycles_out) {
  ENTER(h);
  NvtxRange nvtx_("awfl_time_step");
  if (!(crm_dt > 0.0)) return fail(h, "awfl_time_step: crm_dt must be positive");
  FieldPtrs fp;
  if (make_field_ptrs(h, f, fp)) return 1;
  const Geom &g = h->g;
  const bool need_dt = ncycles_override <= 0;
  if (need_dt) {
    const double inf = INFINITY;
    CK(cudaMemcpyAsync(h->dt_min_d, &inf, sizeof(double), cudaMemcpyHostToDevice, h->stream));
  }
  step_convert_in(h, g, fp, need_dt);
  int ncycles = ncycles_override;
  if (need_dt && step_ncycles(h, crm_dt, &ncycles)) return 1;
  const double dt = crm_dt / ncycles;       // Dycore.h:145
  const int qi = step_advance(h, g, ncycles, dt);
  step_convert_out(h, g, fp, qi);
  CK(cudaGetLastError());
  if (ncycles_out) *ncycles_out = ncycles;
  return 0;
}

int awfl_set_host_slabs(awfl_handle *h, int nslabs) {
  if (nslabs < 0) return fail(h, "awfl_set_host_slabs: need at least one slab (0 = choose from the grid)");
  h->host_slabs = nslabs;
  return 0;
}

// Slab count of the host-buffer entry point. The marching kernels run one-warp blocks whose run time hardly depends on
// how full the SM is, so a slab costs a whole wave however few warps it brings: the slabs in flight (one per compute
// stream) should together fill the resident warps of the state march exactly once -- 148 SMs x 12 warps at the 168-register
// budget -- and not spill into a second, nearly empty wave. Measured on C2: 13 slabs of 10 tiles on 3
// streams (3 x 10 x 58 lines = 1740 of 1776 slots).
static int host_slab_count(const awfl_handle *h) {
  const Geom &g = h->g;
  const int tiles = (g.nens + 31) / 32;
  if (h->host_slabs > 0) return h->host_slabs < tiles ? h->host_slabs : tiles;
  const int nc = h->host_cstreams > 4 ? 4 : (h->host_cstreams < 1 ? 1 : h->host_cstreams);
  const long lines = (long)g.nz * g.ny;                 // x-march blocks per member tile
  long per_slab = (148L * 12) / (nc * lines);
  if (per_slab < 1) per_slab = 1;
  long P = (tiles + per_slab - 1) / per_slab;
  if (P > 16) P = 16;
  return (int)P;
}
int awfl_host_slabs(const awfl_handle *h) { return host_slab_count(h); }

// Dycore::timeStep on HOST fields, pipelined over member slabs (multiples of 32 members; a slab is a strided column
// block of every (cells, nens) field: cudaMemcpy2DAsync, row = the slab's members of one cell).
//   copy-in stream : H2D of slab s, then its conversion kernel (+ CFL minimum)        s = 0..P-1, then the dt read-back
//   compute stream : the SSPRK3 sub-cycles of slab s as soon as its conversion is done
//   copy-out stream: dycore -> coupler conversion of slab s into the RESULT staging, then its D2H
// The reference takes dt from the whole coupler (Dycore.h:141-145), so the sub-cycle count is only known when the LAST
// slab has arrived. To overlap the copies with the kernels anyway the slabs run with a PREDICTED count -- the caller's
// ncycles_override, else the count the previous call derived -- all the way to the host buffers, and the prediction is
// checked against the CFL minimum over all members afterwards. The staged inputs stay untouched on the device (results
// go to a second staging buffer), so a wrong prediction costs one plain, un-overlapped step from those inputs, whose
// results then overwrite what the speculative pass wrote to the host buffers; it never changes a result. Without a
// prediction (first call) the slabs wait for the dt.
// Results are identical to awfl_time_step on the same fields for any slab count (members never interact).
int awfl_time_step_host(awfl_handle *h, double *const *host_fields, double crm_dt, int ncycles_override, int *ncycles_out) {
  ENTER(h);
  NvtxRange nvtx_("awfl_time_step_host");
  if (!(crm_dt > 0.0)) return fail(h, "awfl_time_step_host: crm_dt must be positive");
  // AWFL_HOST_TRACE=1: device timeline of this call on stderr (timing events on the streams; diagnostic only)
  static const bool trace = getenv("AWFL_HOST_TRACE") && atoi(getenv("AWFL_HOST_TRACE")) > 0;
  std::vector<std::pair<std::string, cudaEvent_t>> marks;
  auto mark = [&](const char *what, int s, cudaStream_t st) {
    if (!trace) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    marks.push_back({std::string(what) + "[" + std::to_string(s) + "]", e});
  };
  const Geom &g = h->g;
  const int N = NS + g.nt;
  const size_t fb = sizeof(double) * (size_t)g.ncell;
  if (!h->stage_fields) CK(cudaMalloc((void **)&h->stage_fields, fb * N));
  if (!h->stage_out) CK(cudaMalloc((void **)&h->stage_out, fb * N));
  const int tiles = (g.nens + 31) / 32;
  const int P = host_slab_count(h);
  if (!h->s_in) {
    CK(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
    for (int c = 0; c < 4; c++) CK(cudaStreamCreateWithFlags(&h->s_cmp[c], cudaStreamNonBlocking));
  }
  // events per slab: 3s = inputs on the device, 3s+1 = converted (its CFL minimum is in), 3s+2 = results staged; 3P = start
  while ((int)h->slab_ev.size() < 3 * P + 1) {
    cudaEvent_t e;
    CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    h->slab_ev.push_back(e);
  }
  std::vector<double *> trp(g.nt), trpo(g.nt);
  awfl_fields f;
  f.density_dry = h->stage_fields; f.uvel = h->stage_fields + g.ncell; f.vvel = h->stage_fields + 2 * g.ncell;
  f.wvel = h->stage_fields + 3 * g.ncell; f.temp = h->stage_fields + 4 * g.ncell;
  for (int t = 0; t < g.nt; t++) trp[t] = h->stage_fields + (size_t)(NS + t) * g.ncell;
  f.tracers = trp.data();
  FieldPtrs fp, fpo;
  if (make_field_ptrs(h, &f, fp)) return 1;
  for (int v = 0; v < NS; v++) fpo.f[v] = h->stage_out + (size_t)v * g.ncell;
  for (int t = 0; t < g.nt; t++) trpo[t] = h->stage_out + (size_t)(NS + t) * g.ncell;
  fpo.tr = h->trptr_out_d;
  if (g.nt > 0) CK(cudaMemcpyAsync(h->trptr_out_d, trpo.data(), sizeof(double *) * g.nt, cudaMemcpyHostToDevice, h->stream));
  const bool need_dt = ncycles_override <= 0;
  if (need_dt) {
    const double inf = INFINITY;
    CK(cudaMemcpyAsync(h->dt_min_d, &inf, sizeof(double), cudaMemcpyHostToDevice, h->stream));
  }
  // no stream may run ahead of earlier work on the handle's stream (staging buffers, tracer pointer tables, dt reset)
  const int ev_start = 3 * P;
  CK(cudaEventRecord(h->slab_ev[ev_start], h->stream));
  CK(cudaStreamWaitEvent(h->s_in, h->slab_ev[ev_start], 0));
  CK(cudaStreamWaitEvent(h->s_out, h->slab_ev[ev_start], 0));
  const bool alt = h->host_cstreams > 1 && P > 1;
  const int nc = h->host_cstreams > 4 ? 4 : h->host_cstreams;
  if (alt) for (int c = 0; c < nc; c++) CK(cudaStreamWaitEvent(h->s_cmp[c], h->slab_ev[ev_start], 0));
  // slabs on several streams would overwrite each other's pages of a paged vertical table: they read the global copy
  struct PageGuard {
    awfl_handle *h; bool saved;
    PageGuard(awfl_handle *h_, bool off) : h(h_), saved(h_->cvpage_ok) { if (off) h->cvpage_ok = false; }
    ~PageGuard() { h->cvpage_ok = saved; }
  } page_guard(h, alt);
  cudaStream_t const user_stream = h->stream;
  mark("start", 0, h->s_in);
  const int predicted = need_dt ? h->host_last_ncycles : ncycles_override;
  int qi = 0;   // buffer that holds the advanced state (the same for every slab: it depends on the sub-cycle count only)
  std::vector<Geom> slab(P, g);
  const size_t rows = (size_t)g.nz * g.ny * g.nx, pitch = sizeof(double) * (size_t)g.nens;
  auto slab_stream = [&](int s) { return alt ? h->s_cmp[s % nc] : user_stream; };
  // The two copy streams carry NOTHING but copies: the conversion kernels of a slab run on its compute stream. (They used to
  // sit between the copies; on a GPU filled by three slabs of marching kernels a conversion waits for free block slots, and
  // every such wait was a bubble on the host link -- the end-to-end step was bound by the device-to-host chain at 36 GB/s.)
  // slab s: conversion in (+ CFL minimum) on its compute stream as soon as its inputs have arrived
  auto convert_in_slab = [&](int s) -> int {
    cudaStream_t cs = slab_stream(s);
    CK(cudaStreamWaitEvent(cs, h->slab_ev[3 * s], 0));
    step_convert_in(h, slab[s], fp, need_dt, cs);
    CK(cudaEventRecord(h->slab_ev[3 * s + 1], cs));
    return 0;
  };
  // slab s: its sub-cycles and the conversion into the RESULT staging on its compute stream, then the copies out
  auto advance_slab = [&](int s, int n, double dts) -> int {
    cudaStream_t cs = slab_stream(s);
    h->stream = cs;                                   // the launch helpers enqueue on h->stream
    mark("advance_start", s, cs);
    qi = step_advance(h, slab[s], n, dts);
    h->stream = user_stream;
    mark("advanced", s, cs);
    step_convert_out(h, slab[s], fpo, qi, cs);
    CK(cudaEventRecord(h->slab_ev[3 * s + 2], cs));
    CK(cudaStreamWaitEvent(h->s_out, h->slab_ev[3 * s + 2], 0));
    for (int v = 0; v < N; v++) {
      if (P == 1) CK(cudaMemcpyAsync(host_fields[v], h->stage_out + (size_t)v * g.ncell, fb, cudaMemcpyDeviceToHost, h->s_out));
      else CK(cudaMemcpy2DAsync(host_fields[v] + slab[s].e0, pitch, h->stage_out + (size_t)v * g.ncell + slab[s].e0, pitch,
                                sizeof(double) * (size_t)slab[s].ne, rows, cudaMemcpyDeviceToHost, h->s_out));
    }
    mark("d2h_done", s, h->s_out);
    return 0;
  };
  for (int s = 0; s < P; s++) {
    const int t0 = (int)((long)tiles * s / P), t1 = (int)((long)tiles * (s + 1) / P);
    slab[s].e0 = 32 * t0;
    slab[s].ne = (t1 == tiles ? g.nens : 32 * t1) - slab[s].e0;
    for (int v = 0; v < N; v++) {
      if (P == 1) CK(cudaMemcpyAsync(h->stage_fields + (size_t)v * g.ncell, host_fields[v], fb, cudaMemcpyHostToDevice, h->s_in));
      else CK(cudaMemcpy2DAsync(h->stage_fields + (size_t)v * g.ncell + slab[s].e0, pitch, host_fields[v] + slab[s].e0, pitch,
                                sizeof(double) * (size_t)slab[s].ne, rows, cudaMemcpyHostToDevice, h->s_in));
    }
    CK(cudaEventRecord(h->slab_ev[3 * s], h->s_in));
    mark("h2d_done", s, h->s_in);
    if (convert_in_slab(s)) return 1;
    if (predicted > 0 && advance_slab(s, predicted, crm_dt / predicted)) return 1;
  }
  int ncycles = ncycles_override;
  if (need_dt) {
    // the CFL minimum is complete when every slab has been converted
    for (int s = 0; s < P; s++) CK(cudaStreamWaitEvent(h->s_in, h->slab_ev[3 * s + 1], 0));
    if (step_ncycles(h, crm_dt, &ncycles, h->s_in)) return 1;
  }
  h->host_last_ncycles = ncycles;
  const double dt = crm_dt / ncycles;       // Dycore.h:145
  if (predicted > 0 && predicted != ncycles) {
    // misprediction: let the speculative pass drain, then redo the whole step from the staged inputs (still intact),
    // un-overlapped; its results overwrite the speculative ones in the host buffers
    h->host_mispredictions++;
    CK(cudaStreamSynchronize(h->s_out));
    if (alt) for (int c = 0; c < nc; c++) CK(cudaStreamSynchronize(h->s_cmp[c]));
    step_convert_in(h, g, fp, false);
    qi = step_advance(h, g, ncycles, dt);
    step_convert_out(h, g, fpo, qi);
    CK(cudaEventRecord(h->slab_ev[ev_start], h->stream));
    CK(cudaStreamWaitEvent(h->s_out, h->slab_ev[ev_start], 0));
    for (int v = 0; v < N; v++) CK(cudaMemcpyAsync(host_fields[v], h->stage_out + (size_t)v * g.ncell, fb, cudaMemcpyDeviceToHost, h->s_out));
  } else if (predicted <= 0) {
    for (int s = 0; s < P; s++)
      if (advance_slab(s, ncycles, dt)) return 1;      // the dt needed every slab's conversion: all of them are done by now
  }
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->s_out));
  if (alt) for (int c = 0; c < nc; c++) CK(cudaStreamSynchronize(h->s_cmp[c]));
  CK(cudaStreamSynchronize(h->s_in));
  CK(cudaStreamSynchronize(h->stream));
  if (trace) {
    for (size_t m = 1; m < marks.size(); m++) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, marks[0].second, marks[m].second);
      fprintf(stderr, "awfl_time_step_host %-18s %8.3f ms\n", marks[m].first.c_str(), ms);
    }
    for (auto &m : marks) cudaEventDestroy(m.second);
  }
  if (ncycles_out) *ncycles_out = ncycles;
  return 0;
}

// Debug aid (compute-sanitizer's initcheck is not available on every pool): every scratch buffer of the handle -- the two
// stage states, partial tendencies, mass and tracer fluxes, FCT multipliers, the pressure cache, the host path's staging --
// is filled with NaN bit patterns. A step that read any of them before writing it would return NaN.
int awfl_poison_workspace(awfl_handle *h) {
  ENTER(h);
  const Geom &g = h->g;
  const size_t fb = sizeof(double) * (size_t)g.ncell, lvl = sizeof(double) * (size_t)g.ny * g.nx * g.nens;
  const int N = NS + g.nt;
  for (int b = 0; b < 3; b++) CK(cudaMemsetAsync(h->Q[b], 0xff, fb * N, h->stream));
  CK(cudaMemsetAsync(h->tend, 0xff, fb * NS, h->stream));
  CK(cudaMemsetAsync(h->pcache, 0xff, fb, h->stream));
  for (int d = 0; d < 3; d++) {
    CK(cudaMemsetAsync(h->mflux[d], 0xff, d == 2 ? fb + lvl : fb, h->stream));
    CK(cudaMemsetAsync(h->tflux[d], 0xff, (d == 2 ? fb + lvl : fb) * g.nt, h->stream));
  }
  CK(cudaMemsetAsync(h->mult, 0xff, fb * g.nt, h->stream));
  CK(cudaMemsetAsync(h->pint, 0xff, fb + lvl, h->stream));
  if (h->stage_fields) CK(cudaMemsetAsync(h->stage_fields, 0xff, fb * N, h->stream));
  if (h->stage_out) CK(cudaMemsetAsync(h->stage_out, 0xff, fb * N, h->stream));
  if (h->halo_pressure) CK(cudaMemsetAsync(h->halo_pressure, 0xff, sizeof(double) * (size_t)halo_dims(g, h->hs).nh, h->stream));
  return 0;
}

long awfl_host_mispredictions(const awfl_handle *h) { return h->host_mispredictions; }

int awfl_compute_tendencies(awfl_handle *h, const double *state, const double *tracers, const double *fct_start, double dt,
                            double *state_tend, double *tracers_tend) {
  ENTER(h);
  const Geom &g = h->g;
  const size_t fb = sizeof(double) * (size_t)g.ncell;
  // gather (state | tracers) into one N-field buffer
  CK(cudaMemcpyAsync(h->Q[1], state, fb * NS, cudaMemcpyDeviceToDevice, h->stream));
  if (g.nt > 0) CK(cudaMemcpyAsync(h->Q[1] + (size_t)NS * g.ncell, tracers, fb * g.nt, cudaMemcpyDeviceToDevice, h->stream));
  run_stage(h, h->g, h->Q[1], h->Q[1], nullptr, 0.0, 1.0, dt, dt, 1.0, 0.0, 2, fct_start, state_tend, tracers_tend);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

int awfl_declare_hydrostatic(awfl_handle *h, const awfl_fields *f, const awfl_gcm_columns *gcm) {
  ENTER(h);
  const Geom &g = h->g;
  HydroArgs a;
  memset(&a, 0, sizeof(a));
  a.dz = h->dz; a.vtab = h->vtab; a.pint = h->pint; a.grav_var = h->grav_var;
  if (gcm) {
    a.q = nullptr;
    a.g_rd = gcm->gcm_density_dry; a.g_t = gcm->gcm_temp; a.g_rv = gcm->gcm_water_vapor; a.g_rc = gcm->gcm_cloud_water; a.g_ri = gcm->gcm_cloud_ice;
  } else {
    FieldPtrs fp;
    if (make_field_ptrs(h, f, fp)) return 1;
    const RowMap rm = row_map(g, 256);
    k_coupler_to_dycore<<<row_blocks(g, rm), 256, 0, h->stream>>>(g, h->ph, fp, h->cfg.idWV, h->adds_mass_d, h->positive_d, h->Q[1], 0.8, h->dz, nullptr, rm);
    h->launches++;
    a.q = h->Q[1];
  }
  AWFL_ORD_DISPATCH(hydro_pint(h->stream, h->g, h->ph, a, h->shared_grid));
  h->launches++;
  const long n = (long)g.nz * g.nens;
  k_hydro_mean<<<(unsigned)((n + 127) / 128), 128, 0, h->stream>>>(g, a);
  h->launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

// Idealized initial state into the coupler fields (dyn_st == NULL) or into haloed dynamics arrays.
static int init_idealized(awfl_handle *h, const awfl_fields *f, const double *zint, int data_spec, double *dyn_st, double *dyn_tr) {
  const Geom &g = h->g;
  FieldPtrs fp;
  memset(&fp, 0, sizeof(fp));
  if (f && make_field_ptrs(h, f, fp)) return 1;
  // init-time scratch (not a time-step path): pressure / vapour at the 9 GLL points of every cell + two columns
  const size_t n2 = (size_t)g.nz * g.nens;
  double *scratch = nullptr;
  CK(cudaMalloc((void **)&scratch, sizeof(double) * (18 + 2) * n2));
  InitArgs a;
  a.zint = zint; a.pg = scratch; a.vapg = scratch + 9 * n2; a.hy_dens = scratch + 18 * n2; a.hy_pres = scratch + 19 * n2;
  a.xlen = h->cfg.xlen; a.ylen = h->cfg.ylen; a.cp_d = h->cfg.cp_d; a.p0 = h->cfg.p0;
  if (data_spec == 1) k_init_supercell_column<<<(unsigned)(g.nens + 63) / 64, 64, 0, h->stream>>>(g, h->ph, a);
  else k_init_thermal_column<<<dim3((unsigned)(g.nens + 127) / 128, (unsigned)g.nz), 128, 0, h->stream>>>(g, h->ph, a);
  const RowMap rm = row_map(g, 128);
  k_init_cells<<<row_blocks(g, rm), 128, 0, h->stream>>>(g, h->ph, a, fp, h->cfg.idWV, h->adds_mass_d, data_spec, rm, dyn_st, dyn_tr,
                                                         halo_dims(g, h->hs));
  h->launches += 2;
  cudaError_t e1 = cudaGetLastError(), e2 = cudaStreamSynchronize(h->stream);
  cudaFree(scratch);
  if (e1 != cudaSuccess || e2 != cudaSuccess) return fail(h, std::string("awfl_init_idealized: ") + cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
  return 0;
}

int awfl_init_idealized(awfl_handle *h, const awfl_fields *f, const double *zint, int data_spec) {
  ENTER(h);
  if (data_spec != 0 && data_spec != 1) return fail(h, "awfl_init_idealized: data_spec must be 0 (thermal) or 1 (supercell)");
  if (!zint) return fail(h, "awfl_init_idealized: vertical_interface_height is required");
  return init_idealized(h, f, zint, data_spec, nullptr, nullptr);
}

int awfl_init_supercell(awfl_handle *h, double *state, double *tracers, const double *zint) {
  ENTER(h);
  if (!state || !tracers || !zint) return fail(h, "awfl_init_supercell: NULL pointer");
  return init_idealized(h, nullptr, zint, 1, state, tracers);
}

// ---- the reference's haloed dynamics arrays at the boundary (awfl_haloed.cuh) ----
int awfl_convert_coupler_to_dynamics(awfl_handle *h, const awfl_fields *f, double *state, double *tracers) {
  ENTER(h);
  FieldPtrs fp;
  if (make_field_ptrs(h, f, fp)) return 1;
  const RowMap rm = row_map(h->g, 256);
  k_h_convert_in<<<row_blocks(h->g, rm), 256, 0, h->stream>>>(h->g, h->ph, fp, h->cfg.idWV, h->adds_mass_d, rm, halo_dims(h->g, h->hs), state, tracers);
  h->launches++;
  CK(cudaGetLastError());
  return 0;
}

int awfl_convert_dynamics_to_coupler(awfl_handle *h, const double *state, const double *tracers, const awfl_fields *f) {
  ENTER(h);
  FieldPtrs fp;
  if (make_field_ptrs(h, f, fp)) return 1;
  const RowMap rm = row_map(h->g, 256);
  k_h_convert_out<<<row_blocks(h->g, rm), 256, 0, h->stream>>>(h->g, h->ph, fp, h->cfg.idWV, h->adds_mass_d, rm, halo_dims(h->g, h->hs), state, tracers);
  h->launches++;
  CK(cudaGetLastError());
  return 0;
}

static void launch_halo_exchange(awfl_handle *h, double *state, double *tracers, double *pressure) {
  const Geom &g = h->g;
  const HaloDims d = halo_dims(g, h->hs);
  const unsigned nv = (unsigned)(NS + g.nt + 1);
  const long nx_ = (long)g.nz * g.ny * d.hs * g.nens, ny_ = (long)g.nz * d.hs * g.nx * g.nens, nw = (long)g.ny * g.nx * g.nens;
  k_h_periodic<<<dim3((unsigned)((nx_ + 255) / 256), nv), 256, 0, h->stream>>>(g, d, 0, state, tracers, pressure);
  if (!g.sim2d) k_h_periodic<<<dim3((unsigned)((ny_ + 255) / 256), nv), 256, 0, h->stream>>>(g, d, 1, state, tracers, pressure);
  k_h_wall<<<(unsigned)((nw + 127) / 128), 128, 0, h->stream>>>(g, h->ph, d, h->dz, state, tracers, pressure);
  h->launches += g.sim2d ? 2 : 3;
}

int awfl_halo_exchange(awfl_handle *h, double *state, double *tracers, double *pressure) {
  ENTER(h);
  if (!state || !tracers || !pressure) return fail(h, "awfl_halo_exchange: NULL pointer");
  launch_halo_exchange(h, state, tracers, pressure);
  CK(cudaGetLastError());
  return 0;
}

int awfl_compute_tendencies_haloed(awfl_handle *h, double *state, double *state_tend, double *tracers, double *tracers_tend, double dt) {
  ENTER(h);
  const Geom &g = h->g;
  const HaloDims d = halo_dims(g, h->hs);
  const RowMap rm = row_map(g, 256);
  const unsigned blocks = row_blocks(g, rm);
  // the right-hand side itself: the marching kernels on the gathered interior; the FCT start mass is what the caller left in
  // tracers_tend (Dycore.h:158,173,197 -> :536)
  k_h_gather<<<blocks, 256, 0, h->stream>>>(g, rm, d, state, tracers, h->Q[1]);
  h->launches++;
  run_stage(h, g, h->Q[1], h->Q[1], nullptr, 0.0, 1.0, dt, dt, 1.0, 0.0, 2, tracers_tend, state_tend, tracers_tend);
  // what the reference leaves in the caller's arrays (the SSPRK3 update of timeStep consumes it, Dycore.h:164,184,208):
  // interior divided by density and multiplied back, ghosts filled from the divided values
  if (!h->halo_pressure) CK(cudaMalloc((void **)&h->halo_pressure, sizeof(double) * (size_t)d.nh));
  k_h_divide_multiply<<<blocks, 256, 0, h->stream>>>(g, h->ph, rm, d, state, tracers, h->halo_pressure, 0);
  launch_halo_exchange(h, state, tracers, h->halo_pressure);
  k_h_divide_multiply<<<blocks, 256, 0, h->stream>>>(g, h->ph, rm, d, state, tracers, h->halo_pressure, 1);
  h->launches += 2;
  CK(cudaGetLastError());
  return 0;
}

int awfl_compute_mass_dynamics(awfl_handle *h, const double *state, const double *tracers, double *out_device) {
  ENTER(h);
  const Geom &g = h->g;
  const RowMap rm = row_map(g, 256);
  k_h_gather<<<row_blocks(g, rm), 256, 0, h->stream>>>(g, rm, halo_dims(g, h->hs), state, tracers, h->Q[1]);
  const long n = (long)(g.nt + 2) * g.nens;
  k_mass<<<(unsigned)((n + 127) / 128), 128, 0, h->stream>>>(g, h->Q[1], h->dz, out_device);
  h->launches += 2;
  CK(cudaGetLastError());
  return 0;
}

int awfl_compute_mass(awfl_handle *h, const awfl_fields *f, double *out_host) {
  ENTER(h);
  const Geom &g = h->g;
  FieldPtrs fp;
  if (make_field_ptrs(h, f, fp)) return 1;
  const RowMap rm = row_map(g, 256);
  k_coupler_to_dycore<<<row_blocks(g, rm), 256, 0, h->stream>>>(g, h->ph, fp, h->cfg.idWV, h->adds_mass_d, h->positive_d, h->Q[1], 0.8, h->dz, nullptr, rm);
  h->launches++;
  const long n = (long)(g.nt + 2) * g.nens;
  k_mass<<<(unsigned)((n + 127) / 128), 128, 0, h->stream>>>(g, h->Q[1], h->dz, h->mass_d);
  h->launches++;
  CK(cudaMemcpyAsync(out_host, h->mass_d, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

int awfl_reconstruct(awfl_handle *h, const double *stencils, int n, int ind, int level, int iens, double *out) {
  ENTER(h);
  if (level >= h->g.nz + 2 || iens < 0 || iens >= h->g.nens) return fail(h, "awfl_reconstruct: level/iens out of range");
  AWFL_ORD_DISPATCH(reconstruct(h->stream, stencils, n, ind, level, iens, h->g.nens, h->vtab, h->shared_grid, out));
  h->launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

int awfl_profile_begin(awfl_handle *h) {
  h->prof = true; h->ev_used = 0; h->ev_class.clear();
  return 0;
}

int awfl_profile_end(awfl_handle *h, double *ms_by_class, long *launches_by_class) {
  ENTER(h);
  h->prof = false;
  CK(cudaStreamSynchronize(h->stream));
  for (int c = 0; c < KC_COUNT; c++) { ms_by_class[c] = 0.0; launches_by_class[c] = 0; }
  for (size_t i = 0; i < h->ev_class.size(); i++) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, h->ev_pool[2 * i], h->ev_pool[2 * i + 1]));
    ms_by_class[h->ev_class[i]] += ms;
    launches_by_class[h->ev_class[i]]++;
  }
  h->ev_used = 0; h->ev_class.clear();
  return 0;
}

int awfl_measure_fp64_peak(awfl_handle *h, int iters, double *tflops) {
  ENTER(h);
  int dev = 0, sms = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int blocks = sms * 8, threads = 256;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  k_fp64_peak<<<blocks, threads, 0, h->stream>>>(iters / 8 + 1, 1.0, h->dt_min_d);   // warm-up
  CK(cudaEventRecord(e0, h->stream));
  k_fp64_peak<<<blocks, threads, 0, h->stream>>>(iters, 1.0, h->dt_min_d);
  CK(cudaEventRecord(e1, h->stream));
  CK(cudaEventSynchronize(e1));
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  h->launches += 2;
  const double fl = 2.0 * 64.0 * (double)iters * (double)blocks * threads;
  *tflops = fl / (ms * 1e-3) / 1e12;
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// An entry point that allocates nothing (no counterpart in the reference; added for
// callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// awfl_compute_mass_dynamics gathers the interior of the haloed arrays into the handle's workspace and sums it there; this
// variant needs no handle, and its workspace is the file-scope device array below. Every GPU that loads the library holds
// the array (AWFL_NOALLOC_MAX_DOUBLES * 8 bytes = 16 MiB) until the process exits.
//   * In host code `g_noalloc_scratch` is a host-side stand-in, NOT a device address: the kernels are given the address
//     cudaGetSymbolAddress returns, fetched once per device with that device current (noalloc_scratch below).
//   * Every call on a device shares the array, which keeps the previous call's contents. Nothing has to start at zero:
//     the gather writes every element the sum reads before it reads it (same stream). Calls that overlap on the GPU
//     (different streams, no dependency between them) race on the array.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_noalloc_scratch[AWFL_NOALLOC_MAX_DOUBLES];
namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
double *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};   // device address of g_noalloc_scratch, per device
// device address of the array on the CURRENT device (set-up on the first call there), or NULL with g_create_err set
double *noalloc_scratch() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { g_create_err = std::string("no-alloc entry point: cudaGetDevice: ") + cudaGetErrorString(e); return nullptr; }
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) { g_create_err = "no-alloc entry point: device ordinal beyond 63"; return nullptr; }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    double *p = nullptr;
    e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      g_create_err = std::string("no-alloc entry point: cudaGetSymbolAddress: ") + cudaGetErrorString(e);
      return nullptr;
    }
    g_noalloc_ptr[dev] = p;
  }
  return g_noalloc_ptr[dev];
}
}  // namespace

extern "C" {

unsigned long awfl_noalloc_max_doubles(void) { return (unsigned long)AWFL_NOALLOC_MAX_DOUBLES; }

int awfl_noalloc_setup(void) { return noalloc_scratch() ? 0 : 1; }

int awfl_compute_mass_dynamics_noalloc(int nz, int ny, int nx, int nens, int num_tracers, int ord, const double *state,
                                       const double *tracers, const double *dz, double *out_device, void *stream) {
  if (nz < 1 || ny < 1 || nx < 1 || nens < 1 || num_tracers < 0 || ord < 1 || !(ord & 1)) {
    g_create_err = "awfl_compute_mass_dynamics_noalloc: bad dimensions or order";
    return 1;
  }
  if (!state || !dz || !out_device || (num_tracers > 0 && !tracers)) {
    g_create_err = "awfl_compute_mass_dynamics_noalloc: null array";
    return 1;
  }
  const double need = (double)nz * ny * nx * nens * (double)(NS + num_tracers);
  if (need > (double)AWFL_NOALLOC_MAX_DOUBLES) {
    char msg[256];
    snprintf(msg, sizeof msg, "awfl_compute_mass_dynamics_noalloc: (5 + num_tracers) * nz * ny * nx * nens = %.0f doubles exceeds the "
             "static workspace of %lu doubles (use awfl_compute_mass_dynamics, whose workspace belongs to a handle)", need,
             (unsigned long)AWFL_NOALLOC_MAX_DOUBLES);
    g_create_err = msg;
    return 1;
  }
  double *scratch = noalloc_scratch();
  if (!scratch) return 1;
  Geom g;
  g.nz = nz; g.ny = ny; g.nx = nx; g.nens = nens; g.nt = num_tracers;
  g.sim2d = ny == 1;
  g.ncell = (long)nz * ny * nx * nens;
  g.dx = g.dy = 0.0;                    // not used by the two kernels below
  g.e0 = 0; g.ne = nens;
  cudaStream_t st = (cudaStream_t)stream;
  const RowMap rm = row_map(g, 256);
  k_h_gather<<<row_blocks(g, rm), 256, 0, st>>>(g, rm, halo_dims(g, (ord + 1) / 2), state, tracers, scratch);
  const long n = (long)(num_tracers + 2) * nens;
  k_mass<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(g, scratch, dz, out_device);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { g_create_err = std::string("awfl_compute_mass_dynamics_noalloc: launch: ") + cudaGetErrorString(e); return 1; }
  return 0;
}

}  // extern "C"
