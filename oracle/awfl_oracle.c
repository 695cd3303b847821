/* TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path (pam_b200/).
 *
 * Plain-C CPU restatement of the AWFL dycore time step of E3SM-Project/PAM @ 58b3e77, written from the
 * algorithm (not copied): SSPRK3 + WENO finite-volume compressible Euler on an A-grid with characteristic
 * upwind fluxes and FCT tracer positivity. Every function cites the reference lines it follows
 * (paths relative to /root/reference). Built once per reconstruction order (-DAWFL_ORD=3|5|7|9) by
 * oracle/Makefile into oracle/_lib/libawfl_oracle_ord<N>.so with -O2 -ffp-contract=off, the same numeric
 * flags as the compiled reference in oracle/_ref, so that the two can be compared bit-for-bit.
 *
 * PARITY PINNING: the reference has no golden vectors for this path, so this oracle is
 * pinned against the reference ITSELF, compiled unmodified in oracle/_ref (oracle/ref_build): tests/
 * test_oracle_golden.py requires bit-identical tables, reconstructions, tendencies and time steps, and
 * tests/golden/ holds outputs of the compiled reference for the GPU box.
 *
 * Semantics: the reference's bottom boundary condition has a read/write ordering hazard (Dycore.h:670 vs
 * :692). This oracle implements the INTENDED semantics (ghost theta = theta of the first
 * interior cell), identical to the reference driven with that ghost value pre-filled (ref_harness.cpp).
 *
 * Layout: all 4-D fields are (nz,ny,nx,nens), C order, iens fastest (pam_core/pam_const.h:30-37).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "awfl_tables.h"

#ifndef AWFL_ORD
#define AWFL_ORD 5
#endif
#define ORD AWFL_ORD
#define HS ((ORD + 1) / 2) /* halo width AND number of low-order candidate polynomials (Dycore.h:23) */

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define S2C CAT(AWFL_S2C_, ORD)
#define WRL CAT(AWFL_WRL_, ORD)
#define C2G CAT(AWFL_C2G_, ORD)
#define IDL_RAW CAT(AWFL_IDL_RAW_, ORD)
#define SIGMA CAT(AWFL_SIGMA_, ORD)
#define TVLO_N CAT(AWFL_TVLO_N_, ORD)
#define TVLO_I CAT(AWFL_TVLO_I_, ORD)
#define TVLO_J CAT(AWFL_TVLO_J_, ORD)
#define TVLO_C CAT(AWFL_TVLO_C_, ORD)
#define TVHI_N CAT(AWFL_TVHI_N_, ORD)
#define TVHI_I CAT(AWFL_TVHI_I_, ORD)
#define TVHI_J CAT(AWFL_TVHI_J_, ORD)
#define TVHI_C CAT(AWFL_TVHI_C_, ORD)

enum { idR = 0, idU = 1, idV = 2, idW = 3, idT = 4, NS = 5 };

typedef struct {
  int nz, ny, nx, nens, nt;
  double xlen, ylen, dx, dy;
  double R_d, cp_d, R_v, cp_v, p0, grav, cv_d, gamma, kappa, C0;
  int idWV;
  int *positive, *adds_mass;
  double *dz;       /* (nz,nens)                                  pam_coupler.h:184-202 */
  double *vs2c;     /* (nz+2,ORD,ORD,nens)  vert_sten_to_coefs     Dycore.h:898 */
  double *vwrl;     /* (nz+2,HS,HS,HS,nens) vert_weno_recon_lower  Dycore.h:897 */
  double *grav_var; /* (nz,nens)            variable_gravity       Dycore.h:868 */
  double idl[HS + 1], sigma;
} orc_t;

/* ------------------------------------------------------------------------------------------------ */
/* WENO limiter pieces                                                                              */
/* ------------------------------------------------------------------------------------------------ */

/* WenoLimiter.h:23-29 */
static void convexify(double *w) {
  double sum = 0.;
  const double eps = 1.0e-20;
  for (int i = 0; i < HS + 1; i++) sum += w[i];
  for (int i = 0; i < HS + 1; i++) w[i] /= (sum + eps);
}

/* WenoLimiter.h:11-19 */
static void map_weights(const double *d, double *w) {
  for (int i = 0; i < HS + 1; i++)
    w[i] = w[i] * (d[i] + d[i] * d[i] - 3. * d[i] * w[i] + w[i] * w[i]) / (d[i] * d[i] + w[i] * (1. - 2. * d[i]));
}

/* WenoLimiter.h:32-95 : tuned ideal weights, normalised by convexify */
static void ideal_sigma(double *idl, double *sigma) {
  for (int i = 0; i < HS + 1; i++) idl[i] = IDL_RAW[i];
  *sigma = SIGMA;
  convexify(idl);
}

/* TransformMatrices.h:188,871,2263,4734 (coefs_to_tv): truncated sum of squared-derivative integrals.
 * Term order and association follow the generated expression: c*(a_i*a_i) and (c*a_i)*a_j. */
static double tv_lo(const double *a) {
  double r = 0.;
  for (int t = 0; t < TVLO_N; t++) {
    int i = TVLO_I[t], j = TVLO_J[t];
    double term = (i == j) ? TVLO_C[t] * (a[i] * a[i]) : TVLO_C[t] * a[i] * a[j];
    r = (t == 0) ? term : r + term;
  }
  return r;
}
static double tv_hi(const double *a) {
  double r = 0.;
  for (int t = 0; t < TVHI_N; t++) {
    int i = TVHI_I[t], j = TVHI_J[t];
    double term = (i == j) ? TVHI_C[t] * (a[i] * a[i]) : TVHI_C[t] * a[i] * a[j];
    r = (t == 0) ? term : r + term;
  }
  return r;
}

/* Dycore.h:591-604 (reconstruct) + WenoLimiter.h:98-181 (compute_weno_coefs).
 * u[ORD] stencil of cell averages; hi[s][ii] high-order stencil->coefs; lo[i][s][ii] low-order;
 * ind = 0: value at the left edge of the centre cell, 1: right edge. */
static double reconstruct(const double *u, const double hi[ORD][ORD], const double lo[HS][HS][HS],
                          const double *idl, double sigma, int ind) {
  double a_lo[HS][HS], a_hi[ORD], tv[HS + 1], wts[HS + 1], aw[ORD];
  const double eps = 1.0e-20;
  for (int i = 0; i < HS; i++)
    for (int ii = 0; ii < HS; ii++) {
      double tmp = 0;
      for (int s = 0; s < HS; s++) tmp += lo[i][s][ii] * u[i + s];
      a_lo[i][ii] = tmp;
    }
  for (int ii = 0; ii < ORD; ii++) {
    double tmp = 0;
    for (int s = 0; s < ORD; s++) tmp += hi[s][ii] * u[s];
    a_hi[ii] = tmp;
  }
  /* bridge polynomial */
  for (int i = 0; i < HS; i++)
    for (int ii = 0; ii < HS; ii++) a_hi[ii] -= idl[i] * a_lo[i][ii];
  for (int ii = 0; ii < ORD; ii++) a_hi[ii] /= idl[HS];
  for (int i = 0; i < HS; i++) tv[i] = tv_lo(a_lo[i]);
  tv[HS] = tv_hi(a_hi);
  double lo_avg = 0.;
  for (int i = 0; i < HS; i++) lo_avg += tv[i];
  lo_avg /= HS;
  tv[HS] = lo_avg + (tv[HS] - lo_avg) * sigma;
  for (int i = 0; i < HS + 1; i++) wts[i] = idl[i] / (tv[i] * tv[i] + eps);
  convexify(wts);
  map_weights(idl, wts);
  convexify(wts);
  for (int i = 0; i < ORD; i++) aw[i] = wts[HS] * a_hi[i];
  for (int i = 0; i < HS; i++)
    for (int ii = 0; ii < HS; ii++) aw[ii] += wts[i] * a_lo[i][ii];
  double tmp = 0;
  for (int s = 0; s < ORD; s++) tmp += C2G[s][ind] * aw[s];
  return tmp;
}

/* ------------------------------------------------------------------------------------------------ */
/* Vertical (variable-dz) matrices                                                                  */
/* ------------------------------------------------------------------------------------------------ */

/* externals/YAKL/src/extensions/intrinsics/YAKL_intrinsics_matinv.h:9-55 : un-pivoted Gauss-Jordan */
static void matinv(int n, const double *a, double *inv) {
  double sc[ORD * ORD];
  for (int r = 0; r < n; r++)
    for (int c = 0; c < n; c++) {
      sc[r * n + c] = a[r * n + c];
      inv[r * n + c] = (r == c) ? 1. : 0.;
    }
  for (int d = 0; d < n; d++) {
    double f = 1. / sc[d * n + d];
    for (int c = d; c < n; c++) sc[d * n + c] *= f;
    for (int c = 0; c < n; c++) inv[d * n + c] *= f;
    for (int r = d + 1; r < n; r++) {
      double g = sc[r * n + d];
      for (int c = d; c < n; c++) sc[r * n + c] -= g * sc[d * n + c];
      for (int c = 0; c < n; c++) inv[r * n + c] -= g * inv[d * n + c];
    }
  }
  for (int d = n - 1; d >= 1; d--)
    for (int r = 0; r < d; r++) {
      double g = sc[r * n + d];
      for (int c = r + 1; c < n; c++) sc[r * n + c] -= g * sc[d * n + c];
      for (int c = 0; c < n; c++) inv[r * n + c] -= g * inv[d * n + c];
    }
}

/* TransformMatrices_variable.h:10-48 : cell-average Vandermonde of n cells with edges locs[0..n], inverted */
static void sten_to_coefs_var(int n, const double *locs, double *out) {
  double pw[ORD + 1], c2s[ORD * ORD];
  for (int i = 0; i < n + 1; i++) pw[i] = locs[i];
  for (int j = 0; j < n; j++) c2s[0 * n + j] = 1;
  for (int i = 1; i < n; i++) {
    for (int j = 0; j < n + 1; j++) pw[j] *= locs[j];
    for (int j = 0; j < n; j++) c2s[i * n + j] = 1. / (i + 1.) * (pw[j] - pw[j + 1]) / (locs[j] - locs[j + 1]);
  }
  matinv(n, c2s, out);
}

/* Dycore.h:897-939 : per-(level,iens) matrices; level k <-> cell k-1, window = ORD cells starting AT that
 * cell (clamped), normalised by the first cell's dz, then centred on the middle of the window. */
static void build_vertical_matrices(orc_t *o) {
  int nz = o->nz, nens = o->nens;
  for (int k = 0; k < nz + 2; k++)
    for (int e = 0; e < nens; e++) {
      double dzl[ORD], locs[ORD + 1], hi[ORD * ORD], lo[HS * HS], ll[HS + 1];
      for (int kk = 0; kk < ORD; kk++) {
        int i1 = -1 + k + kk, i2 = -1 + k;
        i1 = i1 < 0 ? 0 : (i1 > nz - 1 ? nz - 1 : i1);
        i2 = i2 < 0 ? 0 : (i2 > nz - 1 ? nz - 1 : i2);
        dzl[kk] = o->dz[i1 * nens + e] / o->dz[i2 * nens + e];
      }
      locs[0] = 0;
      for (int kk = 1; kk < ORD + 1; kk++) locs[kk] = locs[kk - 1] + dzl[kk - 1];
      double mid = (locs[(ORD - 1) / 2] + locs[(ORD + 1) / 2]) / 2;
      for (int kk = 0; kk < ORD + 1; kk++) locs[kk] = locs[kk] - mid;
      sten_to_coefs_var(ORD, locs, hi);
      for (int s = 0; s < ORD; s++)
        for (int ii = 0; ii < ORD; ii++) o->vs2c[((size_t)(k * ORD + s) * ORD + ii) * nens + e] = hi[s * ORD + ii];
      /* TransformMatrices_variable.h:51-69 */
      for (int i = 0; i < HS; i++) {
        for (int ii = 0; ii < HS + 1; ii++) ll[ii] = locs[i + ii];
        sten_to_coefs_var(HS, ll, lo);
        for (int s = 0; s < HS; s++)
          for (int ii = 0; ii < HS; ii++)
            o->vwrl[((size_t)((k * HS + i) * HS + s) * HS + ii) * nens + e] = lo[s * HS + ii];
      }
    }
}

static void load_vert(const orc_t *o, int level, int e, double hi[ORD][ORD], double lo[HS][HS][HS]) {
  int nens = o->nens;
  for (int s = 0; s < ORD; s++)
    for (int ii = 0; ii < ORD; ii++) hi[s][ii] = o->vs2c[((size_t)(level * ORD + s) * ORD + ii) * nens + e];
  for (int i = 0; i < HS; i++)
    for (int s = 0; s < HS; s++)
      for (int ii = 0; ii < HS; ii++) lo[i][s][ii] = o->vwrl[((size_t)((level * HS + i) * HS + s) * HS + ii) * nens + e];
}

/* ------------------------------------------------------------------------------------------------ */
/* Handle                                                                                           */
/* ------------------------------------------------------------------------------------------------ */

int orc_ord(void) { return ORD; }

/* zint: (nz+1) or (nz+1,nens). consts: NULL (Dycore.h:871-876 defaults) or {R_d,cp_d,R_v,cp_v,p0,grav}. */
void *orc_create(int nz, int ny, int nx, int nens, double xlen, double ylen, const double *zint, int zint_per_ens,
                 int nt, const int *positive, const int *adds_mass, int idWV, const double *consts) {
  orc_t *o = (orc_t *)calloc(1, sizeof(orc_t));
  o->nz = nz; o->ny = ny; o->nx = nx; o->nens = nens; o->nt = nt;
  o->xlen = xlen; o->ylen = ylen;
  o->dx = xlen / nx; o->dy = ylen / ny; /* pam_coupler.h get_dx/get_dy */
  o->R_d = consts ? consts[0] : 287.;   o->cp_d = consts ? consts[1] : 1003.;
  o->R_v = consts ? consts[2] : 461.;   o->cp_v = consts ? consts[3] : 1859;
  o->p0 = consts ? consts[4] : 1.e5;    o->grav = consts ? consts[5] : 9.81;
  o->cv_d = o->cp_d - o->R_d;                                    /* Dycore.h:883-891 */
  o->gamma = o->cp_d / o->cv_d;
  o->kappa = o->R_d / o->cp_d;
  o->C0 = pow(o->R_d * pow(o->p0, -o->kappa), o->gamma);
  o->idWV = idWV;
  o->positive = (int *)malloc(sizeof(int) * (nt > 0 ? nt : 1));
  o->adds_mass = (int *)malloc(sizeof(int) * (nt > 0 ? nt : 1));
  for (int t = 0; t < nt; t++) { o->positive[t] = positive[t]; o->adds_mass[t] = adds_mass[t]; }
  o->dz = (double *)malloc(sizeof(double) * nz * nens);
  for (int k = 0; k < nz; k++)
    for (int e = 0; e < nens; e++) {
      double lo = zint_per_ens ? zint[k * nens + e] : zint[k];
      double hi = zint_per_ens ? zint[(k + 1) * nens + e] : zint[k + 1];
      o->dz[k * nens + e] = hi - lo;                             /* pam_coupler.h:196-199 */
    }
  o->vs2c = (double *)malloc(sizeof(double) * (size_t)(nz + 2) * ORD * ORD * nens);
  o->vwrl = (double *)malloc(sizeof(double) * (size_t)(nz + 2) * HS * HS * HS * nens);
  o->grav_var = (double *)calloc((size_t)nz * nens, sizeof(double));
  build_vertical_matrices(o);
  ideal_sigma(o->idl, &o->sigma);
  return o;
}

void orc_destroy(void *h) {
  orc_t *o = (orc_t *)h;
  if (!o) return;
  free(o->positive); free(o->adds_mass); free(o->dz); free(o->vs2c); free(o->vwrl); free(o->grav_var);
  free(o);
}

double orc_get_const(void *h, int which) {
  orc_t *o = (orc_t *)h;
  double v[] = {o->R_d, o->cp_d, o->R_v, o->cp_v, o->p0, o->grav, o->cv_d, o->gamma, o->kappa, o->C0};
  return v[which];
}
void orc_get_tables(double *s2c, double *c2g, double *wrl, double *idl, double *sigma) {
  double d[HS + 1], sg;
  ideal_sigma(d, &sg);
  memcpy(s2c, S2C, sizeof(double) * ORD * ORD);
  memcpy(c2g, C2G, sizeof(double) * ORD * 2);
  memcpy(wrl, WRL, sizeof(double) * HS * HS * HS);
  memcpy(idl, d, sizeof(double) * (HS + 1));
  *sigma = sg;
}
void orc_get_vert_tables(void *h, double *vs2c, double *vwrl) {
  orc_t *o = (orc_t *)h;
  memcpy(vs2c, o->vs2c, sizeof(double) * (size_t)(o->nz + 2) * ORD * ORD * o->nens);
  memcpy(vwrl, o->vwrl, sizeof(double) * (size_t)(o->nz + 2) * HS * HS * HS * o->nens);
}
void orc_get_dz(void *h, double *dz) { orc_t *o = (orc_t *)h; memcpy(dz, o->dz, sizeof(double) * o->nz * o->nens); }
void orc_get_grav_var(void *h, double *g) { orc_t *o = (orc_t *)h; memcpy(g, o->grav_var, sizeof(double) * o->nz * o->nens); }
void orc_set_grav_var(void *h, const double *g) { orc_t *o = (orc_t *)h; memcpy(o->grav_var, g, sizeof(double) * o->nz * o->nens); }

double orc_coefs_to_tv(int n, const double *a) { return n == HS ? tv_lo(a) : (n == ORD ? tv_hi(a) : -1.); }

void orc_reconstruct(const double *stencils, int n, int ind, double *out) {
  double idl[HS + 1], sg;
  ideal_sigma(idl, &sg);
  for (int q = 0; q < n; q++) out[q] = reconstruct(stencils + (size_t)q * ORD, S2C, WRL, idl, sg, ind);
}
void orc_reconstruct_vert(void *h, int level, int e, const double *stencils, int n, int ind, double *out) {
  orc_t *o = (orc_t *)h;
  double hi[ORD][ORD], lo[HS][HS][HS];
  load_vert(o, level, e, hi, lo);
  for (int q = 0; q < n; q++) out[q] = reconstruct(stencils + (size_t)q * ORD, hi, lo, o->idl, o->sigma, ind);
}

/* ------------------------------------------------------------------------------------------------ */
/* Haloed work arrays: [var][nz+2HS][ny+2HS][nx+2HS][nens]                                           */
/* ------------------------------------------------------------------------------------------------ */
#define NZH (o->nz + 2 * HS)
#define NYH (o->ny + 2 * HS)
#define NXH (o->nx + 2 * HS)
#define H4(k, j, i, e) ((((size_t)(k) * NYH + (j)) * NXH + (i)) * o->nens + (e))
#define H5(l, k, j, i, e) ((size_t)(l) * NZH * NYH * NXH * o->nens + H4(k, j, i, e))
#define C4(k, j, i, e) ((((size_t)(k) * o->ny + (j)) * o->nx + (i)) * o->nens + (e))
#define C5(l, k, j, i, e) ((size_t)(l) * o->nz * o->ny * o->nx * o->nens + C4(k, j, i, e))

static size_t halo_count(const orc_t *o) { return (size_t)NZH * NYH * NXH * o->nens; }
static size_t cell_count(const orc_t *o) { return (size_t)o->nz * o->ny * o->nx * o->nens; }

/* Dycore.h:608-711 (halo_exchange): periodic x / y ghosts, solid-wall z ghosts with hydrostatically
 * extrapolated density at constant theta. Intended semantics for the bottom ghost pressure:
 * all ghost copies are made first, then ghost density and pressure. */
static void halo_exchange(const orc_t *o, double *st, double *tr, double *pr) {
  int nz = o->nz, ny = o->ny, nx = o->nx, nens = o->nens, nt = o->nt;
  double gamma = o->gamma, C0 = o->C0, grav = o->grav;
  for (int v = 0; v < NS + nt + 1; v++) {
    double *a = v < NS ? st + H5(v, 0, 0, 0, 0) : (v < NS + nt ? tr + H5(v - NS, 0, 0, 0, 0) : pr);
    for (int k = 0; k < nz; k++)
      for (int j = 0; j < ny; j++)
        for (int ii = 0; ii < HS; ii++)
          for (int e = 0; e < nens; e++) {
            a[H4(HS + k, HS + j, nx + HS + ii, e)] = a[H4(HS + k, HS + j, HS + ii, e)];
            a[H4(HS + k, HS + j, ii, e)] = a[H4(HS + k, HS + j, nx + ii, e)];
          }
    if (ny > 1)
      for (int k = 0; k < nz; k++)
        for (int jj = 0; jj < HS; jj++)
          for (int i = 0; i < nx; i++)
            for (int e = 0; e < nens; e++) {
              a[H4(HS + k, ny + HS + jj, HS + i, e)] = a[H4(HS + k, HS + jj, HS + i, e)];
              a[H4(HS + k, jj, HS + i, e)] = a[H4(HS + k, ny + jj, HS + i, e)];
            }
  }
  for (int kk = 0; kk < HS; kk++)
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          for (int l = 1; l < NS; l++) {
            st[H5(l, kk, HS + j, HS + i, e)] = (l == idW) ? 0. : st[H5(l, HS, HS + j, HS + i, e)];
            st[H5(l, HS + nz + kk, HS + j, HS + i, e)] = (l == idW) ? 0. : st[H5(l, HS + nz - 1, HS + j, HS + i, e)];
          }
          for (int l = 0; l < nt; l++) {
            tr[H5(l, kk, HS + j, HS + i, e)] = tr[H5(l, HS, HS + j, HS + i, e)];
            tr[H5(l, HS + nz + kk, HS + j, HS + i, e)] = tr[H5(l, HS + nz - 1, HS + j, HS + i, e)];
          }
        }
  for (int kk = 0; kk < HS; kk++)
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          { /* bottom: Dycore.h:682-695 */
            int k0 = HS, k = k0 - 1 - kk;
            double rho0 = st[H5(idR, k0, HS + j, HS + i, e)], theta0 = st[H5(idT, k0, HS + j, HS + i, e)];
            double rho0_gm1 = pow(rho0, gamma - 1), theta0_g = pow(theta0, gamma);
            st[H5(idR, k, HS + j, HS + i, e)] =
                pow(rho0_gm1 + grav * (gamma - 1) * o->dz[(k0 - HS) * nens + e] * (kk + 1) / (gamma * C0 * theta0_g),
                    1. / (gamma - 1));
            double rt = st[H5(idR, k, HS + j, HS + i, e)] * st[H5(idT, k, HS + j, HS + i, e)];
            pr[H4(k, HS + j, HS + i, e)] = C0 * pow(rt, gamma);
          }
          { /* top: Dycore.h:696-709 */
            int k0 = HS + nz - 1, k = k0 + 1 + kk;
            double rho0 = st[H5(idR, k0, HS + j, HS + i, e)], theta0 = st[H5(idT, k0, HS + j, HS + i, e)];
            double rho0_gm1 = pow(rho0, gamma - 1), theta0_g = pow(theta0, gamma);
            st[H5(idR, k, HS + j, HS + i, e)] =
                pow(rho0_gm1 - grav * (gamma - 1) * o->dz[(k0 - HS) * nens + e] * (kk + 1) / (gamma * C0 * theta0_g),
                    1. / (gamma - 1));
            double rt = st[H5(idR, k, HS + j, HS + i, e)] * st[H5(idT, k, HS + j, HS + i, e)];
            pr[H4(k, HS + j, HS + i, e)] = C0 * pow(rt, gamma);
          }
        }
}

/* One direction of the flux kernel, Dycore.h:339-386 (x), :391-438 (y), :452-518 (z).
 * base = haloed offset of stencil entry s=0 of the LEFT reconstruction; sd = stride along the direction.
 * mom = index of the momentum normal to the interface. wall = rw forced to 0 (z, k==0||k==nz).
 * Writes NS state fluxes and nt tracer fluxes. */
static void face_flux(const orc_t *o, const double *st, const double *tr, const double *pr, size_t base, size_t sd,
                      int mom, int wall, const double hiL[ORD][ORD], const double loL[HS][HS][HS],
                      const double hiR[ORD][ORD], const double loR[HS][HS][HS], double *fs, double *ft) {
  const double cs = 350;
  const size_t vs = halo_count(o);
  const double *idl = o->idl;
  double sigma = o->sigma, sten[ORD];
  for (int s = 0; s < ORD; s++) sten[s] = st[idR * vs + base + s * sd] * st[mom * vs + base + s * sd];
  double r_L = reconstruct(sten, hiL, loL, idl, sigma, 1);
  if (wall) r_L = 0;
  for (int s = 0; s < ORD; s++) sten[s] = st[idR * vs + base + (1 + s) * sd] * st[mom * vs + base + (1 + s) * sd];
  double r_R = reconstruct(sten, hiR, loR, idl, sigma, 0);
  if (wall) r_R = 0;
  for (int s = 0; s < ORD; s++) sten[s] = pr[base + s * sd];
  double p_L = reconstruct(sten, hiL, loL, idl, sigma, 1);
  for (int s = 0; s < ORD; s++) sten[s] = pr[base + (1 + s) * sd];
  double p_R = reconstruct(sten, hiR, loR, idl, sigma, 0);
  double w1 = 0.5 * (p_R - cs * r_R);
  double w2 = 0.5 * (p_L + cs * r_L);
  double pp = w1 + w2;
  double rn = (w2 - w1) / cs;
  if (wall) rn = 0;
  fs[idR] = rn;
  int upw = rn > 0 ? 0 : 1;
  const double(*hi)[ORD] = upw ? hiR : hiL;
  const double(*lo)[HS][HS] = upw ? loR : loL;
  for (int l = 1; l < NS; l++) {
    for (int s = 0; s < ORD; s++) sten[s] = st[l * vs + base + (upw + s) * sd];
    fs[l] = rn * reconstruct(sten, hi, lo, idl, sigma, 1 - upw);
    if (l == mom) fs[l] = fs[l] + pp;
  }
  for (int t = 0; t < o->nt; t++) {
    for (int s = 0; s < ORD; s++) sten[s] = tr[t * vs + base + (upw + s) * sd];
    ft[t] = rn * reconstruct(sten, hi, lo, idl, sigma, 1 - upw);
  }
}

static double dmax(double a, double b) { return a < b ? b : a; } /* std::max */
static double dmin(double a, double b) { return b < a ? b : a; } /* std::min */

/* Dycore.h:262-586 (compute_tendencies). st/tr haloed, interior filled with conserved variables; on return
 * the interior holds the round-tripped (x/rho)*rho values. tr_tend: in = FCT start mass, out = tendency. */
static void compute_tendencies(const orc_t *o, double *st, double *st_tend, double *tr, double *tr_tend, double dt) {
  int nz = o->nz, ny = o->ny, nx = o->nx, nens = o->nens, nt = o->nt;
  int sim2d = ny == 1;
  double dx = o->dx, dy = o->dy, C0 = o->C0, gamma = o->gamma;
  double *pr = (double *)calloc(halo_count(o), sizeof(double));
  /* Dycore.h:310-321 */
  for (int k = 0; k < nz; k++)
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          double rho = st[H5(idR, HS + k, HS + j, HS + i, e)];
          pr[H4(HS + k, HS + j, HS + i, e)] = C0 * pow(st[H5(idT, HS + k, HS + j, HS + i, e)], gamma);
          for (int l = 1; l < NS; l++) st[H5(l, HS + k, HS + j, HS + i, e)] /= rho;
          for (int t = 0; t < nt; t++) tr[H5(t, HS + k, HS + j, HS + i, e)] /= rho;
        }
  halo_exchange(o, st, tr, pr);
  size_t nfx = (size_t)nz * ny * (nx + 1) * nens, nfy = (size_t)nz * (ny + 1) * nx * nens, nfz = (size_t)(nz + 1) * ny * nx * nens;
  double *sfx = (double *)calloc(NS * nfx, sizeof(double)), *sfy = (double *)calloc(NS * nfy, sizeof(double));
  double *sfz = (double *)calloc(NS * nfz, sizeof(double));
  double *tfx = (double *)calloc((nt + 1) * nfx, sizeof(double)), *tfy = (double *)calloc((nt + 1) * nfy, sizeof(double));
  double *tfz = (double *)calloc((nt + 1) * nfz, sizeof(double));
#define FX(k, j, i, e) ((((size_t)(k) * ny + (j)) * (nx + 1) + (i)) * nens + (e))
#define FY(k, j, i, e) ((((size_t)(k) * (ny + 1) + (j)) * nx + (i)) * nens + (e))
#define FZ(k, j, i, e) ((((size_t)(k) * ny + (j)) * nx + (i)) * nens + (e))
  size_t sx = (size_t)nens, sy = (size_t)NXH * nens, sz = (size_t)NYH * NXH * nens;
  double fs[NS], ft[64];
  for (int k = 0; k < nz + 1; k++)
    for (int j = 0; j < ny + 1; j++)
      for (int i = 0; i < nx + 1; i++)
        for (int e = 0; e < nens; e++) {
          if (j < ny && k < nz) {
            face_flux(o, st, tr, pr, H4(HS + k, HS + j, i, e), sx, idU, 0, S2C, WRL, S2C, WRL, fs, ft);
            for (int l = 0; l < NS; l++) sfx[l * nfx + FX(k, j, i, e)] = fs[l];
            for (int t = 0; t < nt; t++) tfx[t * nfx + FX(k, j, i, e)] = ft[t];
          }
          if (i < nx && k < nz && !sim2d) {
            face_flux(o, st, tr, pr, H4(HS + k, j, HS + i, e), sy, idV, 0, S2C, WRL, S2C, WRL, fs, ft);
            for (int l = 0; l < NS; l++) sfy[l * nfy + FY(k, j, i, e)] = fs[l];
            for (int t = 0; t < nt; t++) tfy[t * nfy + FY(k, j, i, e)] = ft[t];
          } /* sim2d: y fluxes stay 0 (Dycore.h:439-446) */
          if (i < nx && j < ny) {
            double hiL[ORD][ORD], loL[HS][HS][HS], hiR[ORD][ORD], loR[HS][HS][HS];
            load_vert(o, k, e, hiL, loL);
            load_vert(o, k + 1, e, hiR, loR);
            face_flux(o, st, tr, pr, H4(k, HS + j, HS + i, e), sz, idW, (k == 0 || k == nz), hiL, loL, hiR, loR, fs, ft);
            for (int l = 0; l < NS; l++) sfz[l * nfz + FZ(k, j, i, e)] = fs[l];
            for (int t = 0; t < nt; t++) tfz[t * nfz + FZ(k, j, i, e)] = ft[t];
          }
        }
  /* Dycore.h:525-550 : multiply rho back, FCT positivity scaling of outgoing fluxes */
  for (int k = 0; k < nz; k++)
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          double rho = st[H5(idR, HS + k, HS + j, HS + i, e)], dzk = o->dz[k * nens + e];
          for (int l = 1; l < NS; l++) st[H5(l, HS + k, HS + j, HS + i, e)] *= rho;
          for (int t = 0; t < nt; t++) {
            tr[H5(t, HS + k, HS + j, HS + i, e)] *= rho;
            if (!o->positive[t]) continue;
            double *fx = tfx + t * nfx, *fy = tfy + t * nfy, *fz = tfz + t * nfz;
            double avail = dmax(tr_tend[C5(t, k, j, i, e)], 0.) * dx * dy * dzk;
            double ox = (dmax(fx[FX(k, j, i + 1, e)], 0.) - dmin(fx[FX(k, j, i, e)], 0.)) / dx;
            double oy = (dmax(fy[FY(k, j + 1, i, e)], 0.) - dmin(fy[FY(k, j, i, e)], 0.)) / dy;
            double oz = (dmax(fz[FZ(k + 1, j, i, e)], 0.) - dmin(fz[FZ(k, j, i, e)], 0.)) / dzk;
            double out = (ox + oy + oz) * dt * dx * dy * dzk;
            if (out > avail) {
              double mult = avail / out;
              if (fx[FX(k, j, i + 1, e)] > 0) fx[FX(k, j, i + 1, e)] *= mult;
              if (fx[FX(k, j, i, e)] < 0) fx[FX(k, j, i, e)] *= mult;
              if (fy[FY(k, j + 1, i, e)] > 0) fy[FY(k, j + 1, i, e)] *= mult;
              if (fy[FY(k, j, i, e)] < 0) fy[FY(k, j, i, e)] *= mult;
              if (fz[FZ(k + 1, j, i, e)] > 0) fz[FZ(k + 1, j, i, e)] *= mult;
              if (fz[FZ(k, j, i, e)] < 0) fz[FZ(k, j, i, e)] *= mult;
            }
          }
        }
  /* Dycore.h:553-584 : flux divergence + gravity; periodic seam takes the min of the two edge copies */
  for (int k = 0; k < nz; k++)
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          double dzk = o->dz[k * nens + e];
          for (int l = 0; l < NS; l++) {
            double t = -(sfx[l * nfx + FX(k, j, i + 1, e)] - sfx[l * nfx + FX(k, j, i, e)]) / dx -
                       (sfy[l * nfy + FY(k, j + 1, i, e)] - sfy[l * nfy + FY(k, j, i, e)]) / dy -
                       (sfz[l * nfz + FZ(k + 1, j, i, e)] - sfz[l * nfz + FZ(k, j, i, e)]) / dzk;
            if (l == idW) t += -o->grav_var[k * nens + e] * st[H5(idR, HS + k, HS + j, HS + i, e)];
            if (l == idV && sim2d) t = 0;
            st_tend[C5(l, k, j, i, e)] = t;
          }
          for (int t = 0; t < nt; t++) {
            double *fx = tfx + t * nfx, *fy = tfy + t * nfy, *fz = tfz + t * nfz;
            double f0 = fx[FX(k, j, i, e)], f1 = fx[FX(k, j, i + 1, e)];
            double g0 = fy[FY(k, j, i, e)], g1 = fy[FY(k, j + 1, i, e)];
            double h0 = fz[FZ(k, j, i, e)], h1 = fz[FZ(k + 1, j, i, e)];
            if (i == 0) f0 = dmin(f0, fx[FX(k, j, nx, e)]);
            if (i == nx - 1) f1 = dmin(f1, fx[FX(k, j, 0, e)]);
            if (j == 0) g0 = dmin(g0, fy[FY(k, ny, i, e)]);
            if (j == ny - 1) g1 = dmin(g1, fy[FY(k, 0, i, e)]);
            tr_tend[C5(t, k, j, i, e)] = -(f1 - f0) / dx - (g1 - g0) / dy - (h1 - h0) / dzk;
          }
        }
  free(pr); free(sfx); free(sfy); free(sfz); free(tfx); free(tfy); free(tfz);
}

/* un-haloed <-> haloed helpers */
static void to_halo(const orc_t *o, const double *src, double *dst, int nv) {
  for (int l = 0; l < nv; l++)
    for (int k = 0; k < o->nz; k++)
      for (int j = 0; j < o->ny; j++)
        for (int i = 0; i < o->nx; i++)
          memcpy(dst + H5(l, HS + k, HS + j, HS + i, 0), src + C5(l, k, j, i, 0), sizeof(double) * o->nens);
}
static void from_halo(const orc_t *o, const double *src, double *dst, int nv) {
  for (int l = 0; l < nv; l++)
    for (int k = 0; k < o->nz; k++)
      for (int j = 0; j < o->ny; j++)
        for (int i = 0; i < o->nx; i++)
          memcpy(dst + C5(l, k, j, i, 0), src + H5(l, HS + k, HS + j, HS + i, 0), sizeof(double) * o->nens);
}

/* Public: one semi-discrete right-hand side on un-haloed conserved arrays (same contract as
 * ref_compute_tendencies in oracle/ref_build/ref_harness.cpp). */
int orc_compute_tendencies(void *h, const double *state_in, const double *tracers_in, const double *fct_start, double dt,
                           double *state_tend, double *tracers_tend, double *state_rt, double *tracers_rt) {
  orc_t *o = (orc_t *)h;
  size_t nh = halo_count(o), nc = cell_count(o);
  double *st = (double *)calloc(NS * nh, sizeof(double)), *tr = (double *)calloc((o->nt + 1) * nh, sizeof(double));
  double *tt = (double *)calloc((o->nt + 1) * nc, sizeof(double));
  to_halo(o, state_in, st, NS);
  to_halo(o, tracers_in, tr, o->nt);
  memcpy(tt, fct_start, sizeof(double) * o->nt * nc);
  compute_tendencies(o, st, state_tend, tr, tt, dt);
  memcpy(tracers_tend, tt, sizeof(double) * o->nt * nc);
  if (state_rt) from_halo(o, st, state_rt, NS);
  if (tracers_rt) from_halo(o, tr, tracers_rt, o->nt);
  free(st); free(tr); free(tt);
  return 0;
}

/* Dycore.h:1336-1388 (convert_coupler_to_dynamics). f[0..4] = density_dry,uvel,vvel,wvel,temp; f[5+t] tracers */
static void coupler_to_dynamics(const orc_t *o, double *const *f, double *st, double *tr) {
  for (int k = 0; k < o->nz; k++)
    for (int j = 0; j < o->ny; j++)
      for (int i = 0; i < o->nx; i++)
        for (int e = 0; e < o->nens; e++) {
          size_t c = C4(k, j, i, e);
          double rho_d = f[0][c], u = f[1][c], v = f[2][c], w = f[3][c], temp = f[4][c];
          double rho_v = f[5 + o->idWV][c];
          double press = rho_d * o->R_d * temp + rho_v * o->R_v * temp;
          double rho = rho_d;
          for (int t = 0; t < o->nt; t++)
            if (o->adds_mass[t]) rho += f[5 + t][c];
          double theta = pow(press / o->C0, 1. / o->gamma) / rho;
          st[H5(idR, HS + k, HS + j, HS + i, e)] = rho;
          st[H5(idU, HS + k, HS + j, HS + i, e)] = rho * u;
          st[H5(idV, HS + k, HS + j, HS + i, e)] = rho * v;
          st[H5(idW, HS + k, HS + j, HS + i, e)] = rho * w;
          st[H5(idT, HS + k, HS + j, HS + i, e)] = rho * theta;
          for (int t = 0; t < o->nt; t++) tr[H5(t, HS + k, HS + j, HS + i, e)] = f[5 + t][c];
        }
}

/* Dycore.h:1281-1331 (convert_dynamics_to_coupler) */
static void dynamics_to_coupler(const orc_t *o, const double *st, const double *tr, double *const *f) {
  for (int k = 0; k < o->nz; k++)
    for (int j = 0; j < o->ny; j++)
      for (int i = 0; i < o->nx; i++)
        for (int e = 0; e < o->nens; e++) {
          size_t c = C4(k, j, i, e);
          double rho = st[H5(idR, HS + k, HS + j, HS + i, e)];
          double u = st[H5(idU, HS + k, HS + j, HS + i, e)] / rho;
          double v = st[H5(idV, HS + k, HS + j, HS + i, e)] / rho;
          double w = st[H5(idW, HS + k, HS + j, HS + i, e)] / rho;
          double theta = st[H5(idT, HS + k, HS + j, HS + i, e)] / rho;
          double press = o->C0 * pow(rho * theta, o->gamma);
          double rho_v = tr[H5(o->idWV, HS + k, HS + j, HS + i, e)];
          double rho_d = rho;
          for (int t = 0; t < o->nt; t++)
            if (o->adds_mass[t]) rho_d -= tr[H5(t, HS + k, HS + j, HS + i, e)];
          double temp = press / (rho_d * o->R_d + rho_v * o->R_v);
          f[0][c] = rho_d; f[1][c] = u; f[2][c] = v; f[3][c] = w; f[4][c] = temp;
          for (int t = 0; t < o->nt; t++) f[5 + t][c] = tr[H5(t, HS + k, HS + j, HS + i, e)];
        }
}

/* Dycore.h:65-102 (compute_time_step) */
double orc_compute_time_step(void *h, double *const *f, double cfl) {
  orc_t *o = (orc_t *)h;
  double best = INFINITY;
  for (int k = 0; k < o->nz; k++)
    for (int j = 0; j < o->ny; j++)
      for (int i = 0; i < o->nx; i++)
        for (int e = 0; e < o->nens; e++) {
          size_t c = C4(k, j, i, e);
          double rho_d = f[0][c], u = f[1][c], v = f[2][c], w = f[3][c], temp = f[4][c], rho_v = f[5 + o->idWV][c];
          double rho = rho_d + rho_v;
          double p = (rho_d * o->R_d + rho_v * o->R_v) * temp;
          double cs = sqrt(o->gamma * p / rho);
          double dtx = cfl * o->dx / (fabs(u) + cs);
          double dty = cfl * o->dy / (fabs(v) + cs);
          double dtz = cfl * o->dz[k * o->nens + e] / (fabs(w) + cs);
          double m = dmin(dmin(dtx, dty), dtz);
          if (m < best) best = m;
        }
  return best;
}

/* SSPRK3 stage update, Dycore.h:162-176 / :182-200 / :206-221:
 *   out = a*base + b*stage + (b*dt)*tend  written with the reference's association for each stage. */
static void rk_update(const orc_t *o, int stage, double dt, const double *q0, const double *t0, double *q1, double *t1,
                      double *q_out, double *t_out, const double *dq, double *dtr) {
  int nt = o->nt;
  for (int k = 0; k < o->nz; k++)
    for (int j = 0; j < o->ny; j++)
      for (int i = 0; i < o->nx; i++)
        for (int e = 0; e < o->nens; e++) {
          for (int l = 0; l < NS; l++) {
            size_t hh = H5(l, HS + k, HS + j, HS + i, e), cc = C5(l, k, j, i, e);
            if (stage == 1) q_out[hh] = q0[hh] + dt * dq[cc];
            if (stage == 2) q_out[hh] = (3. / 4.) * q0[hh] + (1. / 4.) * q1[hh] + (1. / 4.) * dt * dq[cc];
            if (stage == 3) q_out[hh] = (1. / 3.) * q0[hh] + (2. / 3.) * q1[hh] + (2. / 3.) * dt * dq[cc];
          }
          for (int l = 0; l < nt; l++) {
            size_t hh = H5(l, HS + k, HS + j, HS + i, e), cc = C5(l, k, j, i, e);
            double v = 0;
            if (stage == 1) v = t0[hh] + dt * dtr[cc];
            if (stage == 2) v = (3. / 4.) * t0[hh] + (1. / 4.) * t1[hh] + (1. / 4.) * dt * dtr[cc];
            if (stage == 3) v = (1. / 3.) * t0[hh] + (2. / 3.) * t1[hh] + (2. / 3.) * dt * dtr[cc];
            if (o->positive[l]) v = dmax(0., v);
            t_out[hh] = v;
            if (stage == 1) dtr[cc] = (3. / 4.) * t0[hh] + (1. / 4.) * v;
            if (stage == 2) dtr[cc] = (1. / 3.) * t0[hh] + (2. / 3.) * v;
          }
        }
}

/* Dycore.h:107-255 (timeStep), intended semantics. f = coupler fields (5+nt pointers), updated in place.
 * ncycles_override>0 fixes the sub-cycle count. Returns ncycles. */
int orc_time_step(void *h, double *const *f, double crm_dt, int ncycles_override) {
  orc_t *o = (orc_t *)h;
  int nt = o->nt;
  size_t nh = halo_count(o), nc = cell_count(o);
  double *q = (double *)calloc(NS * nh, sizeof(double)), *tr = (double *)calloc((nt + 1) * nh, sizeof(double));
  double *q1 = (double *)calloc(NS * nh, sizeof(double)), *t1 = (double *)calloc((nt + 1) * nh, sizeof(double));
  double *dq = (double *)calloc(NS * nc, sizeof(double)), *dtr = (double *)calloc((nt + 1) * nc, sizeof(double));
  coupler_to_dynamics(o, f, q, tr);
  for (int l = 0; l < nt; l++)
    if (o->positive[l])
      for (int k = 0; k < o->nz; k++)
        for (int j = 0; j < o->ny; j++)
          for (int i = 0; i < o->nx; i++)
            for (int e = 0; e < o->nens; e++) {
              size_t hh = H5(l, HS + k, HS + j, HS + i, e);
              tr[hh] = dmax(0., tr[hh]);
            }
  double dt_dyn = orc_compute_time_step(h, f, 0.8);
  int ncycles = (int)ceil(crm_dt / dt_dyn);
  if (ncycles_override > 0) ncycles = ncycles_override;
  dt_dyn = crm_dt / ncycles;
  for (int c = 0; c < ncycles; c++) {
    from_halo(o, tr, dtr, nt); /* Dycore.h:156-159 */
    compute_tendencies(o, q, dq, tr, dtr, dt_dyn);
    rk_update(o, 1, dt_dyn, q, tr, q1, t1, q1, t1, dq, dtr);
    compute_tendencies(o, q1, dq, t1, dtr, (1. / 4.) * dt_dyn);
    rk_update(o, 2, dt_dyn, q, tr, q1, t1, q1, t1, dq, dtr);
    compute_tendencies(o, q1, dq, t1, dtr, (2. / 3.) * dt_dyn);
    rk_update(o, 3, dt_dyn, q, tr, q1, t1, q, tr, dq, dtr);
  }
  dynamics_to_coupler(o, q, tr, f);
  free(q); free(tr); free(q1); free(t1); free(dq); free(dtr);
  return ncycles;
}

/* Dycore.h:1392-1504 (declare_current_profile_as_hydrostatic), grav_balance branch, intended semantics;
 * the column mean is accumulated in (j,i) order (the reference uses atomics, order unspecified).
 * gcm = NULL or {gcm_density_dry, gcm_temp, gcm_water_vapor, gcm_cloud_water, gcm_cloud_ice}, each (nz,nens). */
int orc_declare_hydrostatic(void *h, double *const *f, const double *const *gcm) {
  orc_t *o = (orc_t *)h;
  int nz = o->nz, ny = o->ny, nx = o->nx, nens = o->nens, nt = o->nt;
  size_t nh = halo_count(o);
  double *q = (double *)calloc(NS * nh, sizeof(double)), *tr = (double *)calloc((nt + 1) * nh, sizeof(double));
  double *pr = (double *)calloc(nh, sizeof(double));
  if (gcm) {
    for (int k = 0; k < nz; k++)
      for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++)
          for (int e = 0; e < nens; e++) {
            int c = k * nens + e;
            double rho_d = gcm[0][c], rho_v = gcm[2][c];
            double rho = gcm[0][c] + gcm[2][c] + gcm[3][c] + gcm[4][c];
            double p = (rho_d * o->R_d + rho_v * o->R_v) * gcm[1][c];
            q[H5(idR, HS + k, HS + j, HS + i, e)] = rho;
            q[H5(idT, HS + k, HS + j, HS + i, e)] = pow(p / o->C0, 1. / o->gamma);
          }
  } else {
    coupler_to_dynamics(o, f, q, tr);
  }
  for (int k = 0; k < nz; k++)
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          pr[H4(HS + k, HS + j, HS + i, e)] = o->C0 * pow(q[H5(idT, HS + k, HS + j, HS + i, e)], o->gamma);
          q[H5(idT, HS + k, HS + j, HS + i, e)] /= q[H5(idR, HS + k, HS + j, HS + i, e)];
        }
  halo_exchange(o, q, tr, pr);
  double *pint = (double *)calloc((size_t)(nz + 1) * ny * nx * nens, sizeof(double));
  size_t sz = (size_t)NYH * NXH * nens;
  for (int k = 0; k < nz + 1; k++)
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          double hiL[ORD][ORD], loL[HS][HS][HS], hiR[ORD][ORD], loR[HS][HS][HS], sten[ORD];
          load_vert(o, k, e, hiL, loL);
          load_vert(o, k + 1, e, hiR, loR);
          for (int s = 0; s < ORD; s++) sten[s] = pr[H4(k, HS + j, HS + i, e) + s * sz];
          double pL = reconstruct(sten, hiL, loL, o->idl, o->sigma, 1);
          for (int s = 0; s < ORD; s++) sten[s] = pr[H4(k, HS + j, HS + i, e) + (1 + s) * sz];
          double pR = reconstruct(sten, hiR, loR, o->idl, o->sigma, 0);
          pint[(((size_t)k * ny + j) * nx + i) * nens + e] = 0.5 * (pL + pR);
        }
  double r_nx_ny = 1. / (nx * ny);
  for (int k = 0; k < nz; k++)
    for (int e = 0; e < nens; e++) {
      double acc = 0;
      for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++) {
          double dens = q[H5(idR, HS + k, HS + j, HS + i, e)];
          acc += -(pint[(((size_t)(k + 1) * ny + j) * nx + i) * nens + e] - pint[(((size_t)k * ny + j) * nx + i) * nens + e]) /
                 (dens * o->dz[k * nens + e]) * r_nx_ny;
        }
      o->grav_var[k * nens + e] = acc;
    }
  free(q); free(tr); free(pr); free(pint);
  return 0;
}

/* Dycore.h:37-60 (compute_mass): out[(nt+2)*nens] = mean over cells of q*dz for tracers, rho, rho*theta */
int orc_compute_mass(void *h, double *const *f, double *out) {
  orc_t *o = (orc_t *)h;
  int nt = o->nt;
  size_t nh = halo_count(o);
  double *q = (double *)calloc(NS * nh, sizeof(double)), *tr = (double *)calloc((nt + 1) * nh, sizeof(double));
  coupler_to_dynamics(o, f, q, tr);
  for (int v = 0; v < nt + 2; v++)
    for (int e = 0; e < o->nens; e++) {
      double s = 0;
      for (int k = 0; k < o->nz; k++)
        for (int j = 0; j < o->ny; j++)
          for (int i = 0; i < o->nx; i++) {
            size_t hh = H4(HS + k, HS + j, HS + i, e);
            double val = v < nt ? tr[(size_t)v * nh + hh] : (v == nt ? q[idR * nh + hh] : q[idT * nh + hh]);
            s += val * o->dz[k * o->nens + e];
          }
      out[v * o->nens + e] = s / ((double)o->nz * o->ny * o->nx);
    }
  free(q); free(tr);
  return 0;
}

/* ------------------------------------------------------------------------------------------------ */
/* The reference's public methods on its HALOED arrays (Dycore.h:37, :262, :608, :1281, :1336), for the */
/* parity tests of the C ABI's awfl_*_haloed / awfl_convert_* / awfl_halo_exchange entry points.        */
/* st (5,nz+2hs,ny+2hs,nx+2hs,nens), tr (nt,...), pr (nz+2hs,ny+2hs,nx+2hs,nens), caller-owned.         */
/* ------------------------------------------------------------------------------------------------ */
void orc_halo_shape(void *h, int *dims) { orc_t *o = (orc_t *)h; dims[0] = NZH; dims[1] = NYH; dims[2] = NXH; }
int orc_halo_exchange(void *h, double *st, double *tr, double *pr) { halo_exchange((orc_t *)h, st, tr, pr); return 0; }
int orc_compute_tendencies_haloed(void *h, double *st, double *st_tend, double *tr, double *tr_tend, double dt) {
  compute_tendencies((orc_t *)h, st, st_tend, tr, tr_tend, dt);
  return 0;
}
int orc_convert_coupler_to_dynamics(void *h, double *const *f, double *st, double *tr) {
  coupler_to_dynamics((orc_t *)h, f, st, tr);
  return 0;
}
int orc_convert_dynamics_to_coupler(void *h, const double *st, const double *tr, double *const *f) {
  dynamics_to_coupler((orc_t *)h, st, tr, f);
  return 0;
}
/* Dycore.h:37-60 with the reference's arguments */
int orc_compute_mass_dynamics(void *h, const double *q, const double *tr, double *out) {
  orc_t *o = (orc_t *)h;
  int nt = o->nt;
  size_t nh = halo_count(o);
  for (int v = 0; v < nt + 2; v++)
    for (int e = 0; e < o->nens; e++) {
      double s = 0;
      for (int k = 0; k < o->nz; k++)
        for (int j = 0; j < o->ny; j++)
          for (int i = 0; i < o->nx; i++) {
            size_t hh = H4(HS + k, HS + j, HS + i, e);
            double val = v < nt ? tr[(size_t)v * nh + hh] : (v == nt ? q[idR * nh + hh] : q[idT * nh + hh]);
            s += val * o->dz[k * o->nens + e];
          }
      out[v * o->nens + e] = s / ((double)o->nz * o->ny * o->nx);
    }
  return 0;
}
