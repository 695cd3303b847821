/* TEST INFRASTRUCTURE ONLY -- not part of the product path.
 *
 * Plain-C restatement of the AWFL dycore's idealized initial states: the supercell
 * sounding (Dycore::init_supercell, dynamics/awfl/Dycore.h:1096-1276, point functions :774-830) and the rising
 * thermal (Dycore::init, DATA_SPEC_THERMAL branch :1017-1086, point functions :735-769), each followed by
 * convert_dynamics_to_coupler (:1281-1331). Serial, in the reference's loop order. Pinned: bit-identical to the compiled
 * reference (tests/test_init_oracle.py) and to tests/golden/init_*.npz.
 *
 * One intended-semantics point: the reference's cell-average kernel writes hy_dens_cells / hy_pressure_cells for EVERY
 * member from each (k,iens) thread (a shadowed loop variable, Dycore.h:1228-1231), i.e. a race whose serial outcome is
 * "the last member's column everywhere". Every member gets its OWN column here; the two agree whenever the members share
 * one vertical grid, which is all the reference's drivers ever set up (driver.cpp:176: zint is 1-D).
 * Layout: 4-D fields (nz,ny,nx,nens), columns (nz,nens), C order. */
#include <math.h>
#include <stdlib.h>

#include "awfl_tables.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#define IDX4(k, j, i, e) ((((long)(k) * ny + (j)) * nx + (i)) * nens + (e))
#define IDX2(k, e) ((long)(k) * nens + (e))

static const double gll_pts[9] = AWFL_GLL9_PTS_INIT;
static const double gll_wts[9] = AWFL_GLL9_WTS_INIT;

void pamo_gll9(double *pts, double *wts) { for (int i = 0; i < 9; i++) { pts[i] = gll_pts[i]; wts[i] = gll_wts[i]; } }

/* Dycore.h:774-784 */
static double sc_temperature(double z, double z_0, double z_trop, double z_top, double T_0, double T_trop, double T_top) {
  if (z <= z_trop) {
    const double lapse = -(T_trop - T_0) / (z_trop - z_0);
    return T_0 - lapse * (z - z_0);
  } else {
    const double lapse = -(T_top - T_trop) / (z_top - z_trop);
    return T_trop - lapse * (z - z_trop);
  }
}
/* Dycore.h:789-810 */
static double sc_pressure_dry(double z, double z_0, double z_trop, double z_top, double T_0, double T_trop, double T_top, double p_0,
                              double R_d, double grav) {
  if (z <= z_trop) {
    const double lapse = -(T_trop - T_0) / (z_trop - z_0);
    const double T = sc_temperature(z, z_0, z_trop, z_top, T_0, T_trop, T_top);
    return p_0 * pow(T / T_0, grav / (R_d * lapse));
  } else {
    double lapse = -(T_trop - T_0) / (z_trop - z_0);
    const double p_trop = p_0 * pow(T_trop / T_0, grav / (R_d * lapse));
    lapse = -(T_top - T_trop) / (z_top - z_trop);
    if (lapse != 0) {
      const double T = sc_temperature(z, z_0, z_trop, z_top, T_0, T_trop, T_top);
      return p_trop * pow(T / T_trop, grav / (R_d * lapse));
    } else {
      return p_trop * exp(-grav * (z - z_trop) / (R_d * T_trop));
    }
  }
}
/* Dycore.h:815-821 */
static double sc_relhum(double z, double z_0, double z_trop) {
  (void)z_0;
  if (z <= z_trop) return 1.0 - 0.75 * pow(z / z_trop, 1.25);
  return 0.25;
}
/* Dycore.h:826-828 */
static double sc_sat_mix_dry(double press, double T) { return 380 / (press)*exp(17.27 * (T - 273) / (T - 36)); }

/* water-vapour mixing ratio of the sounding at height z (Dycore.h:1158-1163, :1188-1193) */
static double sc_qv(double zloc, double z_top, double R_d, double grav, double *temp_out) {
  const double z_0 = 0, z_trop = 12000, T_0 = 300, T_trop = 213, T_top = 213, p_0 = 100000;
  const double temp = sc_temperature(zloc, z_0, z_trop, z_top, T_0, T_trop, T_top);
  const double press_dry = sc_pressure_dry(zloc, z_0, z_trop, z_top, T_0, T_trop, T_top, p_0, R_d, grav);
  const double qvs = sc_sat_mix_dry(press_dry, temp);
  double relhum = sc_relhum(zloc, z_0, z_trop);
  if (relhum * qvs > 0.014) relhum = 0.014 / qvs;
  const double x = qvs * relhum;
  *temp_out = temp;
  return x < 0.014 ? x : 0.014;      /* std::min(0.014, qvs*relhum): a NaN (pow of a height that rounded below zero) gives 0.014 */
}

/* Dycore.h:1281-1331 for one cell: dycore (rho, rho u, rho v, rho w, rho theta) + tracers -> coupler fields */
static void to_coupler(long q, double rho, double ru, double rv, double rw, double rt, const double *tr, int nt, int idWV,
                       const int *adds_mass, double R_d, double R_v, double gamma, double C0, double **f) {
  const double u = ru / rho, v = rv / rho, w = rw / rho, theta = rt / rho;
  const double press = C0 * pow(rho * theta, gamma);
  const double rho_v = tr[idWV];
  double rho_d = rho;
  for (int t = 0; t < nt; t++) if (adds_mass[t]) rho_d -= tr[t];
  const double temp = press / (rho_d * R_d + rho_v * R_v);
  f[0][q] = rho_d; f[1][q] = u; f[2][q] = v; f[3][q] = w; f[4][q] = temp;
  for (int t = 0; t < nt; t++) f[5 + t][q] = tr[t];
}

/* Dycore::init_supercell + convert_dynamics_to_coupler. fields: density_dry,uvel,vvel,wvel,temp, then nt tracers. */
void pamo_init_supercell(int nz, int ny, int nx, int nens, double xlen, double ylen, const double *zint, double R_d, double R_v,
                         double grav, double gamma, double C0, int idWV, int nt, const int *adds_mass, double **fields) {
  const int ngll = 9;
  const double p_0 = 100000;
  const double dx = xlen / nx, dy = ylen / ny;
  (void)dy;
  double *pg = (double *)malloc(sizeof(double) * nz * ngll), *vapg = (double *)malloc(sizeof(double) * nz * ngll);
  double *tr = (double *)calloc((size_t)(nt > 0 ? nt : 1), sizeof(double));
  for (int e = 0; e < nens; e++) {
    const double z_top = zint[IDX2(nz, e)];
    /* :1146-1182 hydrostatic integration of the pressure at the GLL points of every cell */
    pg[0] = p_0;
    for (int k = 0; k < nz; k++) {
      const double dz = zint[IDX2(k + 1, e)] - zint[IDX2(k, e)];                /* pam_coupler.h set_grid */
      const double zmid = 0.5 * (zint[IDX2(k, e)] + zint[IDX2(k + 1, e)]);
      for (int kk = 0; kk < ngll - 1; kk++) {
        const double b = zmid + gll_pts[kk] * dz, t = zmid + gll_pts[kk + 1] * dz, m = 0.5 * (b + t);
        const double sub = dz * (gll_pts[kk + 1] - gll_pts[kk]);
        double tot = 0;
        for (int kkk = 0; kkk < ngll; kkk++) {
          const double zloc = m + sub * gll_pts[kkk];
          double temp;
          const double qv = sc_qv(zloc, z_top, R_d, grav, &temp);
          tot += (-(1 + qv) * grav / (R_d + qv * R_v) / temp) * gll_wts[kkk];
        }
        tot *= dz * (gll_pts[kk + 1] - gll_pts[kk]);
        pg[k * ngll + kk + 1] = pg[k * ngll + kk] * exp(tot);
        if (kk == ngll - 2 && k < nz - 1) pg[(k + 1) * ngll] = pg[k * ngll + ngll - 1];
      }
    }
    for (int k = 0; k < nz; k++) {
      const double dz = zint[IDX2(k + 1, e)] - zint[IDX2(k, e)];
      const double zmid = 0.5 * (zint[IDX2(k, e)] + zint[IDX2(k + 1, e)]);
      /* :1185-1202 state at the GLL points; :1205-1232 cell averages */
      double press_tot = 0, dens_tot = 0;
      for (int kk = 0; kk < ngll; kk++) {
        const double zloc = zmid + gll_pts[kk] * dz;
        double temp;
        const double qv = sc_qv(zloc, z_top, R_d, grav, &temp);
        const double press = pg[k * ngll + kk];
        const double dens_dry = press / (R_d + qv * R_v) / temp;
        const double dens_vap = qv * dens_dry;
        vapg[k * ngll + kk] = dens_vap;
        press_tot += press * gll_wts[kk];
        dens_tot += (dens_dry + dens_vap) * gll_wts[kk];
      }
      /* :1235-1273 the cell's state */
      const double rho = dens_tot, rt = pow(press_tot / C0, 1.0 / gamma);
      for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++) {
          double ru = 0, rv = 0, rw = 0, wv = 0;
          for (int kk = 0; kk < ngll; kk++)
            for (int jj = 0; jj < ngll; jj++)
              for (int ii = 0; ii < ngll; ii++) {
                const double zloc = zmid + gll_pts[kk] * dz;
                const double zs = 5000, us = 30, uc = 15;
                const double uvel = zloc < zs ? us * (zloc / zs) - uc : us - uc;
                const double factor = gll_wts[ii] * gll_wts[jj] * gll_wts[kk];
                ru += rho * uvel * factor;
                rv += rho * 0.0 * factor;
                rw += rho * 0.0 * factor;
                wv += vapg[k * ngll + kk] * factor;
              }
          for (int t = 0; t < nt; t++) tr[t] = 0;
          tr[idWV] = wv;
          to_coupler(IDX4(k, j, i, e), rho, ru, rv, rw, rt, tr, nt, idWV, adds_mass, R_d, R_v, gamma, C0, fields);
        }
    }
  }
  (void)dx;
  free(pg); free(vapg); free(tr);
}

/* Dycore.h:742-752 */
static void hydro_const_theta(double z, double grav, double C0, double cp, double p0, double gamma, double rd, double *r, double *t) {
  const double theta0 = 300., exner0 = 1.;
  *t = theta0;
  const double exner = exner0 - grav * z / (cp * theta0);
  const double p = p0 * pow(exner, (cp / rd));
  const double rt = pow((p / C0), (1.0 / gamma));
  *r = rt / *t;
}
/* Dycore.h:757-769 */
static double sample_ellipse_cosine(double amp, double x, double y, double z, double x0, double y0, double z0, double xrad, double yrad,
                                    double zrad) {
  const double dist = sqrt(((x - x0) / xrad) * ((x - x0) / xrad) + ((y - y0) / yrad) * ((y - y0) / yrad) +
                           ((z - z0) / zrad) * ((z - z0) / zrad)) * M_PI / 2.;
  if (dist <= M_PI / 2.) return amp * pow(cos(dist), 2.0);
  return 0.;
}

/* Dycore::init, DATA_SPEC_THERMAL (Dycore.h:1017-1086) + convert_dynamics_to_coupler */
void pamo_init_thermal(int nz, int ny, int nx, int nens, double xlen, double ylen, const double *zint, double R_d, double R_v, double cp_d,
                       double p0, double grav, double gamma, double C0, int idWV, int nt, const int *adds_mass, double **fields) {
  const int nq = 9;
  const double dx = xlen / nx, dy = ylen / ny;
  const int sim2d = ny == 1;
  double *tr = (double *)calloc((size_t)(nt > 0 ? nt : 1), sizeof(double));
  for (int k = 0; k < nz; k++)
    for (int e = 0; e < nens; e++) {
      const double dz = zint[IDX2(k + 1, e)] - zint[IDX2(k, e)];
      const double zmid = 0.5 * (zint[IDX2(k, e)] + zint[IDX2(k + 1, e)]);
      double hy_dens = 0, hy_pres = 0;                                          /* :1034-1047 */
      for (int kk = 0; kk < nq; kk++) {
        const double z = zmid + gll_pts[kk] * dz;
        double hr, ht;
        hydro_const_theta(z, grav, C0, cp_d, p0, gamma, R_d, &hr, &ht);
        hy_dens += hr * gll_wts[kk];
        hy_pres += C0 * pow(hr * ht, gamma) * gll_wts[kk];
      }
      for (int j = 0; j < ny; j++)                                              /* :1050-1084 */
        for (int i = 0; i < nx; i++) {
          double rho_s = 0, ru = 0, rv = 0, rw = 0, rt = 0, wv = 0;
          for (int kk = 0; kk < nq; kk++)
            for (int jj = 0; jj < nq; jj++)
              for (int ii = 0; ii < nq; ii++) {
                const double x = (i + 0.5) * dx + gll_pts[ii] * dx;
                double y = (j + 0.5) * dy + gll_pts[jj] * dy;
                if (sim2d) y = ylen / 2;
                const double z = zmid + gll_pts[kk] * dz;
                const double hr = hy_dens, hp = hy_pres;
                const double ht = pow(hp / C0, 1.0 / gamma) / hr;
                const double rho = hr, u = 0, v = 0, w = 0, rho_v = 0;
                const double theta = ht + sample_ellipse_cosine(2.0, x, y, z, xlen / 2, ylen / 2, 2000., 2000., 2000., 2000.);
                const double wt = gll_wts[ii] * gll_wts[jj] * gll_wts[kk];
                rho_s += rho * wt;
                ru += rho * u * wt;
                rv += rho * v * wt;
                rw += rho * w * wt;
                rt += rho * theta * wt;
                wv += rho_v * wt;
              }
          for (int t = 0; t < nt; t++) tr[t] = 0;
          tr[idWV] = wv;
          to_coupler(IDX4(k, j, i, e), rho_s, ru, rv, rw, rt, tr, nt, idWV, adds_mass, R_d, R_v, gamma, C0, fields);
        }
    }
  free(tr);
}
