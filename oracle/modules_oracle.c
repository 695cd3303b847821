/* TEST INFRASTRUCTURE ONLY -- not part of the product path.
 *
 * Plain-C restatement of the coupler modules either side of the dycore, serial, in the loop order of
 * the reference's serial backend (YAKL parallel_for = nested loops, last index fastest), so that the atomicAdd
 * accumulations happen in (j,i) order. Each function cites the reference lines it follows. Pinned: bit-identical to
 * the compiled reference (oracle/_ref, tests/test_modules_oracle.py) and to tests/golden/modules_*.npz.
 * Layout: 4-D fields (nz,ny,nx,nens), columns (nz,nens), C order.
 * Built by oracle/Makefile with -O2 -ffp-contract=off like the compiled reference. */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846   /* glibc's value; -std=c11 hides it */
#endif

#define IDX4(k, j, i, e) ((((long)(k) * ny + (j)) * nx + (i)) * nens + (e))
#define IDX2(k, e) ((long)(k) * nens + (e))

/* modules::sponge_layer, pam_core/modules/sponge_layer.h:9-96. fields: density_dry,uvel,vvel,wvel,temp,tracers. */
void pamo_sponge_layer(int nz, int ny, int nx, int nens, double **fields, int num_fields, const double *zint, const double *zmid,
                       double crm_dt, double time_scale, int num_layers) {
  const int WFLD = 3;                                             /* :32 */
  double *havg = (double *)calloc((size_t)num_fields * nz * nens, sizeof(double));   /* :66-68 */
  const double r_nx_ny = 1.0 / (nx * ny);                         /* :63 */
  for (int ifld = 0; ifld < num_fields; ifld++)                   /* :71-74 */
    for (int kloc = 0; kloc < num_layers; kloc++)
      for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++)
          for (int e = 0; e < nens; e++) {
            const int k = nz - 1 - kloc;
            if (ifld != WFLD) havg[((long)ifld * nz + k) * nens + e] += fields[ifld][IDX4(k, j, i, e)] * r_nx_ny;
          }
  const double time_factor = crm_dt / time_scale;                 /* :82 */
  for (int ifld = 0; ifld < num_fields; ifld++)                   /* :85-92 */
    for (int kloc = 0; kloc < num_layers; kloc++)
      for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++)
          for (int e = 0; e < nens; e++) {
            const int k = nz - 1 - kloc;
            const double rel_dist = (zint[IDX2(nz, e)] - zmid[IDX2(k, e)]) / (zint[IDX2(nz, e)] - zmid[IDX2(nz - 1 - (num_layers - 1), e)]);
            const double space_factor = (cos(M_PI * rel_dist) + 1) / 2;
            const double factor = space_factor * time_factor;
            double *f = &fields[ifld][IDX4(k, j, i, e)];
            *f += (havg[((long)ifld * nz + k) * nens + e] - *f) * factor;
          }
  free(havg);
}

/* Field bundles, in the order of include/pam_modules_b200.h */
typedef struct { double *density_dry, *uvel, *vvel, *temp, *water_vapor, *cloud_water, *ice, *cloud_water_num, *ice_num, *rain_num; } pamo_crm;
typedef struct { const double *gcm_density_dry, *gcm_uvel, *gcm_vvel, *gcm_temp, *gcm_water_vapor, *gcm_cloud_water, *gcm_cloud_ice,
                              *gcm_num_liq, *gcm_num_ice, *gcm_num_rain; } pamo_gcm;
typedef struct { double *rho_d, *uvel, *vvel, *temp, *qtot, *qv, *ql, *qi, *rho_v, *rho_l, *rho_i, *nc, *ni, *nr; } pamo_tend;

/* modules::compute_gcm_forcing_tendencies, pam_core/modules/gcm_forcing.h:18-205 */
void pamo_compute_gcm_forcing_tendencies(int nz, int ny, int nx, int nens, const pamo_crm *c, const pamo_gcm *g, const pamo_tend *t,
                                         double dt_gcm) {
  const long n2 = (long)nz * nens;
  double *col = (double *)calloc((size_t)10 * n2, sizeof(double));        /* :62-100 */
  double *a_rd = col, *a_u = col + n2, *a_v = col + 2 * n2, *a_t = col + 3 * n2, *a_qv = col + 4 * n2, *a_ql = col + 5 * n2,
         *a_qi = col + 6 * n2, *a_nc = col + 7 * n2, *a_ni = col + 8 * n2, *a_nr = col + 9 * n2;
  const double r_nx_ny = 1.0 / (nx * ny);                                 /* :102 */
  for (int k = 0; k < nz; k++)                                            /* :103-131 */
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          const long q = IDX4(k, j, i, e), ke = IDX2(k, e);
          a_rd[ke] += c->density_dry[q] * r_nx_ny;
          a_u[ke] += c->uvel[q] * r_nx_ny;
          a_v[ke] += c->vvel[q] * r_nx_ny;
          a_t[ke] += c->temp[q] * r_nx_ny;
          const double tmp_qv = c->water_vapor[q] / (c->density_dry[q] + c->water_vapor[q]);
          const double tmp_ql = c->cloud_water[q] / (c->density_dry[q] + c->water_vapor[q]);
          const double tmp_qi = c->ice[q] / (c->density_dry[q] + c->water_vapor[q]);
          a_qv[ke] += tmp_qv * r_nx_ny;
          a_ql[ke] += tmp_ql * r_nx_ny;
          a_qi[ke] += tmp_qi * r_nx_ny;
          a_nc[ke] += c->cloud_water_num[q] * r_nx_ny;
          a_ni[ke] += c->ice_num[q] * r_nx_ny;
          a_nr[ke] += c->rain_num[q] * r_nx_ny;
        }
  const double r_dt_gcm = 1.0 / dt_gcm;                                   /* :167 */
  for (int k = 0; k < nz; k++)                                            /* :171-203 */
    for (int e = 0; e < nens; e++) {
      const long ke = IDX2(k, e);
      t->rho_d[ke] = (g->gcm_density_dry[ke] - a_rd[ke]) * r_dt_gcm;
      t->uvel[ke] = (g->gcm_uvel[ke] - a_u[ke]) * r_dt_gcm;
      t->vvel[ke] = (g->gcm_vvel[ke] - a_v[ke]) * r_dt_gcm;
      t->temp[ke] = (g->gcm_temp[ke] - a_t[ke]) * r_dt_gcm;
      const double tmp_qv_gcm = g->gcm_water_vapor[ke] / (g->gcm_density_dry[ke] + g->gcm_water_vapor[ke]);
      const double tmp_ql_gcm = g->gcm_cloud_water[ke] / (g->gcm_density_dry[ke] + g->gcm_water_vapor[ke]);
      const double tmp_qi_gcm = g->gcm_cloud_ice[ke] / (g->gcm_density_dry[ke] + g->gcm_water_vapor[ke]);
      t->qv[ke] = (tmp_qv_gcm - a_qv[ke]) * r_dt_gcm;
      t->ql[ke] = (tmp_ql_gcm - a_ql[ke]) * r_dt_gcm;
      t->qi[ke] = (tmp_qi_gcm - a_qi[ke]) * r_dt_gcm;
      t->nc[ke] = (g->gcm_num_liq[ke] - a_nc[ke]) * r_dt_gcm;
      t->ni[ke] = (g->gcm_num_ice[ke] - a_ni[ke]) * r_dt_gcm;
      t->nr[ke] = (g->gcm_num_rain[ke] - a_nr[ke]) * r_dt_gcm;
      t->qtot[ke] = t->qv[ke] + t->ql[ke] + t->qi[ke];
    }
  free(col);
}

/* fill_holes, pam_core/modules/gcm_forcing.h:208-289 */
static void fill_holes(int nz, int ny, int nx, int nens, double *rho_x, const double *neg_mass, const double *dz) {
  const long n2 = (long)nz * nens;
  double *pos_mass = (double *)calloc((size_t)n2, sizeof(double));                         /* :225-230 */
  for (int k = 0; k < nz; k++)                                                              /* :233-235 */
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++)
          if (rho_x[IDX4(k, j, i, e)] > 0) pos_mass[IDX2(k, e)] += rho_x[IDX4(k, j, i, e)] * dz[IDX2(k, e)];
  int neg_too_large = 0;                                                                    /* :238 */
  for (int k = 0; k < nz; k++)                                                              /* :241-251 */
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          if (i == 0 && j == 0) { if (neg_mass[IDX2(k, e)] > pos_mass[IDX2(k, e)]) neg_too_large = 1; }
          if (pos_mass[IDX2(k, e)] > 0) {
            const long q = IDX4(k, j, i, e);
            const double factor = rho_x[q] * dz[IDX2(k, e)] / pos_mass[IDX2(k, e)];
            const double v = rho_x[q] - (neg_mass[IDX2(k, e)] * factor) / dz[IDX2(k, e)];
            rho_x[q] = 0.0 < v ? v : 0.0;   /* yakl::max = std::max (YAKL.h:36): a NaN (0/0 for a member without mass) gives 0 */
          }
        }
  if (neg_too_large) {                                                                      /* :253-287 */
    double *neg_glob = (double *)calloc((size_t)nens, sizeof(double)), *pos_glob = (double *)calloc((size_t)nens, sizeof(double));
    for (int k = 0; k < nz; k++)
      for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++)
          for (int e = 0; e < nens; e++) {
            if (i == 0 && j == 0) {
              const double d = neg_mass[IDX2(k, e)] - pos_mass[IDX2(k, e)];
              neg_glob[e] += 0.0 < d ? d : 0.0;
            }
            pos_glob[e] += rho_x[IDX4(k, j, i, e)] * dz[IDX2(k, e)];
          }
    for (int k = 0; k < nz; k++)
      for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++)
          for (int e = 0; e < nens; e++) {
            const long q = IDX4(k, j, i, e);
            const double factor = rho_x[q] * dz[IDX2(k, e)] / pos_glob[e];
            const double v = rho_x[q] - (neg_glob[e] * factor) / dz[IDX2(k, e)];
            rho_x[q] = 0.0 < v ? v : 0.0;   /* yakl::max = std::max (YAKL.h:36): a NaN (0/0 for a member without mass) gives 0 */
          }
    free(neg_glob); free(pos_glob);
  }
  free(pos_mass);
}

/* modules::apply_gcm_forcing_tendencies, pam_core/modules/gcm_forcing.h:294-436 */
void pamo_apply_gcm_forcing_tendencies(int nz, int ny, int nx, int nens, const pamo_crm *c, const pamo_gcm *g, const pamo_tend *t,
                                       const double *dz, double dt, double dt_gcm) {
  const long n2 = (long)nz * nens;
  double *w = (double *)calloc((size_t)6 * n2, sizeof(double));                             /* :343-364 */
  double *neg_v = w, *neg_l = w + n2, *neg_i = w + 2 * n2, *avg_v = w + 3 * n2, *avg_l = w + 4 * n2, *avg_i = w + 5 * n2;
  for (long q = 0; q < n2; q++) { t->rho_v[q] = 0; t->rho_l[q] = 0; t->rho_i[q] = 0; }
  const double r_nx_ny = 1.0 / (nx * ny);                                                   /* :366 */
  for (int k = 0; k < nz; k++)                                                              /* :367-414 */
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          const long q = IDX4(k, j, i, e), ke = IDX2(k, e);
          const double rho_d_old = c->density_dry[q];
          c->density_dry[q] += t->rho_d[ke] * dt;
          c->uvel[q] += t->uvel[ke] * dt;
          c->vvel[q] += t->vvel[ke] * dt;
          c->temp[q] += t->temp[ke] * dt;
          const double tmp_qv_old = c->water_vapor[q] / (rho_d_old + c->water_vapor[q]);
          const double tmp_ql_old = c->cloud_water[q] / (rho_d_old + c->water_vapor[q]);
          const double tmp_qi_old = c->ice[q] / (rho_d_old + c->water_vapor[q]);
          const double tmp_qv_new = (tmp_qv_old + t->qv[ke] * dt);
          const double tmp_ql_new = (tmp_ql_old + t->ql[ke] * dt);
          const double tmp_qi_new = (tmp_qi_old + t->qi[ke] * dt);
          c->water_vapor[q] = tmp_qv_new * c->density_dry[q] / (1 - tmp_qv_new);
          c->cloud_water[q] = tmp_ql_new * (c->density_dry[q] + c->water_vapor[q]);
          c->ice[q] = tmp_qi_new * (c->density_dry[q] + c->water_vapor[q]);
          c->cloud_water_num[q] += t->nc[ke] * dt;
          c->ice_num[q] += t->ni[ke] * dt;
          c->rain_num[q] += t->nr[ke] * dt;
          if (c->cloud_water_num[q] < 0) c->cloud_water_num[q] = 0;
          if (c->ice_num[q] < 0) c->ice_num[q] = 0;
          if (c->rain_num[q] < 0) c->rain_num[q] = 0;
          avg_v[ke] += c->water_vapor[q] * r_nx_ny;
          avg_l[ke] += c->cloud_water[q] * r_nx_ny;
          avg_i[ke] += c->ice[q] * r_nx_ny;
          if (c->water_vapor[q] < 0) { neg_v[ke] += -c->water_vapor[q] * dz[ke]; c->water_vapor[q] = 0; }
          if (c->cloud_water[q] < 0) { neg_l[ke] += -c->cloud_water[q] * dz[ke]; c->cloud_water[q] = 0; }
          if (c->ice[q] < 0) { neg_i[ke] += -c->ice[q] * dz[ke]; c->ice[q] = 0; }
        }
  const double r_dt_gcm = 1.0 / dt_gcm;                                                     /* :418-424 */
  for (long q = 0; q < n2; q++) {
    t->rho_v[q] = (g->gcm_water_vapor[q] - avg_v[q]) * r_dt_gcm;
    t->rho_l[q] = (g->gcm_cloud_water[q] - avg_l[q]) * r_dt_gcm;
    t->rho_i[q] = (g->gcm_cloud_ice[q] - avg_i[q]) * r_dt_gcm;
  }
  double sv = 0, sl = 0, si = 0;                                                            /* :427-431 (yakl::intrinsics::sum) */
  for (long q = 0; q < n2; q++) { sv += neg_v[q]; sl += neg_l[q]; si += neg_i[q]; }
  if (sv > 0) fill_holes(nz, ny, nx, nens, c->water_vapor, neg_v, dz);
  if (sl > 0) fill_holes(nz, ny, nx, nens, c->cloud_water, neg_l, dz);
  if (si > 0) fill_holes(nz, ny, nx, nens, c->ice, neg_i, dz);
  free(w);
}

/* yakl::Random = B. Jenkins' public-domain "smallprng" (externals/YAKL/src/YAKL_random.h:24-74) */
static double smallprng_first(unsigned long long seed) {
  unsigned long long a = 0xf1ea5eedULL, b = seed, c = seed, d = seed;
  for (int i = 0; i < 21; i++) {           /* set_seed: 20 warm-up rounds; then the draw */
    const unsigned long long e = a - ((b << 7) | (b >> 57));
    a = b ^ ((c << 13) | (c >> 51));
    b = c + ((d << 37) | (d >> 27));
    c = d + e;
    d = e + a;
  }
  return (double)d / (double)18446744073709551615ULL;
}

/* modules::perturb_temperature, pam_core/modules/perturb_temperature.h:10-63 */
void pamo_perturb_temperature(int nz, int ny, int nx, int nens, double *temp, const int *id, double magnitude) {
  const int num_levels = nz / 4;                                                            /* :22 */
  const long n2 = (long)nz * nens;
  double *m1 = (double *)calloc((size_t)n2, sizeof(double)), *m2 = (double *)calloc((size_t)n2, sizeof(double));
  const double r_nx_ny = 1.0 / (nx * ny);
  for (int k = 0; k < nz; k++)                                                              /* :38-41 */
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) m1[IDX2(k, e)] += temp[IDX4(k, j, i, e)] * r_nx_ny;
  for (int k = 0; k < num_levels; k++)                                                      /* :43-54 */
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) {
          const int seed = id[e] * num_levels * ny * nx + k * ny * nx + j * nx + i;
          double r = smallprng_first((unsigned long long)(seed * 1e3)) * 2.0 - 1.0;
          r = r < 1.0 ? r : 1.0;
          r = r > -1.0 ? r : -1.0;
          const double scaling = (num_levels - (double)k) / num_levels;
          temp[IDX4(k, j, i, e)] += r * magnitude * scaling;
          m2[IDX2(k, e)] += temp[IDX4(k, j, i, e)] * r_nx_ny;
        }
  for (int k = 0; k < num_levels; k++)                                                      /* :57-61 */
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int e = 0; e < nens; e++) temp[IDX4(k, j, i, e)] = temp[IDX4(k, j, i, e)] * m1[IDX2(k, e)] / m2[IDX2(k, e)];
  free(m1); free(m2);
}
