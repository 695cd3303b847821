/* pam_modules_b200.h -- C ABI of the coupler modules that run either side of the dycore in PAM's step loop
 * (standalone/mmf_simplified/driver.cpp:246-250), as hand-written sm_100a CUDA kernels inside libawfl_b200.so.
 * SURVEY.md 8(f) row 1 and row 4. Plain pointers and sizes; every array is a DEVICE pointer to FP64 data in the
 * coupler's layout (C order, ensemble member fastest): 4-D fields (nz,ny,nx,nens), column arrays (nz,nens).
 * All entry points return 0 on success; pamm_last_error() holds the message otherwise. Work is enqueued on the
 * workspace's stream and the calls do not synchronise the host unless stated.
 *
 * The reference accumulates its column means with atomicAdd (order-dependent on a GPU); these kernels sum every
 * (level, member) column in (j,i) order -- the order of the reference's serial CPU backend -- so results are
 * reproducible run to run. */
#ifndef PAM_MODULES_B200_H
#define PAM_MODULES_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pamm_workspace pamm_workspace;

/* Scratch for the modules on one coupler shape (no allocation inside a call). stream: cudaStream_t or NULL. */
int pamm_create(int nz, int ny, int nx, int nens, int device, void *stream, pamm_workspace **out);
void pamm_destroy(pamm_workspace *w);
const char *pamm_last_error(const pamm_workspace *w);
long pamm_launch_count(const pamm_workspace *w);

/* modules::sponge_layer (pam_core/modules/sponge_layer.h:9-96): relax the top num_layers levels of every field
 * towards its horizontal mean (w towards 0) with strength crm_dt/time_scale * ((cos(pi*rel_dist)+1)/2).
 * fields: HOST array of num_fields device pointers in the reference's order: density_dry, uvel, vvel, wvel, temp,
 * then the tracers. zint (nz+1,nens), zmid (nz,nens). */
int pamm_sponge_layer(pamm_workspace *w, double *const *fields, int num_fields, const double *zint, const double *zmid,
                      double crm_dt, double time_scale, int num_layers);

/* CRM fields read (compute) or updated in place (apply) by the GCM forcing. */
typedef struct {
  double *density_dry, *uvel, *vvel, *temp, *water_vapor, *cloud_water, *ice, *cloud_water_num, *ice_num, *rain_num;
} pamm_crm_fields;

/* GCM column state (nz,nens). */
typedef struct {
  const double *gcm_density_dry, *gcm_uvel, *gcm_vvel, *gcm_temp, *gcm_water_vapor, *gcm_cloud_water, *gcm_cloud_ice;
  const double *gcm_num_liq, *gcm_num_ice, *gcm_num_rain;
} pamm_gcm_columns;

/* The DataManager entries "gcm_forcing_tend_*" (nz,nens), gcm_forcing.h:134-149. */
typedef struct {
  double *rho_d, *uvel, *vvel, *temp, *qtot, *qv, *ql, *qi, *rho_v, *rho_l, *rho_i, *nc, *ni, *nr;
} pamm_forcing_tend;

/* modules::compute_gcm_forcing_tendencies (gcm_forcing.h:18-205): (GCM column - CRM column mean) / gcm_physics_dt. */
int pamm_compute_gcm_forcing_tendencies(pamm_workspace *w, const pamm_crm_fields *crm, const pamm_gcm_columns *gcm,
                                        const pamm_forcing_tend *tend, double gcm_physics_dt);

/* modules::apply_gcm_forcing_tendencies + fill_holes (gcm_forcing.h:208-436): apply the tendencies over crm_dt,
 * clip the number concentrations, diagnose the density forcings, fill negative water masses level by level and,
 * where a level cannot pay for its holes, over the whole CRM. dz (nz,nens). No host synchronisation: the
 * reference's two host-side decisions (any negative mass? level budget exceeded?) are taken on the device, per
 * (level, member) resp. per member: where the reference's global decision would run a stage on a level or member
 * that does not need it, that stage is the identity, so the results are the same. */
int pamm_apply_gcm_forcing_tendencies(pamm_workspace *w, const pamm_crm_fields *crm, const pamm_gcm_columns *gcm,
                                      const pamm_forcing_tend *tend, const double *dz, double crm_dt, double gcm_physics_dt);

/* modules::broadcast_initial_gcm_column (pam_core/modules/broadcast_initial_gcm_column.h:7-41): fill density_dry,
 * uvel, vvel, wvel, temp, water_vapor with the GCM columns gcm_density_dry, gcm_uvel, gcm_vvel, gcm_wvel, gcm_temp,
 * gcm_water_vapor. Both arguments: HOST arrays of 6 device pointers in that order. */
int pamm_broadcast_initial_gcm_column(pamm_workspace *w, const double *const *gcm_columns, double *const *fields);

/* modules::perturb_temperature (pam_core/modules/perturb_temperature.h:10-63): add a random perturbation of at most
 * +-magnitude*(num_levels-k)/num_levels K (num_levels = nz/4) to temp in the lowest num_levels levels, then rescale
 * each perturbed level to its original horizontal mean. The random numbers are the reference's: yakl::Random (the
 * public-domain "smallprng" of B. Jenkins, externals/YAKL/src/YAKL_random.h:24-74) seeded per cell with
 * (id(iens)*num_levels*ny*nx + k*ny*nx + j*nx + i)*1e3, first draw. id: DEVICE array of nens ints. */
int pamm_perturb_temperature(pamm_workspace *w, double *temp, const int *id, double magnitude);

#ifdef __cplusplus
}
#endif
#endif
