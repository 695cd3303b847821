// Coupler modules either side of the dycore in PAM's step loop, as sm_100a CUDA kernels behind the C ABI of
// include/pam_modules_b200.h (part of libawfl_b200.so). Reference behaviour: pam_core/modules/sponge_layer.h,
// gcm_forcing.h, perturb_temperature.h (cited per kernel).
//
// Layout: ensemble member fastest, so one warp = 32 consecutive members of one (level[, field]) column set and every
// global access is a coalesced 256-byte line. All of these modules reduce over the horizontal (j,i) of one level and
// member: a thread owns one (level, member) column and walks its ny*nx cells in (j,i) order, which is the summation
// order of the reference's serial backend (its GPU backend uses atomicAdd, i.e. no defined order). They are
// HBM-bound streaming kernels; there is NO CPU fallback.
#include <cuda_runtime.h>
#include <math.h>
#include <string>
#include <vector>

#include "../../include/pam_modules_b200.h"

struct pamm_workspace {
  int nz, ny, nx, nens, device;
  cudaStream_t stream;
  std::string err;
  long launches;
  double *scratch;      // 6 (nz,nens) arrays: negative / positive mass of the three water species
  int *flags;           // 3 ints: level budget exceeded (per species)
  double **fptr_d;      // device copy of the sponge layer's field pointer list
  int fptr_cap;
};

static std::string g_pamm_err;

#define PCK(call)                                                              \
  do {                                                                         \
    cudaError_t e_ = (call);                                                   \
    if (e_ != cudaSuccess) {                                                   \
      w->err = std::string(#call) + ": " + cudaGetErrorString(e_);             \
      return 1;                                                                \
    }                                                                          \
  } while (0)

namespace {

struct Dims { int nz, ny, nx, nens; };

// ---------------------------------------------------------------------------------------------------------------
// modules::sponge_layer (sponge_layer.h:9-96). One thread = (field, sponge level, member): horizontal mean in (j,i)
// order (:71-74; zero for the w field, :32), then the relaxation of every cell of that level (:86-92).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_sponge(const Dims d, double *const *fields, int num_fields, int num_layers,
                                                const double *zint, const double *zmid, double time_factor) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int kloc = blockIdx.y, ifld = blockIdx.z;
  if (e >= d.nens) return;
  const int k = d.nz - 1 - kloc;
  double *f = fields[ifld] + (long)k * d.ny * d.nx * d.nens + e;
  const int nh = d.ny * d.nx;
  const double r_nx_ny = 1.0 / (d.nx * d.ny);
  double havg = 0.0;
  if (ifld != 3) {
    for (int c = 0; c < nh; c++) havg += f[(long)c * d.nens] * r_nx_ny;
  }
  const double ztop = zint[(long)d.nz * d.nens + e];
  const double rel_dist = (ztop - zmid[(long)k * d.nens + e]) / (ztop - zmid[(long)(d.nz - 1 - (num_layers - 1)) * d.nens + e]);
  const double space_factor = (cos(M_PI * rel_dist) + 1) / 2;
  const double factor = space_factor * time_factor;
  for (int c = 0; c < nh; c++) {
    const double x = f[(long)c * d.nens];
    f[(long)c * d.nens] = x + (havg - x) * factor;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// modules::compute_gcm_forcing_tendencies (gcm_forcing.h:18-205). One thread = (level, member): the ten column
// means (:103-131) and the fourteen tendencies (:171-203).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_gcm_forcing_compute(const Dims d, const pamm_crm_fields c, const pamm_gcm_columns g,
                                                             const pamm_forcing_tend t, double r_dt_gcm) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y;
  if (e >= d.nens) return;
  const long base = (long)k * d.ny * d.nx * d.nens + e;
  const int nh = d.ny * d.nx;
  const double r_nx_ny = 1.0 / (d.nx * d.ny);
  double a_rd = 0, a_u = 0, a_v = 0, a_t = 0, a_qv = 0, a_ql = 0, a_qi = 0, a_nc = 0, a_ni = 0, a_nr = 0;
  for (int cc = 0; cc < nh; cc++) {
    const long i = base + (long)cc * d.nens;
    const double rd = c.density_dry[i], rv = c.water_vapor[i];
    a_rd += rd * r_nx_ny;
    a_u += c.uvel[i] * r_nx_ny;
    a_v += c.vvel[i] * r_nx_ny;
    a_t += c.temp[i] * r_nx_ny;
    a_qv += (rv / (rd + rv)) * r_nx_ny;
    a_ql += (c.cloud_water[i] / (rd + rv)) * r_nx_ny;
    a_qi += (c.ice[i] / (rd + rv)) * r_nx_ny;
    a_nc += c.cloud_water_num[i] * r_nx_ny;
    a_ni += c.ice_num[i] * r_nx_ny;
    a_nr += c.rain_num[i] * r_nx_ny;
  }
  const long ke = (long)k * d.nens + e;
  const double grd = g.gcm_density_dry[ke], grv = g.gcm_water_vapor[ke];
  t.rho_d[ke] = (grd - a_rd) * r_dt_gcm;
  t.uvel[ke] = (g.gcm_uvel[ke] - a_u) * r_dt_gcm;
  t.vvel[ke] = (g.gcm_vvel[ke] - a_v) * r_dt_gcm;
  t.temp[ke] = (g.gcm_temp[ke] - a_t) * r_dt_gcm;
  const double fqv = (grv / (grd + grv) - a_qv) * r_dt_gcm;
  const double fql = (g.gcm_cloud_water[ke] / (grd + grv) - a_ql) * r_dt_gcm;
  const double fqi = (g.gcm_cloud_ice[ke] / (grd + grv) - a_qi) * r_dt_gcm;
  t.qv[ke] = fqv; t.ql[ke] = fql; t.qi[ke] = fqi;
  t.nc[ke] = (g.gcm_num_liq[ke] - a_nc) * r_dt_gcm;
  t.ni[ke] = (g.gcm_num_ice[ke] - a_ni) * r_dt_gcm;
  t.nr[ke] = (g.gcm_num_rain[ke] - a_nr) * r_dt_gcm;
  t.qtot[ke] = fqv + fql + fqi;
}

// ---------------------------------------------------------------------------------------------------------------
// modules::apply_gcm_forcing_tendencies (gcm_forcing.h:294-436) with the level stage of fill_holes (:208-255) fused in.
// One thread = (level, member). Pass 1 applies the forcing to every cell of the level (:367-414) and accumulates the
// diagnostic column means and the negative mass of the three water species; the density forcings are diagnosed
// (:420-424); then, per species with negative mass on this level, the available positive mass is summed (:233-235)
// and taken away proportionally (:241-251). The reference decides on the host whether to call fill_holes at all
// (sum of the negative mass over ALL levels and members > 0, :427-431); for a (level, member) without negative mass
// the fill is the identity (x - 0*factor/dz), so deciding per thread gives the same result without a host sync.
// neg/pos are kept for the whole-CRM stage below.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void level_fill(double *x, long base, int nh, int nens, double dzk, double neg, double &pos_out) {
  double pos = 0.0;
  for (int cc = 0; cc < nh; cc++) {
    const double v = x[base + (long)cc * nens];
    if (v > 0) pos += v * dzk;
  }
  pos_out = pos;
  if (pos > 0) {
    for (int cc = 0; cc < nh; cc++) {
      const long i = base + (long)cc * nens;
      const double v = x[i];
      const double factor = v * dzk / pos;
      x[i] = fmax(0.0, v - (neg * factor) / dzk);
    }
  }
}

__global__ void __launch_bounds__(128) k_gcm_forcing_apply(const Dims d, const pamm_crm_fields c, const pamm_gcm_columns g,
                                                           const pamm_forcing_tend t, const double *dz, double dt, double r_dt_gcm,
                                                           double *scratch) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y;
  if (e >= d.nens) return;
  const long base = (long)k * d.ny * d.nx * d.nens + e;
  const long ke = (long)k * d.nens + e;
  const int nh = d.ny * d.nx;
  const double r_nx_ny = 1.0 / (d.nx * d.ny);
  const double f_rd = t.rho_d[ke], f_u = t.uvel[ke], f_v = t.vvel[ke], f_t = t.temp[ke];
  const double f_qv = t.qv[ke], f_ql = t.ql[ke], f_qi = t.qi[ke], f_nc = t.nc[ke], f_ni = t.ni[ke], f_nr = t.nr[ke];
  const double dzk = dz[ke];
  double avg_v = 0, avg_l = 0, avg_i = 0, neg_v = 0, neg_l = 0, neg_i = 0;
  for (int cc = 0; cc < nh; cc++) {
    const long i = base + (long)cc * d.nens;
    const double rho_d_old = c.density_dry[i];
    const double rd = rho_d_old + f_rd * dt;
    c.density_dry[i] = rd;
    c.uvel[i] += f_u * dt;
    c.vvel[i] += f_v * dt;
    c.temp[i] += f_t * dt;
    double rv = c.water_vapor[i], rl = c.cloud_water[i], ri = c.ice[i];
    const double qv_new = rv / (rho_d_old + rv) + f_qv * dt;
    const double ql_new = rl / (rho_d_old + rv) + f_ql * dt;
    const double qi_new = ri / (rho_d_old + rv) + f_qi * dt;
    rv = qv_new * rd / (1 - qv_new);
    rl = ql_new * (rd + rv);
    ri = qi_new * (rd + rv);
    double xc = c.cloud_water_num[i] + f_nc * dt, xi = c.ice_num[i] + f_ni * dt, xr = c.rain_num[i] + f_nr * dt;
    if (xc < 0) xc = 0;
    if (xi < 0) xi = 0;
    if (xr < 0) xr = 0;
    c.cloud_water_num[i] = xc; c.ice_num[i] = xi; c.rain_num[i] = xr;
    avg_v += rv * r_nx_ny; avg_l += rl * r_nx_ny; avg_i += ri * r_nx_ny;
    if (rv < 0) { neg_v += -rv * dzk; rv = 0; }
    if (rl < 0) { neg_l += -rl * dzk; rl = 0; }
    if (ri < 0) { neg_i += -ri * dzk; ri = 0; }
    c.water_vapor[i] = rv; c.cloud_water[i] = rl; c.ice[i] = ri;
  }
  t.rho_v[ke] = (g.gcm_water_vapor[ke] - avg_v) * r_dt_gcm;
  t.rho_l[ke] = (g.gcm_cloud_water[ke] - avg_l) * r_dt_gcm;
  t.rho_i[ke] = (g.gcm_cloud_ice[ke] - avg_i) * r_dt_gcm;
  const long nke = (long)d.nz * d.nens;
  double pos_v = 0, pos_l = 0, pos_i = 0;
  if (neg_v > 0) level_fill(c.water_vapor, base, nh, d.nens, dzk, neg_v, pos_v);
  if (neg_l > 0) level_fill(c.cloud_water, base, nh, d.nens, dzk, neg_l, pos_l);
  if (neg_i > 0) level_fill(c.ice, base, nh, d.nens, dzk, neg_i, pos_i);
  scratch[0 * nke + ke] = neg_v; scratch[1 * nke + ke] = pos_v;
  scratch[2 * nke + ke] = neg_l; scratch[3 * nke + ke] = pos_l;
  scratch[4 * nke + ke] = neg_i; scratch[5 * nke + ke] = pos_i;
}

// Whole-CRM stage of fill_holes (gcm_forcing.h:257-289): one thread = (member, species). The mass a level could not pay
// for, max(0, neg - pos), is summed over the levels and removed from every cell of the member in proportion to its mass.
// The reference runs this stage for all members as soon as ANY level of ANY member is short (host-side flag, :238-244,
// :257); for a member that is not short itself the stage is the identity, so each thread decides for its own member.
// A member without any mass of the species is left alone (the reference evaluates std::max(0, 0 - x*NaN) = 0 there: the same).
__global__ void __launch_bounds__(128) k_fill_holes_global(const Dims d, double *xv, double *xl, double *xi, const double *dz,
                                                           const double *scratch) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int sp = blockIdx.y;
  if (e >= d.nens) return;
  double *x = sp == 0 ? xv : (sp == 1 ? xl : xi);
  const long nke = (long)d.nz * d.nens;
  const double *neg = scratch + (2 * sp) * nke, *pos = scratch + (2 * sp + 1) * nke;
  double neg_glob = 0.0;
  for (int k = 0; k < d.nz; k++) neg_glob += fmax(0.0, neg[(long)k * d.nens + e] - pos[(long)k * d.nens + e]);
  if (!(neg_glob > 0)) return;
  const int nh = d.ny * d.nx;
  double pos_glob = 0.0;
  for (int k = 0; k < d.nz; k++) {
    const double dzk = dz[(long)k * d.nens + e];
    for (int cc = 0; cc < nh; cc++) pos_glob += x[((long)k * nh + cc) * d.nens + e] * dzk;
  }
  if (!(pos_glob > 0)) return;
  for (int k = 0; k < d.nz; k++) {
    const double dzk = dz[(long)k * d.nens + e];
    for (int cc = 0; cc < nh; cc++) {
      const long i = ((long)k * nh + cc) * d.nens + e;
      const double v = x[i];
      const double factor = v * dzk / pos_glob;
      x[i] = fmax(0.0, v - (neg_glob * factor) / dzk);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// modules::perturb_temperature (perturb_temperature.h:10-63). One thread = (level < nz/4, member): horizontal mean
// before (:34-37), perturbation with the reference's generator (:42-50), mean after, energy-conserving rescale (:54-58).
// yakl::Random = B. Jenkins' public-domain "smallprng" (YAKL_random.h:24-74): state (0xf1ea5eed, s, s, s), 20 warm-up
// rounds, first draw / 2^64.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long rot64(unsigned long long x, int k) { return (x << k) | (x >> (64 - k)); }
__device__ __forceinline__ double smallprng_first(unsigned long long seed) {
  unsigned long long a = 0xf1ea5eedULL, b = seed, c = seed, dd = seed;
  for (int i = 0; i < 21; i++) {       // 20 warm-up rounds + the draw
    const unsigned long long e = a - rot64(b, 7);
    a = b ^ rot64(c, 13);
    b = c + rot64(dd, 37);
    c = dd + e;
    dd = e + a;
  }
  return (double)dd / 18446744073709551615.0;
}

__global__ void __launch_bounds__(128) k_perturb_temperature(const Dims d, double *temp, const int *id, int num_levels, double magnitude) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y;
  if (e >= d.nens) return;
  const long base = (long)k * d.ny * d.nx * d.nens + e;
  const double r_nx_ny = 1.0 / (d.nx * d.ny);
  double m1 = 0.0, m2 = 0.0;
  const int ide = id[e];
  const double scaling = (num_levels - (double)k) / num_levels;
  for (int j = 0; j < d.ny; j++)
    for (int i = 0; i < d.nx; i++) {
      const long q = base + ((long)j * d.nx + i) * d.nens;
      double x = temp[q];
      m1 += x * r_nx_ny;
      const int seed = ide * num_levels * d.ny * d.nx + k * d.ny * d.nx + j * d.nx + i;     // int arithmetic, as the reference
      double r = smallprng_first((unsigned long long)(seed * 1e3)) * 2.0 - 1.0;
      r = fmin(r, 1.0);
      r = fmax(r, -1.0);
      x += r * magnitude * scaling;
      temp[q] = x;
      m2 += x * r_nx_ny;
    }
  for (int c = 0; c < d.ny * d.nx; c++) {
    const long q = base + (long)c * d.nens;
    temp[q] = temp[q] * m1 / m2;
  }
}

// modules::broadcast_initial_gcm_column (pam_core/modules/broadcast_initial_gcm_column.h:7-41): every cell of a level
// takes the GCM column's value. One thread = (level, member) writing its ny*nx cells of the six fields.
struct BroadcastArgs { const double *col[6]; double *fld[6]; };
__global__ void __launch_bounds__(128) k_broadcast_column(const Dims d, const BroadcastArgs a) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y, f = blockIdx.z;
  if (e >= d.nens) return;
  const double v = a.col[f][(long)k * d.nens + e];
  double *x = a.fld[f] + (long)k * d.ny * d.nx * d.nens + e;
  for (int c = 0; c < d.ny * d.nx; c++) x[(long)c * d.nens] = v;
}

}  // namespace

// =================================================================================================================
// C ABI
// =================================================================================================================
extern "C" {

int pamm_create(int nz, int ny, int nx, int nens, int device, void *stream, pamm_workspace **out) {
  *out = nullptr;
  if (nz < 1 || ny < 1 || nx < 1 || nens < 1) { g_pamm_err = "pamm_create: bad dimensions"; return 1; }
  if (cudaSetDevice(device) != cudaSuccess) { g_pamm_err = "pamm_create: cudaSetDevice failed (no CUDA device: this library has no CPU fallback)"; return 1; }
  pamm_workspace *w = new pamm_workspace;
  w->nz = nz; w->ny = ny; w->nx = nx; w->nens = nens; w->device = device;
  w->stream = (cudaStream_t)stream; w->launches = 0; w->scratch = nullptr; w->flags = nullptr; w->fptr_d = nullptr; w->fptr_cap = 0;
  if (cudaMalloc((void **)&w->scratch, sizeof(double) * 6 * (size_t)nz * nens) != cudaSuccess) {
    g_pamm_err = "pamm_create: cudaMalloc failed"; delete w; return 1;
  }
  *out = w;
  return 0;
}

void pamm_destroy(pamm_workspace *w) {
  if (!w) return;
  cudaFree(w->scratch); cudaFree(w->fptr_d);
  delete w;
}

const char *pamm_last_error(const pamm_workspace *w) { return w ? w->err.c_str() : g_pamm_err.c_str(); }
long pamm_launch_count(const pamm_workspace *w) { return w->launches; }

int pamm_sponge_layer(pamm_workspace *w, double *const *fields, int num_fields, const double *zint, const double *zmid,
                      double crm_dt, double time_scale, int num_layers) {
  if (num_fields < 5) { w->err = "pamm_sponge_layer: need at least the five state fields"; return 1; }
  if (num_layers < 1 || num_layers > w->nz) { w->err = "pamm_sponge_layer: sponge_num_layers out of range"; return 1; }
  if (!(time_scale > 0.0)) { w->err = "pamm_sponge_layer: sponge_time_scale must be positive"; return 1; }
  if (num_fields > w->fptr_cap) {
    cudaFree(w->fptr_d);
    PCK(cudaMalloc((void **)&w->fptr_d, sizeof(double *) * num_fields));
    w->fptr_cap = num_fields;
  }
  PCK(cudaMemcpyAsync(w->fptr_d, fields, sizeof(double *) * num_fields, cudaMemcpyHostToDevice, w->stream));
  const Dims d{w->nz, w->ny, w->nx, w->nens};
  const dim3 grid((unsigned)(w->nens + 127) / 128, (unsigned)num_layers, (unsigned)num_fields);
  k_sponge<<<grid, 128, 0, w->stream>>>(d, w->fptr_d, num_fields, num_layers, zint, zmid, crm_dt / time_scale);
  w->launches++;
  PCK(cudaGetLastError());
  return 0;
}

int pamm_compute_gcm_forcing_tendencies(pamm_workspace *w, const pamm_crm_fields *crm, const pamm_gcm_columns *gcm,
                                        const pamm_forcing_tend *tend, double gcm_physics_dt) {
  if (!(gcm_physics_dt > 0.0)) { w->err = "pamm_compute_gcm_forcing_tendencies: gcm_physics_dt must be positive"; return 1; }
  const Dims d{w->nz, w->ny, w->nx, w->nens};
  const dim3 grid((unsigned)(w->nens + 127) / 128, (unsigned)w->nz);
  k_gcm_forcing_compute<<<grid, 128, 0, w->stream>>>(d, *crm, *gcm, *tend, 1.0 / gcm_physics_dt);
  w->launches++;
  PCK(cudaGetLastError());
  return 0;
}

int pamm_apply_gcm_forcing_tendencies(pamm_workspace *w, const pamm_crm_fields *crm, const pamm_gcm_columns *gcm,
                                      const pamm_forcing_tend *tend, const double *dz, double crm_dt, double gcm_physics_dt) {
  if (!(gcm_physics_dt > 0.0)) { w->err = "pamm_apply_gcm_forcing_tendencies: gcm_physics_dt must be positive"; return 1; }
  const Dims d{w->nz, w->ny, w->nx, w->nens};
  const dim3 grid((unsigned)(w->nens + 127) / 128, (unsigned)w->nz);
  k_gcm_forcing_apply<<<grid, 128, 0, w->stream>>>(d, *crm, *gcm, *tend, dz, crm_dt, 1.0 / gcm_physics_dt, w->scratch);
  k_fill_holes_global<<<dim3((unsigned)(w->nens + 127) / 128, 3), 128, 0, w->stream>>>(d, crm->water_vapor, crm->cloud_water, crm->ice, dz,
                                                                                       w->scratch);
  w->launches += 2;
  PCK(cudaGetLastError());
  return 0;
}

int pamm_broadcast_initial_gcm_column(pamm_workspace *w, const double *const *gcm_columns, double *const *fields) {
  BroadcastArgs a;
  for (int f = 0; f < 6; f++) { a.col[f] = gcm_columns[f]; a.fld[f] = fields[f]; }
  const Dims d{w->nz, w->ny, w->nx, w->nens};
  k_broadcast_column<<<dim3((unsigned)(w->nens + 127) / 128, (unsigned)w->nz, 6), 128, 0, w->stream>>>(d, a);
  w->launches++;
  PCK(cudaGetLastError());
  return 0;
}

int pamm_perturb_temperature(pamm_workspace *w, double *temp, const int *id, double magnitude) {
  const int num_levels = w->nz / 4;
  if (num_levels < 1) return 0;
  const Dims d{w->nz, w->ny, w->nx, w->nens};
  const dim3 grid((unsigned)(w->nens + 127) / 128, (unsigned)num_levels);
  k_perturb_temperature<<<grid, 128, 0, w->stream>>>(d, temp, id, num_levels, magnitude);
  w->launches++;
  PCK(cudaGetLastError());
  return 0;
}

}  // extern "C"
