// Coupler modules either side of the dycore in PAM's step loop, as sm_100a CUDA kernels behind the C ABI of
// include/pam_modules_b200.h (part of libawfl_b200.so). Reference behaviour: pam_core/modules/sponge_layer.h,
// gcm_forcing.h, perturb_temperature.h (cited per kernel).
//
// Layout: ensemble member fastest, so one warp = 32 consecutive members of one (level[, field]) column set and every
// global access is a coalesced 256-byte line. All of these modules reduce over the horizontal (j,i) of one level and
// member: a block is 32 members x S segments of the column's ny*nx cells, each thread walks its run of cells in (j,i)
// order and the S partial sums are combined in segment order (seg_combine): deterministic, no atomics -- the reference's
// GPU backend uses atomicAdd, i.e. no defined order. With S = 1 (2-D CRMs) this is the summation order of the
// reference's serial backend. They are HBM-bound streaming kernels; there is NO CPU fallback.
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

// Work decomposition of the reducing kernels. A block is 32 consecutive members x S horizontal segments (threadIdx.x =
// member lane: coalesced; threadIdx.y = segment): a thread walks a contiguous run of the ny*nx cells of its (level, member)
// column in (j,i) order, the S partial sums of a column are combined in segment order by every thread of the column
// (deterministic, no atomics). S = 1 (2-D CRMs: a column has 64 cells) is the summation order of the reference's serial
// backend, bit for bit; a 3-D CRM (1024 cells per column, and perhaps only 128 members per GPU) gets S = 16, i.e. 16 x the
// threads -- with one thread per column the C3 shard ran these kernels at 3-4 % of the HBM rate.
constexpr int SEG_MAX = 16;
static int seg_count(int nh) {
  int s = nh / 64;
  return s < 1 ? 1 : (s > SEG_MAX ? SEG_MAX : s);
}
struct SegItem { int e, c0, c1; bool active; };
__device__ __forceinline__ SegItem seg_item(const Dims &d) {
  SegItem it;
  const int e = blockIdx.x * 32 + threadIdx.x, nh = d.ny * d.nx, S = blockDim.y, sg = threadIdx.y;
  it.active = e < d.nens;
  it.e = it.active ? e : d.nens - 1;       // lanes past the last member shadow it (they take part in the barriers, write nothing)
  it.c0 = (int)((long)nh * sg / S);
  it.c1 = (int)((long)nh * (sg + 1) / S);
  return it;
}
// v[0..NV) -> the sums over the S segments of this lane's column, in segment order, in every thread of the column
template <int NV>
__device__ __forceinline__ void seg_combine(double (&v)[NV], double *sm) {
  const int S = blockDim.y;
  if (S == 1) return;
  __syncthreads();                            // sm may still be read from a previous combine
#pragma unroll
  for (int q = 0; q < NV; q++) sm[(q * SEG_MAX + threadIdx.y) * 32 + threadIdx.x] = v[q];
  __syncthreads();
#pragma unroll
  for (int q = 0; q < NV; q++) {
    double t = sm[(q * SEG_MAX) * 32 + threadIdx.x];
    for (int sg = 1; sg < S; sg++) t += sm[(q * SEG_MAX + sg) * 32 + threadIdx.x];
    v[q] = t;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// modules::sponge_layer (sponge_layer.h:9-96). One thread = (field, sponge level, member): horizontal mean in (j,i)
// order (:71-74; zero for the w field, :32), then the relaxation of every cell of that level (:86-92).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32 * SEG_MAX) k_sponge(const Dims d, double *const *fields, int num_fields, int num_layers,
                                                         const double *zint, const double *zmid, double time_factor) {
  __shared__ double sm[SEG_MAX * 32];
  const SegItem it = seg_item(d);
  const int e = it.e, kloc = blockIdx.y, ifld = blockIdx.z;
  const int k = d.nz - 1 - kloc;
  double *f = fields[ifld] + (long)k * d.ny * d.nx * d.nens + e;
  const double r_nx_ny = 1.0 / (d.nx * d.ny);
  double havg[1] = {0.0};
  if (ifld != 3) {
    for (int c = it.c0; c < it.c1; c++) havg[0] += f[(long)c * d.nens] * r_nx_ny;
  }
  seg_combine<1>(havg, sm);
  const double ztop = zint[(long)d.nz * d.nens + e];
  const double rel_dist = (ztop - zmid[(long)k * d.nens + e]) / (ztop - zmid[(long)(d.nz - 1 - (num_layers - 1)) * d.nens + e]);
  const double space_factor = (cos(M_PI * rel_dist) + 1) / 2;
  const double factor = space_factor * time_factor;
  if (!it.active) return;
  for (int c = it.c0; c < it.c1; c++) {
    const double x = f[(long)c * d.nens];
    f[(long)c * d.nens] = x + (havg[0] - x) * factor;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// modules::compute_gcm_forcing_tendencies (gcm_forcing.h:18-205). One thread = (level, member): the ten column
// means (:103-131) and the fourteen tendencies (:171-203).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32 * SEG_MAX) k_gcm_forcing_compute(const Dims d, const pamm_crm_fields c, const pamm_gcm_columns g,
                                                                      const pamm_forcing_tend t, double r_dt_gcm) {
  __shared__ double sm[10 * SEG_MAX * 32];
  const SegItem it = seg_item(d);
  const int e = it.e, k = blockIdx.y;
  const long base = (long)k * d.ny * d.nx * d.nens + e;
  const double r_nx_ny = 1.0 / (d.nx * d.ny);
  double a[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};      // rd, u, v, t, qv, ql, qi, nc, ni, nr
  for (int cc = it.c0; cc < it.c1; cc++) {
    const long i = base + (long)cc * d.nens;
    const double rd = c.density_dry[i], rv = c.water_vapor[i];
    a[0] += rd * r_nx_ny;
    a[1] += c.uvel[i] * r_nx_ny;
    a[2] += c.vvel[i] * r_nx_ny;
    a[3] += c.temp[i] * r_nx_ny;
    a[4] += (rv / (rd + rv)) * r_nx_ny;
    a[5] += (c.cloud_water[i] / (rd + rv)) * r_nx_ny;
    a[6] += (c.ice[i] / (rd + rv)) * r_nx_ny;
    a[7] += c.cloud_water_num[i] * r_nx_ny;
    a[8] += c.ice_num[i] * r_nx_ny;
    a[9] += c.rain_num[i] * r_nx_ny;
  }
  seg_combine<10>(a, sm);
  if (!it.active || threadIdx.y != 0) return;
  const long ke = (long)k * d.nens + e;
  const double grd = g.gcm_density_dry[ke], grv = g.gcm_water_vapor[ke];
  t.rho_d[ke] = (grd - a[0]) * r_dt_gcm;
  t.uvel[ke] = (g.gcm_uvel[ke] - a[1]) * r_dt_gcm;
  t.vvel[ke] = (g.gcm_vvel[ke] - a[2]) * r_dt_gcm;
  t.temp[ke] = (g.gcm_temp[ke] - a[3]) * r_dt_gcm;
  const double fqv = (grv / (grd + grv) - a[4]) * r_dt_gcm;
  const double fql = (g.gcm_cloud_water[ke] / (grd + grv) - a[5]) * r_dt_gcm;
  const double fqi = (g.gcm_cloud_ice[ke] / (grd + grv) - a[6]) * r_dt_gcm;
  t.qv[ke] = fqv; t.ql[ke] = fql; t.qi[ke] = fqi;
  t.nc[ke] = (g.gcm_num_liq[ke] - a[7]) * r_dt_gcm;
  t.ni[ke] = (g.gcm_num_ice[ke] - a[8]) * r_dt_gcm;
  t.nr[ke] = (g.gcm_num_rain[ke] - a[9]) * r_dt_gcm;
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
__global__ void __launch_bounds__(32 * SEG_MAX) k_gcm_forcing_apply(const Dims d, const pamm_crm_fields c, const pamm_gcm_columns g,
                                                                    const pamm_forcing_tend t, const double *dz, double dt, double r_dt_gcm,
                                                                    double *scratch) {
  __shared__ double sm[6 * SEG_MAX * 32];
  const SegItem it = seg_item(d);
  const int e = it.e, k = blockIdx.y;
  const long base = (long)k * d.ny * d.nx * d.nens + e;
  const long ke = (long)k * d.nens + e;
  const double r_nx_ny = 1.0 / (d.nx * d.ny);
  const double f_rd = t.rho_d[ke], f_u = t.uvel[ke], f_v = t.vvel[ke], f_t = t.temp[ke];
  const double f_qv = t.qv[ke], f_ql = t.ql[ke], f_qi = t.qi[ke], f_nc = t.nc[ke], f_ni = t.ni[ke], f_nr = t.nr[ke];
  const double dzk = dz[ke];
  double a[6] = {0, 0, 0, 0, 0, 0};       // avg_v, avg_l, avg_i, neg_v, neg_l, neg_i
  for (int cc = it.c0; cc < it.c1; cc++) {
    const long i = base + (long)cc * d.nens;
    const double rho_d_old = c.density_dry[i];
    const double rd = rho_d_old + f_rd * dt;
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
    a[0] += rv * r_nx_ny; a[1] += rl * r_nx_ny; a[2] += ri * r_nx_ny;
    if (rv < 0) { a[3] += -rv * dzk; rv = 0; }
    if (rl < 0) { a[4] += -rl * dzk; rl = 0; }
    if (ri < 0) { a[5] += -ri * dzk; ri = 0; }
    if (it.active) {
      c.density_dry[i] = rd;
      c.uvel[i] += f_u * dt;
      c.vvel[i] += f_v * dt;
      c.temp[i] += f_t * dt;
      c.cloud_water_num[i] = xc; c.ice_num[i] = xi; c.rain_num[i] = xr;
      c.water_vapor[i] = rv; c.cloud_water[i] = rl; c.ice[i] = ri;
    }
  }
  seg_combine<6>(a, sm);
  const bool lead = it.active && threadIdx.y == 0;
  if (lead) {
    t.rho_v[ke] = (g.gcm_water_vapor[ke] - a[0]) * r_dt_gcm;
    t.rho_l[ke] = (g.gcm_cloud_water[ke] - a[1]) * r_dt_gcm;
    t.rho_i[ke] = (g.gcm_cloud_ice[ke] - a[2]) * r_dt_gcm;
  }
  // level stage of fill_holes (gcm_forcing.h:208-255) for the species that went negative somewhere on this (level, member):
  // the positive mass available on the level, then every cell gives up its share. Whether ANY column of the block needs it
  // is a block-wide decision (the barriers inside seg_combine); a column that does not need it only lends its threads.
  double *const xs[3] = {c.water_vapor, c.cloud_water, c.ice};
  double pos[3] = {0, 0, 0};
  const bool need_any = a[3] > 0 || a[4] > 0 || a[5] > 0;
  if (__syncthreads_or(need_any ? 1 : 0)) {
#pragma unroll
    for (int sp = 0; sp < 3; sp++) {
      if (!(a[3 + sp] > 0)) continue;
      for (int cc = it.c0; cc < it.c1; cc++) {
        const double v = xs[sp][base + (long)cc * d.nens];
        if (v > 0) pos[sp] += v * dzk;
      }
    }
    seg_combine<3>(pos, sm);
    if (it.active) {
#pragma unroll
      for (int sp = 0; sp < 3; sp++) {
        if (!(a[3 + sp] > 0) || !(pos[sp] > 0)) continue;
        for (int cc = it.c0; cc < it.c1; cc++) {
          const long i = base + (long)cc * d.nens;
          const double v = xs[sp][i];
          const double factor = v * dzk / pos[sp];
          xs[sp][i] = fmax(0.0, v - (a[3 + sp] * factor) / dzk);
        }
      }
    }
  }
  if (lead) {
    const long nke = (long)d.nz * d.nens;
    scratch[0 * nke + ke] = a[3]; scratch[1 * nke + ke] = pos[0];
    scratch[2 * nke + ke] = a[4]; scratch[3 * nke + ke] = pos[1];
    scratch[4 * nke + ke] = a[5]; scratch[5 * nke + ke] = pos[2];
  }
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
  const dim3 grid((unsigned)(w->nens + 31) / 32, (unsigned)num_layers, (unsigned)num_fields);
  k_sponge<<<grid, dim3(32, seg_count(w->ny * w->nx)), 0, w->stream>>>(d, w->fptr_d, num_fields, num_layers, zint, zmid, crm_dt / time_scale);
  w->launches++;
  PCK(cudaGetLastError());
  return 0;
}

int pamm_compute_gcm_forcing_tendencies(pamm_workspace *w, const pamm_crm_fields *crm, const pamm_gcm_columns *gcm,
                                        const pamm_forcing_tend *tend, double gcm_physics_dt) {
  if (!(gcm_physics_dt > 0.0)) { w->err = "pamm_compute_gcm_forcing_tendencies: gcm_physics_dt must be positive"; return 1; }
  const Dims d{w->nz, w->ny, w->nx, w->nens};
  const dim3 grid((unsigned)(w->nens + 31) / 32, (unsigned)w->nz);
  k_gcm_forcing_compute<<<grid, dim3(32, seg_count(w->ny * w->nx)), 0, w->stream>>>(d, *crm, *gcm, *tend, 1.0 / gcm_physics_dt);
  w->launches++;
  PCK(cudaGetLastError());
  return 0;
}

int pamm_apply_gcm_forcing_tendencies(pamm_workspace *w, const pamm_crm_fields *crm, const pamm_gcm_columns *gcm,
                                      const pamm_forcing_tend *tend, const double *dz, double crm_dt, double gcm_physics_dt) {
  if (!(gcm_physics_dt > 0.0)) { w->err = "pamm_apply_gcm_forcing_tendencies: gcm_physics_dt must be positive"; return 1; }
  const Dims d{w->nz, w->ny, w->nx, w->nens};
  const dim3 grid((unsigned)(w->nens + 31) / 32, (unsigned)w->nz);
  k_gcm_forcing_apply<<<grid, dim3(32, seg_count(w->ny * w->nx)), 0, w->stream>>>(d, *crm, *gcm, *tend, dz, crm_dt, 1.0 / gcm_physics_dt, w->scratch);
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
