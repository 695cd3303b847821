// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// C-ABI harness around the UNMODIFIED reference AWFL dycore, compiled in place from the sources under
// /root/reference (never copied into this repo) by oracle/ref_build/Makefile, output into oracle/_ref/.
// It is used to (1) pin the C restatement in oracle/awfl_oracle.c, (2) generate tests/golden/*.npz,
// (3) serve as the timed CPU baseline ("kind": "reference") in bench.py.
//
// What it wraps (reference file:line):
//   pam::PamCoupler::allocate_coupler_state / set_grid / add_tracer   pam_core/pam_coupler.h:184-356
//   Dycore::init                                                      dynamics/awfl/Dycore.h:835
//   Dycore::timeStep            (raw, informational: SURVEY F2)       dynamics/awfl/Dycore.h:107
//   Dycore::compute_tendencies  (oracle of record, F2 neutralised)    dynamics/awfl/Dycore.h:262
//   Dycore::halo_exchange / reconstruct / convert_* / compute_time_step
//   TransformMatrices::* / weno::wenoSetIdealSigma constant tables
//
// "Intended semantics" (SURVEY.md F2): the reference's bottom boundary condition reads ghost theta written
// by another thread (Dycore.h:692 vs :670); in serial order that read is of stale memory. The harness owns
// the haloed arrays and pre-fills every bottom ghost theta with the value the racing thread would write
// (theta of the first interior cell) before each compute_tendencies call, which makes the reference
// deterministic and equal to what the code intends.

#include "pam_coupler.h"
#include "Dycore.h"
#include "sponge_layer.h"
#include "gcm_forcing.h"
#include "perturb_temperature.h"
#include <cstring>
#include <sstream>
#include <chrono>

namespace pam { std::mutex data_manager_mutex; }   // pam_core/pam_coupler_globals.cpp:4-6

namespace {

int constexpr ORD = Dycore::ord;
int constexpr HS  = Dycore::hs;
int constexpr idR = Dycore::idR, idU = Dycore::idU, idV = Dycore::idV, idW = Dycore::idW, idT = Dycore::idT;

struct Handle {
  pam::PamCoupler coupler;
  Dycore          dycore;
  int nz, ny, nx, nens, nt;
  std::vector<std::string> tracer_names;
  std::string err;
};

struct CoutSilencer {
  std::streambuf *old;
  std::ostringstream sink;
  CoutSilencer()  { old = std::cout.rdbuf(sink.rdbuf()); }
  ~CoutSilencer() { std::cout.rdbuf(old); }
};

void ensure_yakl() {
  if (!yakl::isInitialized()) {
    CoutSilencer s;
    size_t mb = 1024;
    if (const char *e = std::getenv("PAM_REF_POOL_MB")) mb = (size_t) atol(e);
    yakl::init(yakl::InitConfig().set_pool_initial_mb(mb).set_pool_grow_mb(mb));
  }
}

// Pre-fill the bottom ghost theta. Must be called when idT holds rho*theta (before the
// reference's in-place divide), so the quotient is bit-identical to what Dycore.h:319 + :670 produce.
void prefill_bottom_ghost_theta(Handle &h, real5d const &state) {
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens;
  for (int kk=0; kk<HS; kk++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++) {
    state(idT,kk,HS+j,HS+i,e) = state(idT,HS,HS+j,HS+i,e) / state(idR,HS,HS+j,HS+i,e);
  }
}

template <class A> void fill_from(A &a, const double *src) { std::memcpy(a.data(), src, a.size()*sizeof(double)); }
template <class A> void copy_to  (A const &a, double *dst) { std::memcpy(dst, a.data(), a.size()*sizeof(double)); }

void tend_intended(Handle &h, real5d const &state, real5d const &state_tend,
                   real5d const &tracers, real5d const &tracers_tend, real dt) {
  prefill_bottom_ghost_theta(h, state);
  h.dycore.compute_tendencies(h.coupler, state, state_tend, tracers, tracers_tend, dt);
}

} // namespace


extern "C" {

int ref_ord() { return ORD; }

// Finalise YAKL quietly: its pool destructor otherwise prints high-water statistics to stdout at process exit,
// which would follow bench.py's single JSON line.
void ref_shutdown() {
  if (yakl::isInitialized()) {
    CoutSilencer s;
    yakl::finalize();
  }
}

int ref_openmp() {
  #ifdef YAKL_ARCH_OPENMP
    return 1;
  #else
    return 0;
  #endif
}

// zint: (nz+1) values if zint_per_ens==0, else (nz+1)*nens (k-major, iens fastest).
// tracer_names: num_tracers C strings; positive/adds_mass: num_tracers ints.
// consts: NULL for the dycore's defaults, else {R_d, cp_d, R_v, cp_v, p0, grav}.
void *ref_create(int nz, int ny, int nx, int nens, double xlen, double ylen,
                 const double *zint, int zint_per_ens,
                 int num_tracers, const char **tracer_names, const int *positive, const int *adds_mass,
                 const double *consts, double crm_dt) {
  ensure_yakl();
  Handle *h = new Handle;
  try {
    h->nz=nz; h->ny=ny; h->nx=nx; h->nens=nens; h->nt=num_tracers;
    auto &c = h->coupler;
    c.set_option<real>("crm_dt", crm_dt);
    if (consts) {
      c.set_option<real>("R_d" , consts[0]);
      c.set_option<real>("cp_d", consts[1]);
      c.set_option<real>("R_v" , consts[2]);
      c.set_option<real>("cp_v", consts[3]);
      c.set_option<real>("p0"  , consts[4]);
      c.set_option<real>("grav", consts[5]);
    }
    c.allocate_coupler_state(nz, ny, nx, nens);
    real2d z2("zint_in", nz+1, nens);
    for (int k=0;k<=nz;k++) for (int e=0;e<nens;e++) z2(k,e) = zint_per_ens ? zint[k*nens+e] : zint[k];
    c.set_grid(xlen, ylen, realConst2d(z2));
    for (int t=0;t<num_tracers;t++) {
      c.add_tracer(tracer_names[t], tracer_names[t], positive[t]!=0, adds_mass[t]!=0);
      h->tracer_names.push_back(tracer_names[t]);
      // add_tracer does not zero the field
      auto f = c.get_data_manager_device_readwrite().get<real,4>(tracer_names[t]);
      f = 0;
    }
    {
      CoutSilencer s;   // init prints a leftover debug solve (Dycore.h:851-864)
      h->dycore.init(c);
    }
    // variable_gravity is allocated but never filled by init
    auto g = c.get_data_manager_device_readwrite().get<real,2>("variable_gravity");
    g = 0;
  } catch (std::string &s) { std::cerr << "ref_create: " << s << std::endl; delete h; return nullptr; }
    catch (...)            { std::cerr << "ref_create: exception" << std::endl; delete h; return nullptr; }
  return h;
}

void ref_destroy(void *hv) { delete (Handle*) hv; }

void ref_set_crm_dt(void *hv, double dt) { ((Handle*)hv)->coupler.set_option<real>("crm_dt", dt); }

double ref_get_option(void *hv, const char *name) { return ((Handle*)hv)->coupler.get_option<real>(name); }
int    ref_get_idWV  (void *hv) { return ((Handle*)hv)->coupler.get_option<int>("idWV"); }

// Generic access to any real-typed DataManager entry (flattened, C order).
long ref_field_size(void *hv, const char *name) {
  auto &dm = ((Handle*)hv)->coupler.get_data_manager_device_readwrite();
  return (long) dm.get_collapsed<real>(name).size();
}
int ref_set_field(void *hv, const char *name, const double *src, long n) {
  auto &dm = ((Handle*)hv)->coupler.get_data_manager_device_readwrite();
  auto a = dm.get_collapsed<real>(name);
  if ((long)a.size() != n) return 1;
  std::memcpy(a.data(), src, n*sizeof(double));
  return 0;
}
int ref_get_field(void *hv, const char *name, double *dst, long n) {
  auto &dm = ((Handle*)hv)->coupler.get_data_manager_device_readwrite();
  auto a = dm.get_collapsed<real>(name);
  if ((long)a.size() != n) return 1;
  std::memcpy(dst, a.data(), n*sizeof(double));
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// Constant tables (dynamics/awfl/TransformMatrices.h, WenoLimiter.h:32-95) for table-level tests.
// ---------------------------------------------------------------------------------------------------
void ref_get_tables(double *sten_to_coefs /*ORD*ORD*/, double *coefs_to_gll /*ORD*2*/,
                    double *weno_recon_lower /*HS^3*/, double *idl /*HS+1*/, double *sigma) {
  SArray<real,2,ORD,ORD> s2c;   TransformMatrices::sten_to_coefs(s2c);
  SArray<real,2,ORD,2>   c2g;   TransformMatrices::coefs_to_gll_lower(c2g);
  SArray<real,3,HS,HS,HS> wrl;  TransformMatrices::weno_lower_sten_to_coefs(wrl);
  SArray<real,1,HS+1> idl_s; real sig;
  weno::wenoSetIdealSigma<ORD>(idl_s, sig);
  for (int i=0;i<ORD;i++) for (int j=0;j<ORD;j++) sten_to_coefs[i*ORD+j] = s2c(i,j);
  for (int i=0;i<ORD;i++) for (int j=0;j<2;j++)   coefs_to_gll[i*2+j]    = c2g(i,j);
  for (int i=0;i<HS;i++) for (int j=0;j<HS;j++) for (int k=0;k<HS;k++) weno_recon_lower[(i*HS+j)*HS+k] = wrl(i,j,k);
  for (int i=0;i<HS+1;i++) idl[i] = idl_s(i);
  *sigma = sig;
}

// coefs_to_tv for n = HS (low-order candidates) or n = ORD (high-order)
double ref_coefs_to_tv(int n, const double *a) {
  if (n == HS)  { SArray<real,1,HS>  s; for (int i=0;i<HS;i++)  s(i)=a[i]; return TransformMatrices::coefs_to_tv(s); }
  if (n == ORD) { SArray<real,1,ORD> s; for (int i=0;i<ORD;i++) s(i)=a[i]; return TransformMatrices::coefs_to_tv(s); }
  return -1;
}

// Dycore::reconstruct with the constant (horizontal) tables, n stencils of ORD values each.
void ref_reconstruct(const double *stencils, int n, int ind, double *out) {
  SArray<real,2,ORD,ORD> s2c;   TransformMatrices::sten_to_coefs(s2c);
  SArray<real,2,ORD,2>   c2g;   TransformMatrices::coefs_to_gll_lower(c2g);
  SArray<real,3,HS,HS,HS> wrl;  TransformMatrices::weno_lower_sten_to_coefs(wrl);
  SArray<real,1,HS+1> idl; real sigma;
  weno::wenoSetIdealSigma<ORD>(idl, sigma);
  for (int q=0;q<n;q++) {
    SArray<real,1,ORD> st; for (int s=0;s<ORD;s++) st(s) = stencils[q*ORD+s];
    out[q] = Dycore::reconstruct(st, c2g, s2c, wrl, idl, sigma, ind);
  }
}

// Same with the per-level vertical matrices of (level, iens) from the DataManager.
void ref_reconstruct_vert(void *hv, int level, int iens, const double *stencils, int n, int ind, double *out) {
  Handle &h = *(Handle*)hv;
  auto &dm = h.coupler.get_data_manager_device_readwrite();
  auto vwrl = dm.get<real const,5>("vert_weno_recon_lower");
  auto vs2c = dm.get<real const,4>("vert_sten_to_coefs");
  SArray<real,2,ORD,ORD> s2c;  SArray<real,3,HS,HS,HS> wrl;
  for (int a=0;a<ORD;a++) for (int b=0;b<ORD;b++) s2c(a,b) = vs2c(level,a,b,iens);
  for (int a=0;a<HS;a++) for (int b=0;b<HS;b++) for (int c=0;c<HS;c++) wrl(a,b,c) = vwrl(level,a,b,c,iens);
  SArray<real,2,ORD,2>   c2g;   TransformMatrices::coefs_to_gll_lower(c2g);
  SArray<real,1,HS+1> idl; real sigma;
  weno::wenoSetIdealSigma<ORD>(idl, sigma);
  for (int q=0;q<n;q++) {
    SArray<real,1,ORD> st; for (int s=0;s<ORD;s++) st(s) = stencils[q*ORD+s];
    out[q] = Dycore::reconstruct(st, c2g, s2c, wrl, idl, sigma, ind);
  }
}

// ---------------------------------------------------------------------------------------------------
// Dycore entry points
// ---------------------------------------------------------------------------------------------------
double ref_compute_time_step(void *hv, double cfl) {
  Handle &h = *(Handle*)hv;
  return h.dycore.compute_time_step(h.coupler, cfl);
}

// Unmodified Dycore::timeStep (informational only; depends on stale pool memory).
int ref_timestep_raw(void *hv) {
  Handle &h = *(Handle*)hv;
  try { h.dycore.timeStep(h.coupler); } catch (...) { return 1; }
  return 0;
}

// Unmodified declare_current_profile_as_hydrostatic (informational; same F2 dependence).
int ref_declare_hydrostatic_raw(void *hv, int use_gcm) {
  Handle &h = *(Handle*)hv;
  try { h.dycore.declare_current_profile_as_hydrostatic(h.coupler, use_gcm!=0); } catch (...) { return 1; }
  return 0;
}

// One semi-discrete RHS with harness-owned arrays (oracle of record for tendency-level parity).
//   state_in   : 5 x (nz,ny,nx,nens) un-haloed conserved state (rho, rho u, rho v, rho w, rho theta)
//   tracers_in : nt x (nz,ny,nx,nens)
//   fct_start  : nt x (nz,ny,nx,nens) FCT starting mass (what timeStep stashes in tracers_tend)
// outputs: state_tend (5 x ...), tracers_tend (nt x ...), and the round-tripped state/tracers
// ((x/rho)*rho, Dycore.h:316-320,527-532) which the RK update consumes.
int ref_compute_tendencies(void *hv, const double *state_in, const double *tracers_in, const double *fct_start,
                           double dt, double *state_tend_out, double *tracers_tend_out,
                           double *state_rt_out, double *tracers_rt_out) {
  Handle &h = *(Handle*)hv;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real5d state_tend  ("state_tend"  ,5 ,nz,ny,nx,nens);
  real5d tracers_tend("tracers_tend",nt,nz,ny,nx,nens);
  state = 0; if (nt>0) tracers = 0;
  long n4 = (long)nz*ny*nx*nens;
  for (int l=0;l<5;l++)  for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++)
    state(l,HS+k,HS+j,HS+i,e) = state_in[l*n4 + (((long)k*ny+j)*nx+i)*nens+e];
  for (int l=0;l<nt;l++) for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++) {
    long o = l*n4 + (((long)k*ny+j)*nx+i)*nens+e;
    tracers(l,HS+k,HS+j,HS+i,e) = tracers_in[o];
    tracers_tend(l,k,j,i,e)     = fct_start[o];
  }
  try { tend_intended(h, state, state_tend, tracers, tracers_tend, dt); } catch (...) { return 1; }
  std::memcpy(state_tend_out, state_tend.data(), 5*n4*sizeof(double));
  if (nt>0) std::memcpy(tracers_tend_out, tracers_tend.data(), nt*n4*sizeof(double));
  for (int l=0;l<5;l++)  for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++)
    if (state_rt_out) state_rt_out[l*n4 + (((long)k*ny+j)*nx+i)*nens+e] = state(l,HS+k,HS+j,HS+i,e);
  for (int l=0;l<nt;l++) for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++)
    if (tracers_rt_out) tracers_rt_out[l*n4 + (((long)k*ny+j)*nx+i)*nens+e] = tracers(l,HS+k,HS+j,HS+i,e);
  return 0;
}

// Dycore::timeStep with the reference's own conversion, dt and tendency routines, the SSPRK3 staging
// of Dycore.h:147-222 restated here so the harness owns the haloed arrays (F2 neutralised).
// ncycles_override > 0 fixes the number of sub-cycles; returns the ncycles used (or -1 on error).
int ref_timestep_intended(void *hv, int ncycles_override) {
  Handle &h = *(Handle*)hv;
  auto &c = h.coupler;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  try {
    auto positive = c.get_data_manager_device_readonly().get<bool const,1>("tracer_positive");
    real dt_phys = c.get_option<real>("crm_dt");
    real5d q ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d tr("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    q = 0; if (nt>0) tr = 0;
    h.dycore.convert_coupler_to_dynamics(c, q, tr);
    for (int l=0;l<nt;l++) if (positive(l))
      for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++)
        tr(l,HS+k,HS+j,HS+i,e) = std::max(0., tr(l,HS+k,HS+j,HS+i,e));
    real dt_dyn = h.dycore.compute_time_step(c);
    int ncycles = (int) std::ceil(dt_phys / dt_dyn);
    if (ncycles_override > 0) ncycles = ncycles_override;
    dt_dyn = dt_phys / ncycles;
    real5d q1 ("state_tmp"   ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d t1 ("tracers_tmp" ,nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d dq ("state_tend"  ,5 ,nz,ny,nx,nens);
    real5d dt_("tracers_tend",nt,nz,ny,nx,nens);
    q1 = 0; if (nt>0) t1 = 0;
    // stage coefficients: new = a*base + b*stage + b*dt*tend ; next FCT start = fa*base + fb*new
    for (int icycle=0; icycle<ncycles; icycle++) {
      for (int l=0;l<nt;l++) for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++)
        dt_(l,k,j,i,e) = tr(l,HS+k,HS+j,HS+i,e);
      // ---- stage 1 ----
      tend_intended(h, q, dq, tr, dt_, dt_dyn);
      for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++) {
        for (int l=0;l<5;l++) q1(l,HS+k,HS+j,HS+i,e) = q(l,HS+k,HS+j,HS+i,e) + dt_dyn * dq(l,k,j,i,e);
        for (int l=0;l<nt;l++) {
          real v = tr(l,HS+k,HS+j,HS+i,e) + dt_dyn * dt_(l,k,j,i,e);
          if (positive(l)) v = std::max(0., v);
          t1(l,HS+k,HS+j,HS+i,e) = v;
          dt_(l,k,j,i,e) = (3._fp/4._fp) * tr(l,HS+k,HS+j,HS+i,e) + (1._fp/4._fp) * v;
        }
      }
      // ---- stage 2 ----
      tend_intended(h, q1, dq, t1, dt_, (1._fp/4._fp) * dt_dyn);
      for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++) {
        for (int l=0;l<5;l++) q1(l,HS+k,HS+j,HS+i,e) = (3._fp/4._fp) * q (l,HS+k,HS+j,HS+i,e) +
                                                       (1._fp/4._fp) * q1(l,HS+k,HS+j,HS+i,e) +
                                                       (1._fp/4._fp) * dt_dyn * dq(l,k,j,i,e);
        for (int l=0;l<nt;l++) {
          real v = (3._fp/4._fp) * tr(l,HS+k,HS+j,HS+i,e) + (1._fp/4._fp) * t1(l,HS+k,HS+j,HS+i,e) +
                   (1._fp/4._fp) * dt_dyn * dt_(l,k,j,i,e);
          if (positive(l)) v = std::max(0., v);
          t1(l,HS+k,HS+j,HS+i,e) = v;
          dt_(l,k,j,i,e) = (1._fp/3._fp) * tr(l,HS+k,HS+j,HS+i,e) + (2._fp/3._fp) * v;
        }
      }
      // ---- stage 3 ----
      tend_intended(h, q1, dq, t1, dt_, (2._fp/3._fp) * dt_dyn);
      for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++) {
        for (int l=0;l<5;l++) q(l,HS+k,HS+j,HS+i,e) = (1._fp/3._fp) * q (l,HS+k,HS+j,HS+i,e) +
                                                      (2._fp/3._fp) * q1(l,HS+k,HS+j,HS+i,e) +
                                                      (2._fp/3._fp) * dt_dyn * dq(l,k,j,i,e);
        for (int l=0;l<nt;l++) {
          real v = (1._fp/3._fp) * tr(l,HS+k,HS+j,HS+i,e) + (2._fp/3._fp) * t1(l,HS+k,HS+j,HS+i,e) +
                   (2._fp/3._fp) * dt_dyn * dt_(l,k,j,i,e);
          if (positive(l)) v = std::max(0., v);
          tr(l,HS+k,HS+j,HS+i,e) = v;
        }
      }
    }
    h.dycore.convert_dynamics_to_coupler(c, q, tr);
    return ncycles;
  } catch (std::string &s) { std::cerr << "ref_timestep_intended: " << s << std::endl; return -1; }
    catch (...) { return -1; }
}

// declare_current_profile_as_hydrostatic (Dycore.h:1392-1504, grav_balance branch) driven through the
// reference's own convert / halo_exchange / reconstruct, with the bottom ghost theta pre-filled (F2)
// and the column mean accumulated in (j,i) order instead of atomics.
int ref_declare_hydrostatic_intended(void *hv, int use_gcm) {
  Handle &h = *(Handle*)hv;
  auto &c = h.coupler;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  try {
    auto &dm = c.get_data_manager_device_readwrite();
    real gamma = c.get_option<real>("gamma_d"), C0 = c.get_option<real>("C0");
    real R_d = c.get_option<real>("R_d"), R_v = c.get_option<real>("R_v");
    auto dz   = dm.get<real const,2>("vertical_cell_dz");
    auto vwrl = dm.get<real const,5>("vert_weno_recon_lower");
    auto vs2c = dm.get<real const,4>("vert_sten_to_coefs");
    auto grav_var = dm.get<real,2>("variable_gravity");
    real5d q ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d tr("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    q = 0; if (nt>0) tr = 0;
    if (use_gcm) {
      auto g_rd = dm.get<real const,2>("gcm_density_dry"); auto g_t  = dm.get<real const,2>("gcm_temp");
      auto g_rv = dm.get<real const,2>("gcm_water_vapor"); auto g_rc = dm.get<real const,2>("gcm_cloud_water");
      auto g_ri = dm.get<real const,2>("gcm_cloud_ice");
      for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++) {
        real rho = g_rd(k,e) + g_rv(k,e) + g_rc(k,e) + g_ri(k,e);
        real p   = (g_rd(k,e)*R_d + g_rv(k,e)*R_v)*g_t(k,e);
        q(idR,HS+k,HS+j,HS+i,e) = rho;
        q(idT,HS+k,HS+j,HS+i,e) = std::pow( p/C0 , 1._fp/gamma );
      }
    } else {
      h.dycore.convert_coupler_to_dynamics(c, q, tr);
    }
    real4d pressure("pressure",nz+2*HS,ny+2*HS,nx+2*HS,nens);
    pressure = 0;
    prefill_bottom_ghost_theta(h, q);
    for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++) {
      pressure(HS+k,HS+j,HS+i,e) = C0*std::pow(q(idT,HS+k,HS+j,HS+i,e),gamma);
      q(idT,HS+k,HS+j,HS+i,e) /= q(idR,HS+k,HS+j,HS+i,e);
    }
    h.dycore.halo_exchange(c, q, tr, pressure);
    SArray<real,2,ORD,2> c2g;  TransformMatrices::coefs_to_gll_lower(c2g);
    SArray<real,1,HS+1> idl; real sigma;
    weno::wenoSetIdealSigma<ORD>(idl, sigma);
    real4d pint("pint",nz+1,ny,nx,nens);
    for (int k=0;k<=nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++) {
      real edge[2];
      for (int side=0; side<2; side++) {
        SArray<real,2,ORD,ORD> s2c;  SArray<real,3,HS,HS,HS> wrl;  SArray<real,1,ORD> st;
        for (int a=0;a<ORD;a++) for (int b=0;b<ORD;b++) s2c(a,b) = vs2c(k+side,a,b,e);
        for (int a=0;a<HS;a++) for (int b=0;b<HS;b++) for (int d=0;d<HS;d++) wrl(a,b,d) = vwrl(k+side,a,b,d,e);
        for (int s=0;s<ORD;s++) st(s) = pressure(k+side+s,HS+j,HS+i,e);
        edge[side] = Dycore::reconstruct(st, c2g, s2c, wrl, idl, sigma, 1-side);
      }
      pint(k,j,i,e) = 0.5_fp * (edge[0] + edge[1]);
    }
    real r_nx_ny = 1./(nx*ny);
    for (int k=0;k<nz;k++) for (int e=0;e<nens;e++) {
      real acc = 0;
      for (int j=0;j<ny;j++) for (int i=0;i<nx;i++)
        acc += -(pint(k+1,j,i,e)-pint(k,j,i,e))/(q(idR,HS+k,HS+j,HS+i,e)*dz(k,e))*r_nx_ny;
      grav_var(k,e) = acc;
    }
    return 0;
  } catch (std::string &s) { std::cerr << "ref_declare_hydrostatic_intended: " << s << std::endl; return 1; }
    catch (...) { return 1; }
}

// Dycore::compute_mass (Dycore.h:37-60) on the coupler's current fields: out[(nt+2)*nens]
int ref_compute_mass(void *hv, double *out) {
  Handle &h = *(Handle*)hv;
  auto &c = h.coupler;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  try {
    real5d q ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d tr("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    q = 0; if (nt>0) tr = 0;
    h.dycore.convert_coupler_to_dynamics(c, q, tr);
    auto m = h.dycore.compute_mass(c, q, tr);
    std::memcpy(out, m.data(), (size_t)(nt+2)*nens*sizeof(double));
    return 0;
  } catch (...) { return 1; }
}

// ---------------------------------------------------------------------------------------------------
// Coupler modules either side of the dycore: the reference's own functions on the
// handle's coupler. Options they read are set through ref_set_option_real / _int.
//   modules::sponge_layer                    pam_core/modules/sponge_layer.h:9-96
//   modules::compute_gcm_forcing_tendencies  pam_core/modules/gcm_forcing.h:18-205
//   modules::apply_gcm_forcing_tendencies    pam_core/modules/gcm_forcing.h:294-436 (+ fill_holes :208-289)
//   modules::perturb_temperature             pam_core/modules/perturb_temperature.h:10-63
// ---------------------------------------------------------------------------------------------------
void ref_set_option_real(void *hv, const char *name, double v) { ((Handle*)hv)->coupler.set_option<real>(name, v); }
void ref_set_option_int (void *hv, const char *name, int v)    { ((Handle*)hv)->coupler.set_option<int>(name, v); }
int ref_entry_exists(void *hv, const char *name) {
  return ((Handle*)hv)->coupler.get_data_manager_device_readwrite().entry_exists(name) ? 1 : 0;
}
int ref_sponge_layer(void *hv) {
  try { modules::sponge_layer(((Handle*)hv)->coupler); return 0; }
  catch (std::string &s) { std::cerr << "ref_sponge_layer: " << s << std::endl; return 1; } catch (...) { return 1; }
}
int ref_compute_gcm_forcing_tendencies(void *hv) {
  try { modules::compute_gcm_forcing_tendencies(((Handle*)hv)->coupler); return 0; }
  catch (std::string &s) { std::cerr << "ref_compute_gcm_forcing_tendencies: " << s << std::endl; return 1; } catch (...) { return 1; }
}
int ref_apply_gcm_forcing_tendencies(void *hv) {
  try { modules::apply_gcm_forcing_tendencies(((Handle*)hv)->coupler); return 0; }
  catch (std::string &s) { std::cerr << "ref_apply_gcm_forcing_tendencies: " << s << std::endl; return 1; } catch (...) { return 1; }
}
int ref_perturb_temperature(void *hv, const int *ids, double magnitude) {
  Handle &h = *(Handle*)hv;
  try {
    int1d id("id", h.nens);
    for (int e=0;e<h.nens;e++) id(e) = ids[e];
    modules::perturb_temperature(h.coupler, id, magnitude);
    return 0;
  } catch (std::string &s) { std::cerr << "ref_perturb_temperature: " << s << std::endl; return 1; } catch (...) { return 1; }
}

// ---------------------------------------------------------------------------------------------------
// Idealized initial states.
//   ref_init_supercell : the reference's own Dycore::init_supercell (Dycore.h:1096-1276) followed by
//                        convert_dynamics_to_coupler (:1281-1331), exactly what Dycore::init does for
//                        initData = supercell (:1016, :1087).
//   ref_init_thermal   : the DATA_SPEC_THERMAL branch of Dycore::init (:1017-1086) sits behind
//                        PAM_STANDALONE + a YAML file and cannot be called on its own, so its two loops are
//                        re-driven here around the reference's own point functions Dycore::hydro_const_theta
//                        (:742) and Dycore::sample_ellipse_cosine (:757) and its GLL tables, then handed to
//                        the reference's convert_dynamics_to_coupler.
//   ref_gll9           : TransformMatrices::get_gll_points / get_gll_weights, 9 points.
// ---------------------------------------------------------------------------------------------------
void ref_gll9(double *pts, double *wts) {
  SArray<real,1,9> p, w;
  TransformMatrices::get_gll_points(p); TransformMatrices::get_gll_weights(w);
  for (int i=0;i<9;i++) { pts[i] = p(i); wts[i] = w(i); }
}
int ref_init_supercell(void *hv) {
  Handle &h = *(Handle*)hv;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  try {
    real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    state = 0; tracers = 0;
    h.dycore.init_supercell(h.coupler, state, tracers);
    h.dycore.convert_dynamics_to_coupler(h.coupler, state, tracers);
    return 0;
  } catch (std::string &s) { std::cerr << "ref_init_supercell: " << s << std::endl; return 1; } catch (...) { return 1; }
}
int ref_init_thermal(void *hv) {
  Handle &h = *(Handle*)hv;
  auto &c = h.coupler;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  try {
    real R_d = c.get_option<real>("R_d"), cp_d = c.get_option<real>("cp_d"), p0 = c.get_option<real>("p0");
    real grav = c.get_option<real>("grav"), gamma = c.get_option<real>("gamma_d"), C0 = c.get_option<real>("C0");
    int idWV = c.get_option<int>("idWV");
    real xlen = c.get_xlen(), ylen = c.get_ylen(), dx = c.get_dx(), dy = c.get_dy();
    bool sim2d = ny == 1;
    auto &dm = c.get_data_manager_device_readwrite();
    auto dz   = dm.get<real const,2>("vertical_cell_dz");
    auto zmid = dm.get<real const,2>("vertical_midpoint_height");
    real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    state = 0; tracers = 0;
    real2d hy_dens_cells("hy_dens_cells",nz,nens), hy_pressure_cells("hy_pressure_cells",nz,nens);
    const int nqpoints = 9;
    SArray<real,1,nqpoints> qpoints, qweights;
    TransformMatrices::get_gll_points(qpoints); TransformMatrices::get_gll_weights(qweights);
    for (int k=0;k<nz;k++) for (int iens=0;iens<nens;iens++) {
      hy_dens_cells(k,iens) = 0.; hy_pressure_cells(k,iens) = 0.;
      for (int kk=0; kk<nqpoints; kk++) {
        real z = zmid(k,iens) + qpoints(kk)*dz(k,iens);
        real hr, ht;
        Dycore::hydro_const_theta(z,grav,C0,cp_d,p0,gamma,R_d,hr,ht);
        hy_dens_cells    (k,iens) += hr                       * qweights(kk);
        hy_pressure_cells(k,iens) += C0*std::pow(hr*ht,gamma) * qweights(kk);
      }
    }
    for (int k=0;k<nz;k++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int iens=0;iens<nens;iens++) {
      for (int kk=0; kk<nqpoints; kk++) for (int jj=0; jj<nqpoints; jj++) for (int ii=0; ii<nqpoints; ii++) {
        real x = (i+0.5)*dx + qpoints(ii)*dx;
        real y = (j+0.5)*dy + qpoints(jj)*dy;   if (sim2d) y = ylen/2;
        real z = zmid(k,iens) + qpoints(kk)*dz(k,iens);
        real hr = hy_dens_cells(k,iens), hp = hy_pressure_cells(k,iens);
        real ht = std::pow( hp/C0 , 1._fp/gamma ) / hr;
        real rho = hr, u = 0, v = 0, w = 0, rho_v = 0;
        real theta = ht + Dycore::sample_ellipse_cosine(2._fp , x,y,z , xlen/2,ylen/2,2000. , 2000.,2000.,2000.);
        real wt = qweights(ii)*qweights(jj)*qweights(kk);
        state(idR,HS+k,HS+j,HS+i,iens) += rho       * wt;
        state(idU,HS+k,HS+j,HS+i,iens) += rho*u     * wt;
        state(idV,HS+k,HS+j,HS+i,iens) += rho*v     * wt;
        state(idW,HS+k,HS+j,HS+i,iens) += rho*w     * wt;
        state(idT,HS+k,HS+j,HS+i,iens) += rho*theta * wt;
        if (nt > 0) tracers(idWV,HS+k,HS+j,HS+i,iens) += rho_v * wt;
      }
    }
    h.dycore.convert_dynamics_to_coupler(c, state, tracers);
    return 0;
  } catch (std::string &s) { std::cerr << "ref_init_thermal: " << s << std::endl; return 1; } catch (...) { return 1; }
}

// Wall-clock seconds for `nsteps` raw reference time steps (CPU baseline timing).
double ref_time_steps_raw(void *hv, int nsteps) {
  Handle &h = *(Handle*)hv;
  auto t0 = std::chrono::steady_clock::now();
  for (int s=0;s<nsteps;s++) h.dycore.timeStep(h.coupler);
  yakl::fence();
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1-t0).count();
}

// ---------------------------------------------------------------------------------------------------
// The reference's public methods on caller-owned HALOED arrays, flattened C order (pins the oracle's
// orc_*_haloed / orc_convert_* / orc_halo_exchange, which in turn check the C ABI's entry points).
// ---------------------------------------------------------------------------------------------------
int ref_halo_exchange(void *hv, double *st, double *tr, double *pr) {
  Handle &h = *(Handle*)hv;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real4d pressure("pressure",nz+2*HS,ny+2*HS,nx+2*HS,nens);
  fill_from(state, st); if (nt>0) fill_from(tracers, tr); fill_from(pressure, pr);
  // F2: the bottom ghost theta the racing thread would have written (here the array already holds theta)
  for (int kk=0; kk<HS; kk++) for (int j=0;j<ny;j++) for (int i=0;i<nx;i++) for (int e=0;e<nens;e++)
    state(idT,kk,HS+j,HS+i,e) = state(idT,HS,HS+j,HS+i,e);
  try { h.dycore.halo_exchange(h.coupler, state, tracers, pressure); } catch (...) { return 1; }
  copy_to(state, st); if (nt>0) copy_to(tracers, tr); copy_to(pressure, pr);
  return 0;
}
int ref_compute_tendencies_haloed(void *hv, double *st, double *st_tend, double *tr, double *tr_tend, double dt) {
  Handle &h = *(Handle*)hv;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real5d state_tend  ("state_tend"  ,5 ,nz,ny,nx,nens);
  real5d tracers_tend("tracers_tend",nt,nz,ny,nx,nens);
  fill_from(state, st); if (nt>0) { fill_from(tracers, tr); fill_from(tracers_tend, tr_tend); }
  try { tend_intended(h, state, state_tend, tracers, tracers_tend, dt); } catch (...) { return 1; }
  copy_to(state, st); copy_to(state_tend, st_tend);
  if (nt>0) { copy_to(tracers, tr); copy_to(tracers_tend, tr_tend); }
  return 0;
}
int ref_convert_coupler_to_dynamics(void *hv, double *st, double *tr) {
  Handle &h = *(Handle*)hv;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  fill_from(state, st); if (nt>0) fill_from(tracers, tr);
  try { h.dycore.convert_coupler_to_dynamics(h.coupler, state, tracers); } catch (...) { return 1; }
  copy_to(state, st); if (nt>0) copy_to(tracers, tr);
  return 0;
}
int ref_convert_dynamics_to_coupler(void *hv, const double *st, const double *tr) {
  Handle &h = *(Handle*)hv;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  fill_from(state, st); if (nt>0) fill_from(tracers, tr);
  try { h.dycore.convert_dynamics_to_coupler(h.coupler, state, tracers); } catch (...) { return 1; }
  return 0;
}
int ref_compute_mass_dynamics(void *hv, const double *st, const double *tr, double *out) {
  Handle &h = *(Handle*)hv;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
  fill_from(state, st); if (nt>0) fill_from(tracers, tr);
  try {
    auto m = h.dycore.compute_mass(h.coupler, state, tracers);
    std::memcpy(out, m.data(), (size_t)(nt+2)*nens*sizeof(double));
  } catch (...) { return 1; }
  return 0;
}
// Dycore::init_supercell proper: the haloed arrays it fills (interior), flattened
int ref_init_supercell_dynamics(void *hv, double *st, double *tr) {
  Handle &h = *(Handle*)hv;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  try {
    real5d state  ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d tracers("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    state = 0; tracers = 0;
    h.dycore.init_supercell(h.coupler, state, tracers);
    copy_to(state, st); copy_to(tracers, tr);
    return 0;
  } catch (...) { return 1; }
}

} // extern "C"
