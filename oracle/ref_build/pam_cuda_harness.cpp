// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// One driver translation unit over the REAL pam::PamCoupler (pam_core/pam_coupler.h:13-396) and its YAKL-backed
// DataManager (pam_core/DataManager.h:90-153,285-312), compiled with nvcc -DYAKL_ARCH_CUDA for sm_100a from the sources
// where they lie under /root/reference (never copied), twice -- the include path decides which "Dycore.h" it sees, the
// way the reference's build picks a dycore directory (-DPAM_DYCORE=<dir>, dynamics/CMakeLists.txt:5-17):
//
//   oracle/_ref/libpam_ref_ord<N>_cuda.so     -I$(REF)/dynamics/awfl : the UNMODIFIED reference dycore, i.e. its generic
//                                             YAKL lambdas launched through cudaKernelVal
//                                             (externals/YAKL/src/YAKL_parallel_for_common.h:54-76,498-561). The
//                                             same-GPU baseline (bench.py "gpu_reference") and a second parity anchor.
//   oracle/_ref/libpam_dropin_ord<N>_cuda.so  -Ipam_b200/cxx, links libawfl_b200.so : the drop-in test proper -- THIS
//                                             repo's Dycore.h driven by the reference's own coupler on YAKL-allocated
//                                             device fields, coupler.run_module("dycore", ...) as driver.cpp:248 does.
//
// Everything that touches array contents does so on the device (YAKL parallel_for) or through cudaMemcpy: with the CUDA
// backend the DataManager's fields are device memory.
#include "pam_coupler.h"
#include "Dycore.h"
#include <cuda_runtime.h>
#include <cstring>
#include <sstream>

namespace pam { std::mutex data_manager_mutex; }   // pam_core/pam_coupler_globals.cpp:4-6

namespace {

int constexpr ORD = Dycore::ord;
int constexpr HS  = Dycore::hs;
int constexpr idR = Dycore::idR, idT = Dycore::idT;

struct Handle {
  pam::PamCoupler coupler;
  Dycore          dycore;
  int nz, ny, nx, nens, nt;
  std::vector<std::string> tracer_names;
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
    size_t mb = 2048;
    if (const char *e = std::getenv("PAM_REF_POOL_MB")) mb = (size_t) atol(e);
    yakl::init(yakl::InitConfig().set_pool_initial_mb(mb).set_pool_grow_mb(mb));
  }
}

} // namespace


extern "C" {

int pc_ord() { return ORD; }

int pc_is_dropin() {
  #ifdef PAM_B200_DROPIN
    return 1;
  #else
    return 0;
  #endif
}

void pc_shutdown() {
  if (yakl::isInitialized()) { CoutSilencer s; yakl::finalize(); }
}

// zint: host, (nz+1) values if zint_per_ens==0, else (nz+1)*nens (k-major, iens fastest). consts: NULL or
// {R_d, cp_d, R_v, cp_v, p0, grav}.
void *pc_create(int nz, int ny, int nx, int nens, double xlen, double ylen, const double *zint, int zint_per_ens,
                int num_tracers, const char **tracer_names, const int *positive, const int *adds_mass,
                const double *consts, double crm_dt) {
  ensure_yakl();
  Handle *h = new Handle;
  try {
    h->nz=nz; h->ny=ny; h->nx=nx; h->nens=nens; h->nt=num_tracers;
    auto &c = h->coupler;
    c.set_option<real>("crm_dt", crm_dt);
    if (consts) {
      c.set_option<real>("R_d" , consts[0]);  c.set_option<real>("cp_d", consts[1]);
      c.set_option<real>("R_v" , consts[2]);  c.set_option<real>("cp_v", consts[3]);
      c.set_option<real>("p0"  , consts[4]);  c.set_option<real>("grav", consts[5]);
    }
    c.allocate_coupler_state(nz, ny, nx, nens);
    realHost2d zh("zint_host", nz+1, nens);
    for (int k=0;k<=nz;k++) for (int e=0;e<nens;e++) zh(k,e) = zint_per_ens ? zint[k*nens+e] : zint[k];
    c.set_grid(xlen, ylen, realConst2d(zh.createDeviceCopy()));
    for (int t=0;t<num_tracers;t++) {
      c.add_tracer(tracer_names[t], tracer_names[t], positive[t]!=0, adds_mass[t]!=0);
      h->tracer_names.push_back(tracer_names[t]);
      auto f = c.get_data_manager_device_readwrite().get<real,4>(tracer_names[t]);
      f = 0;                                              // add_tracer does not zero the field
    }
    {
      CoutSilencer s;                                     // the reference's init prints a leftover debug solve (Dycore.h:851-864)
      h->dycore.init(c);
    }
    auto g = c.get_data_manager_device_readwrite().get<real,2>("variable_gravity");
    g = 0;                                                // allocated but never filled by init
    yakl::fence();
  } catch (std::string &s) { std::cerr << "pc_create: " << s << std::endl; delete h; return nullptr; }
    catch (...)            { std::cerr << "pc_create: exception" << std::endl; delete h; return nullptr; }
  return h;
}

void pc_destroy(void *hv) {
  Handle *h = (Handle*) hv;
  yakl::fence();
  h->dycore.finalize(h->coupler);
  delete h;
}

const char *pc_dycore_name(void *hv) {
  static std::string name;
  name = ((Handle*)hv)->dycore.dycore_name();
  return name.c_str();
}

void pc_set_crm_dt(void *hv, double dt) { ((Handle*)hv)->coupler.set_option<real>("crm_dt", dt); }
double pc_get_option(void *hv, const char *name) { return ((Handle*)hv)->coupler.get_option<real>(name); }
int pc_entry_exists(void *hv, const char *name) {
  return ((Handle*)hv)->coupler.get_data_manager_device_readwrite().entry_exists(name) ? 1 : 0;
}

// Any real-typed DataManager entry, flattened C order; host <-> device copies.
long pc_field_size(void *hv, const char *name) {
  auto &dm = ((Handle*)hv)->coupler.get_data_manager_device_readwrite();
  return (long) dm.get_collapsed<real>(name).size();
}
int pc_set_field(void *hv, const char *name, const double *src, long n) {
  auto &dm = ((Handle*)hv)->coupler.get_data_manager_device_readwrite();
  auto a = dm.get_collapsed<real>(name);
  if ((long)a.size() != n) return 1;
  yakl::fence();
  return cudaMemcpy(a.data(), src, n*sizeof(double), cudaMemcpyHostToDevice) == cudaSuccess ? 0 : 2;
}
int pc_get_field(void *hv, const char *name, double *dst, long n) {
  auto &dm = ((Handle*)hv)->coupler.get_data_manager_device_readonly();
  auto a = dm.get_collapsed<real const>(name);
  if ((long)a.size() != n) return 1;
  yakl::fence();
  return cudaMemcpy(dst, a.data(), n*sizeof(double), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : 2;
}
// bool-typed entries (tracer_positive, tracer_adds_mass) as 0/1 ints
int pc_get_flags(void *hv, const char *name, int *dst, int n) {
  try {
    auto a = ((Handle*)hv)->coupler.get_data_manager_device_readonly().get<bool const,1>(name).createHostCopy();
    if ((int)a.size() != n) return 1;
    for (int i=0;i<n;i++) dst[i] = a(i) ? 1 : 0;
    return 0;
  } catch (...) { return 2; }
}

double pc_compute_time_step(void *hv, double cfl) {
  Handle &h = *(Handle*)hv;
  return h.dycore.compute_time_step(h.coupler, cfl);
}

int pc_declare_hydrostatic(void *hv, int use_gcm) {
  Handle &h = *(Handle*)hv;
  try { h.dycore.declare_current_profile_as_hydrostatic(h.coupler, use_gcm!=0); yakl::fence(); }
  catch (std::string &s) { std::cerr << "pc_declare_hydrostatic: " << s << std::endl; return 1; } catch (...) { return 1; }
  return 0;
}

// The call of the standalone driver (standalone/mmf_simplified/driver.cpp:248)
int pc_time_step(void *hv) {
  Handle &h = *(Handle*)hv;
  try { h.coupler.run_module("dycore", [&] (pam::PamCoupler &c) { h.dycore.timeStep(c); }); yakl::fence(); }
  catch (std::string &s) { std::cerr << "pc_time_step: " << s << std::endl; return 1; } catch (...) { return 1; }
  return 0;
}

// nsteps of the same call between two CUDA events on the default stream (where YAKL and the shim enqueue): milliseconds
double pc_time_steps_ms(void *hv, int nsteps) {
  Handle &h = *(Handle*)hv;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  yakl::fence();
  cudaEventRecord(e0, 0);
  for (int s=0;s<nsteps;s++) h.coupler.run_module("dycore", [&] (pam::PamCoupler &c) { h.dycore.timeStep(c); });
  cudaEventRecord(e1, 0);
  cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return (double) ms;
}

// Dycore::timeStep (Dycore.h:107-255) restated around the class's PUBLIC methods -- convert_coupler_to_dynamics,
// compute_time_step, compute_tendencies, convert_dynamics_to_coupler -- on harness-owned haloed YAKL arrays, all loops on
// the device. For the reference build this is the "intended semantics" step on the GPU: the bottom ghost theta is
// pre-filled before every compute_tendencies call, which removes the read/write race of Dycore.h:670 vs :692 (SURVEY.md
// F2) the same way oracle/ref_build/ref_harness.cpp does on the CPU. For the drop-in build it exercises the shim's
// haloed-array methods on real YAKL arrays. Returns the sub-cycle count (or -1).
int pc_time_step_public(void *hv, int ncycles_override) {
  using yakl::c::parallel_for;
  using yakl::c::SimpleBounds;
  Handle &h = *(Handle*)hv;
  auto &c = h.coupler;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  try {
    auto positive = c.get_data_manager_device_readonly().get<bool const,1>("tracer_positive");
    real dt_phys = c.get_option<real>("crm_dt");
    real5d q ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d tr("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    q = 0; tr = 0;
    h.dycore.convert_coupler_to_dynamics(c, q, tr);
    parallel_for( SimpleBounds<5>(nt,nz,ny,nx,nens) , YAKL_LAMBDA (int l, int k, int j, int i, int e) {
      if (positive(l)) tr(l,HS+k,HS+j,HS+i,e) = std::max(0., tr(l,HS+k,HS+j,HS+i,e));
    });
    real dt_dyn = h.dycore.compute_time_step(c);
    int ncycles = (int) std::ceil(dt_phys / dt_dyn);
    if (ncycles_override > 0) ncycles = ncycles_override;
    dt_dyn = dt_phys / ncycles;
    real5d q1 ("state_tmp"   ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d t1 ("tracers_tmp" ,nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d dq ("state_tend"  ,5 ,nz,ny,nx,nens);
    real5d dt_("tracers_tend",nt,nz,ny,nx,nens);
    q1 = 0; t1 = 0;
    auto prefill = [&] (real5d const &s) {     // F2: ghost theta = theta of the first interior cell
      parallel_for( SimpleBounds<4>(HS,ny,nx,nens) , YAKL_LAMBDA (int kk, int j, int i, int e) {
        s(idT,kk,HS+j,HS+i,e) = s(idT,HS,HS+j,HS+i,e) / s(idR,HS,HS+j,HS+i,e);
      });
    };
    for (int icycle=0; icycle<ncycles; icycle++) {
      parallel_for( SimpleBounds<5>(nt,nz,ny,nx,nens) , YAKL_LAMBDA (int l, int k, int j, int i, int e) {
        dt_(l,k,j,i,e) = tr(l,HS+k,HS+j,HS+i,e);
      });
      // stage 1 (Dycore.h:156-175)
      prefill(q);
      h.dycore.compute_tendencies(c, q, dq, tr, dt_, dt_dyn);
      parallel_for( SimpleBounds<4>(nz,ny,nx,nens) , YAKL_LAMBDA (int k, int j, int i, int e) {
        for (int l=0;l<5;l++) q1(l,HS+k,HS+j,HS+i,e) = q(l,HS+k,HS+j,HS+i,e) + dt_dyn * dq(l,k,j,i,e);
        for (int l=0;l<nt;l++) {
          real v = tr(l,HS+k,HS+j,HS+i,e) + dt_dyn * dt_(l,k,j,i,e);
          if (positive(l)) v = std::max(0., v);
          t1(l,HS+k,HS+j,HS+i,e) = v;
          dt_(l,k,j,i,e) = (3._fp/4._fp) * tr(l,HS+k,HS+j,HS+i,e) + (1._fp/4._fp) * v;
        }
      });
      // stage 2 (Dycore.h:177-199)
      prefill(q1);
      h.dycore.compute_tendencies(c, q1, dq, t1, dt_, (1._fp/4._fp) * dt_dyn);
      parallel_for( SimpleBounds<4>(nz,ny,nx,nens) , YAKL_LAMBDA (int k, int j, int i, int e) {
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
      });
      // stage 3 (Dycore.h:201-221)
      prefill(q1);
      h.dycore.compute_tendencies(c, q1, dq, t1, dt_, (2._fp/3._fp) * dt_dyn);
      parallel_for( SimpleBounds<4>(nz,ny,nx,nens) , YAKL_LAMBDA (int k, int j, int i, int e) {
        for (int l=0;l<5;l++) q(l,HS+k,HS+j,HS+i,e) = (1._fp/3._fp) * q (l,HS+k,HS+j,HS+i,e) +
                                                      (2._fp/3._fp) * q1(l,HS+k,HS+j,HS+i,e) +
                                                      (2._fp/3._fp) * dt_dyn * dq(l,k,j,i,e);
        for (int l=0;l<nt;l++) {
          real v = (1._fp/3._fp) * tr(l,HS+k,HS+j,HS+i,e) + (2._fp/3._fp) * t1(l,HS+k,HS+j,HS+i,e) +
                   (2._fp/3._fp) * dt_dyn * dt_(l,k,j,i,e);
          if (positive(l)) v = std::max(0., v);
          tr(l,HS+k,HS+j,HS+i,e) = v;
        }
      });
    }
    h.dycore.convert_dynamics_to_coupler(c, q, tr);
    yakl::fence();
    return ncycles;
  } catch (std::string &s) { std::cerr << "pc_time_step_public: " << s << std::endl; return -1; }
    catch (...) { return -1; }
}

// Dycore::compute_mass(coupler, state, tracers) (Dycore.h:37-60) on the coupler's current fields: out[(nt+2)*nens], host
int pc_compute_mass(void *hv, double *out) {
  Handle &h = *(Handle*)hv;
  auto &c = h.coupler;
  int nz=h.nz, ny=h.ny, nx=h.nx, nens=h.nens, nt=h.nt;
  try {
    real5d q ("state"  ,5 ,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    real5d tr("tracers",nt,nz+2*HS,ny+2*HS,nx+2*HS,nens);
    q = 0; tr = 0;
    h.dycore.convert_coupler_to_dynamics(c, q, tr);
    auto m = h.dycore.compute_mass(c, q, tr).createHostCopy();
    std::memcpy(out, m.data(), (size_t)(nt+2)*nens*sizeof(double));
    return 0;
  } catch (...) { return 1; }
}

} // extern "C"
