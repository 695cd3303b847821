// Dycore.h -- drop-in replacement for E3SM-Project/PAM's dynamics/awfl/Dycore.h (class Dycore, lines 13-1550).
//
// A PAM build selects its dycore by directory (`-DPAM_DYCORE=<dir>`, dynamics/CMakeLists.txt:5-17): point it at this
// directory and link libawfl_b200.so. The class below keeps the reference's public method names, argument meaning and
// error behaviour; every compute method pulls raw device pointers out of the coupler's DataManager and calls the C ABI
// of include/awfl_b200.h, whose entry points launch hand-written sm_100a kernels. There is no YAKL parallel_for and no
// CPU fallback behind this header.
//
// The class is templated on the coupler type so that it accepts both the real `pam::PamCoupler` (YAKL-backed
// DataManager: `dm.get<real,4>(name).data()`) and the header-only look-alike of pam_coupler_b200.h (cudaMalloc-backed).
// Required coupler surface (pam_core/pam_coupler.h:59-251): get_nx/ny/nz/nens, get_xlen/ylen, get_option<T>,
// set_option<T>, option_exists, get_tracer_names, get_tracer_info, get_num_tracers,
// get_data_manager_device_readwrite/readonly with get<T,N>(name) -> array with .data(), register_and_allocate<T>,
// entry_exists.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/awfl_b200.h"
#include "../csrc/awfl_tables.h"

#ifndef PAM_B200_REAL_DEFINED
#define PAM_B200_REAL_DEFINED
typedef double real;   // pam_core/pam_const.h:22
#endif

namespace awfl_b200_detail {
// pam_core/pam_const.h:249-252: print, then throw the message
[[noreturn]] inline void endrun(const std::string &msg) {
  std::fprintf(stderr, "%s\n", msg.c_str());
  throw std::string(msg);
}

// Dycore::reconstruct is YAKL_INLINE in the reference (callable from host and device code alike, Dycore.h:591)
#if defined(YAKL_INLINE)
#define AWFL_B200_INLINE YAKL_INLINE
#elif defined(__CUDACC__)
#define AWFL_B200_INLINE __host__ __device__ inline
#else
#define AWFL_B200_INLINE inline
#endif

// TransformMatrices::coefs_to_tv (TransformMatrices.h:188,871,2263,4734): the truncated total variation of a polynomial as
// a sparse quadratic form in its coefficients (tables generated from exact rationals by tools/gen_tables.py)
template <int N> struct TvForm;
#define AWFL_B200_TVFORM(O)                                                                                          \
  template <> struct TvForm<(O + 1) / 2> {                                                                           \
    template <class R, class A> AWFL_B200_INLINE static R eval(A const &a) {                                         \
      const int I[] = AWFL_TVLO_I_##O##_INIT, J[] = AWFL_TVLO_J_##O##_INIT;                                          \
      const double C[] = AWFL_TVLO_C_##O##_INIT;                                                                     \
      R r = 0.;                                                                                                      \
      for (int t = 0; t < AWFL_TVLO_N_##O; t++) r = r + (I[t] == J[t] ? C[t] * (a[I[t]] * a[I[t]]) : C[t] * a[I[t]] * a[J[t]]); \
      return r;                                                                                                      \
    }                                                                                                                \
  };                                                                                                                 \
  template <> struct TvForm<O> {                                                                                     \
    template <class R, class A> AWFL_B200_INLINE static R eval(A const &a) {                                         \
      const int I[] = AWFL_TVHI_I_##O##_INIT, J[] = AWFL_TVHI_J_##O##_INIT;                                          \
      const double C[] = AWFL_TVHI_C_##O##_INIT;                                                                     \
      R r = 0.;                                                                                                      \
      for (int t = 0; t < AWFL_TVHI_N_##O; t++) r = r + (I[t] == J[t] ? C[t] * (a[I[t]] * a[I[t]]) : C[t] * a[I[t]] * a[J[t]]); \
      return r;                                                                                                      \
    }                                                                                                                \
  };
#ifdef MW_ORD
#if MW_ORD == 3
AWFL_B200_TVFORM(3)
#elif MW_ORD == 7
AWFL_B200_TVFORM(7)
#elif MW_ORD == 9
AWFL_B200_TVFORM(9)
#else
AWFL_B200_TVFORM(5)
#endif
#else
AWFL_B200_TVFORM(5)
#endif

// The algorithm of Dycore::reconstruct + weno::compute_weno_coefs, operation for operation as the reference writes it
// (WenoLimiter.h:98-181, Dycore.h:599-603), generic in the scalar type R: `real` for the class method below, a counting type
// for tools/count_flops.cpp (the instrumented flop count the roofline numerator is frozen from).
template <int ord, class R, class S1, class C2G, class S2C, class WRL, class IDL>
AWFL_B200_INLINE R reconstruct_generic(S1 const &stencil, C2G const &coefs_to_gll, S2C const &sten_to_coefs, WRL const &weno_recon_lower,
                                      IDL const &idl, R sigma, int ind) {
  constexpr int hs = (ord + 1) / 2;
  const R eps = 1.0e-20;
  R a_lo[hs][hs], a_hi[ord], tv[hs + 1], wts[hs + 1];
  for (int i = 0; i < hs; i++)
    for (int ii = 0; ii < hs; ii++) {
      R tmp = 0;
      for (int s = 0; s < hs; s++) tmp = tmp + weno_recon_lower(i, s, ii) * stencil(i + s);
      a_lo[i][ii] = tmp;
    }
  for (int ii = 0; ii < ord; ii++) {
    R tmp = 0;
    for (int s = 0; s < ord; s++) tmp = tmp + sten_to_coefs(s, ii) * stencil(s);
    a_hi[ii] = tmp;
  }
  for (int i = 0; i < hs; i++)                                   // the "bridge" polynomial, WenoLimiter.h:131-138
    for (int ii = 0; ii < hs; ii++) a_hi[ii] = a_hi[ii] - idl(i) * a_lo[i][ii];
  for (int ii = 0; ii < ord; ii++) a_hi[ii] = a_hi[ii] / idl(hs);
  R lo_avg = 0;
  for (int i = 0; i < hs; i++) { tv[i] = TvForm<hs>::template eval<R>(a_lo[i]); lo_avg = lo_avg + tv[i]; }
  lo_avg = lo_avg / R(hs);
  tv[hs] = lo_avg + (TvForm<ord>::template eval<R>(a_hi) - lo_avg) * sigma;
  R sum = 0;
  for (int i = 0; i < hs + 1; i++) { wts[i] = idl(i) / (tv[i] * tv[i] + eps); sum = sum + wts[i]; }
  for (int i = 0; i < hs + 1; i++) wts[i] = wts[i] / (sum + eps);  // convexify, WenoLimiter.h:23-29
  sum = 0;
  for (int i = 0; i < hs + 1; i++) {                             // map_weights, WenoLimiter.h:11-19
    const R d = idl(i), w = wts[i];
    wts[i] = w * (d + d * d - R(3) * d * w + w * w) / (d * d + w * (R(1) - R(2) * d));
    sum = sum + wts[i];
  }
  for (int i = 0; i < hs + 1; i++) wts[i] = wts[i] / (sum + eps);
  R val = 0;
  for (int s = 0; s < ord; s++) {
    R aw = wts[hs] * a_hi[s];
    if (s < hs) for (int i = 0; i < hs; i++) aw = aw + wts[i] * a_lo[i][s];
    val = val + coefs_to_gll(s, ind) * aw;
  }
  return val;
}

// the (num_tracers+2, nens) device array Dycore::compute_mass returns, by the type of the dynamics arrays passed in:
// the reference's real2d inside a PAM build (YAKL), the coupler look-alike's own array type otherwise
template <class A5> struct MassArray;
#ifdef YAKL_INLINE
template <class T> struct MassArray<yakl::Array<T, 5, yakl::memDevice, yakl::styleC>> {
  typedef yakl::Array<double, 2, yakl::memDevice, yakl::styleC> type;
  static type make(int nv, int nens) { return type("mass", nv, nens); }
};
#endif
}  // namespace awfl_b200_detail

class Dycore {
 public:
  // The reference fixes the order at compile time with MW_ORD (Dycore.h:17-21); the same macro is honoured here.
#ifdef MW_ORD
  static int const ord = MW_ORD;
#else
  static int const ord = 5;
#endif
  static int const hs = (ord + 1) / 2;
  static int const num_state = 5;
  static int const idR = 0, idU = 1, idV = 2, idW = 3, idT = 4;

  Dycore() : h_(nullptr), nt_(0) {}
  ~Dycore() { release(); }
  Dycore(const Dycore &) = delete;
  Dycore &operator=(const Dycore &) = delete;

  // Dycore::dycore_name (Dycore.h:1544)
  const char *dycore_name() const { return awfl_dycore_name(); }

  // Dycore::init (Dycore.h:835-1091)
  template <class Coupler>
  void init(Coupler &coupler, bool verbose = false) {
    using awfl_b200_detail::endrun;
    (void)verbose;
    release();
    const int nz = coupler.get_nz(), ny = coupler.get_ny(), nx = coupler.get_nx(), nens = coupler.get_nens();
    coupler.template set_option<bool>("balance_hydrostasis_with_gravity", true);             // Dycore.h:866
    set_default<real>(coupler, "R_d", 287.);  set_default<real>(coupler, "cp_d", 1003.);     // Dycore.h:871-876
    set_default<real>(coupler, "R_v", 461.);  set_default<real>(coupler, "cp_v", 1859.);
    set_default<real>(coupler, "p0", 1.e5);   set_default<real>(coupler, "grav", 9.81);
    const std::vector<std::string> names = coupler.get_tracer_names();
    nt_ = (int)names.size();
    std::vector<unsigned char> pos(nt_ > 0 ? nt_ : 1), adds(nt_ > 0 ? nt_ : 1);
    int idWV = -1;
    for (int t = 0; t < nt_; t++) {                                                          // Dycore.h:952-981
      std::string desc; bool found = false, p = false, a = false;
      coupler.get_tracer_info(names[t], desc, found, p, a);
      pos[t] = p; adds[t] = a;
      if (names[t] == "water_vapor") idWV = t;
    }
    if (idWV < 0) endrun("ERROR: the AWFL dycore needs a tracer named water_vapor (Dycore.h:83,974)");
    auto &dm = coupler.get_data_manager_device_readwrite();
    std::vector<double> dz((size_t)nz * nens);
    if (awfl_copy_to_host(dz.data(), dm.template get<real const, 2>("vertical_cell_dz").data(), sizeof(double) * dz.size()))
      endrun("ERROR: could not read vertical_cell_dz from the device");
    awfl_config cfg{};
    cfg.nz = nz; cfg.ny = ny; cfg.nx = nx; cfg.nens = nens; cfg.num_tracers = nt_; cfg.ord = ord;
    cfg.xlen = coupler.get_xlen(); cfg.ylen = coupler.get_ylen();
    cfg.R_d = coupler.template get_option<real>("R_d");   cfg.cp_d = coupler.template get_option<real>("cp_d");
    cfg.R_v = coupler.template get_option<real>("R_v");   cfg.cp_v = coupler.template get_option<real>("cp_v");
    cfg.p0 = coupler.template get_option<real>("p0");     cfg.grav = coupler.template get_option<real>("grav");
    cfg.gamma_d = coupler.option_exists("gamma_d") ? coupler.template get_option<real>("gamma_d") : 0.;
    cfg.kappa_d = coupler.option_exists("kappa_d") ? coupler.template get_option<real>("kappa_d") : 0.;
    cfg.C0 = coupler.option_exists("C0") ? coupler.template get_option<real>("C0") : 0.;
    cfg.idWV = idWV; cfg.tracer_positive = pos.data(); cfg.tracer_adds_mass = adds.data();
    cfg.dz_host = dz.data(); cfg.device = current_device(); cfg.stream = nullptr;   // YAKL's default stream
    if (awfl_create(&cfg, &h_) != 0) endrun(std::string("awfl_create: ") + awfl_last_error(nullptr));
    // options the reference sets if absent (Dycore.h:883-891,974)
    set_default<real>(coupler, "cv_d", awfl_get_option(h_, "cv_d"));
    set_default<real>(coupler, "gamma_d", awfl_get_option(h_, "gamma_d"));
    set_default<real>(coupler, "kappa_d", awfl_get_option(h_, "kappa_d"));
    set_default<real>(coupler, "cv_v", coupler.template get_option<real>("R_v") - coupler.template get_option<real>("cp_v"));   // Dycore.h:886 (sic)
    set_default<real>(coupler, "C0", awfl_get_option(h_, "C0"));
    coupler.template set_option<int>("idWV", idWV);
    // dycore-owned DataManager entries other modules may read (Dycore.h:868,897-898,975-984)
    if (!dm.entry_exists("variable_gravity")) dm.template register_and_allocate<real>("variable_gravity", "", {nz, nens});
    if (!dm.entry_exists("vert_sten_to_coefs")) dm.template register_and_allocate<real>("vert_sten_to_coefs", "", {nz + 2, ord, ord, nens});
    if (!dm.entry_exists("vert_weno_recon_lower")) dm.template register_and_allocate<real>("vert_weno_recon_lower", "", {nz + 2, hs, hs, hs, nens});
    if (!dm.entry_exists("tracer_adds_mass")) dm.template register_and_allocate<bool>("tracer_adds_mass", "", {nt_});   // Dycore.h:975-981
    if (!dm.entry_exists("tracer_positive")) dm.template register_and_allocate<bool>("tracer_positive", "", {nt_});
    {
      std::vector<char> fa(nt_), fp(nt_);     // one byte per flag, the device layout of a bool array
      for (int t = 0; t < nt_; t++) { bool a = adds[t] != 0, p = pos[t] != 0; std::memcpy(&fa[t], &a, 1); std::memcpy(&fp[t], &p, 1); }
      static_assert(sizeof(bool) == 1, "bool flags are copied bytewise");
      if (awfl_copy_to_device(dm.template get<bool, 1>("tracer_adds_mass").data(), fa.data(), (unsigned long)nt_) ||
          awfl_copy_to_device(dm.template get<bool, 1>("tracer_positive").data(), fp.data(), (unsigned long)nt_))
        endrun("ERROR: could not write the tracer flags to the device");
    }
    if (!dm.entry_exists("hy_dens_cells")) dm.template register_and_allocate<real>("hy_dens_cells", "", {nz, nens});
    if (!dm.entry_exists("hy_pressure_cells")) dm.template register_and_allocate<real>("hy_pressure_cells", "", {nz, nens});
    check(awfl_export_tables(h_, dm.template get<real, 2>("variable_gravity").data(), dm.template get<real, 4>("vert_sten_to_coefs").data(),
                             dm.template get<real, 5>("vert_weno_recon_lower").data()), "awfl_export_tables");
    // idealized initial data, Dycore.h:985-1091: "initData" of the standalone driver's YAML file (the reference parses it
    // with yaml-cpp under PAM_STANDALONE; one "key : value" line is all that is read, so it is scanned for here)
    if (coupler.option_exists("standalone_input_file")) {
      const std::string data = yaml_scalar(coupler.template get_option<std::string>("standalone_input_file"), "initData");
      if (data == "thermal") init_idealized_state_and_tracers(coupler, 0);
      else if (data == "supercell") init_idealized_state_and_tracers(coupler, 1);
      else if (data != "external") endrun("ERROR: Invalid data_spec");                       // Dycore.h:1001
    }
  }

  // The DATA_SPEC_THERMAL (0) / DATA_SPEC_SUPERCELL (1) branches of Dycore::init (Dycore.h:1007-1088; init_supercell
  // :1096-1276): overwrite the coupler fields with the idealized state.
  template <class Coupler>
  void init_idealized_state_and_tracers(Coupler &coupler, int data_spec) {
    Fields<Coupler> f(coupler, nt_);
    auto &dm = coupler.get_data_manager_device_readwrite();
    check(awfl_init_idealized(h_, &f.f, dm.template get<real const, 2>("vertical_interface_height").data(), data_spec), "awfl_init_idealized");
  }

  // Dycore::timeStep (Dycore.h:107-255)
  template <class Coupler>
  void timeStep(Coupler &coupler) {
    Fields<Coupler> f(coupler, nt_);
    auto &dm = coupler.get_data_manager_device_readwrite();
    check(awfl_set_variable_gravity(h_, dm.template get<real const, 2>("variable_gravity").data()), "awfl_set_variable_gravity");
#ifdef PAM_DEBUG
    const std::vector<real> mass_init = compute_mass(coupler);                               // Dycore.h:136-138
#endif
    check(awfl_time_step(h_, &f.f, coupler.template get_option<real>("crm_dt"), 0, nullptr), "awfl_time_step");
#ifdef PAM_DEBUG
    // Dycore.h:224-251: warn where a (variable, member) mass changed by more than 1e-10 both relatively and absolutely
    const std::vector<real> mass_final = compute_mass(coupler);
    const int nens = coupler.get_nens();
    for (size_t q = 0; q < mass_init.size(); q++) {
      const double diff = mass_final[q] - mass_init[q], rel = std::fabs(diff) / (std::fabs(mass_init[q]) + 1.e-20);
      if (rel > 1.e-10 && std::fabs(diff) > 1.e-10)
        std::printf("WARNING: conservation violated variable,ensemble,rel_diff,mass_diff,init,final: %d , %d , %10.6e , %10.6e , %10.6e , %10.6e\n",
                    (int)(q / nens), (int)(q % nens), rel, diff, mass_init[q], mass_final[q]);
    }
#endif
  }

  // Dycore::compute_time_step (Dycore.h:65-102)
  template <class Coupler>
  real compute_time_step(Coupler &coupler, real cfl_in = 0.8) {
    Fields<Coupler> f(coupler, nt_);
    double dt = 0.;
    check(awfl_compute_time_step(h_, &f.f, cfl_in, &dt), "awfl_compute_time_step");
    return dt;
  }

  // Dycore::declare_current_profile_as_hydrostatic (Dycore.h:1392-1504)
  template <class Coupler>
  void declare_current_profile_as_hydrostatic(Coupler &coupler, bool use_gcm_data = false) {
    Fields<Coupler> f(coupler, nt_);
    auto &dm = coupler.get_data_manager_device_readwrite();
    awfl_gcm_columns g{};
    if (use_gcm_data) {                                                                      // Dycore.h:1416-1420
      g.gcm_density_dry = dm.template get<real const, 2>("gcm_density_dry").data();
      g.gcm_temp = dm.template get<real const, 2>("gcm_temp").data();
      g.gcm_water_vapor = dm.template get<real const, 2>("gcm_water_vapor").data();
      g.gcm_cloud_water = dm.template get<real const, 2>("gcm_cloud_water").data();
      g.gcm_cloud_ice = dm.template get<real const, 2>("gcm_cloud_ice").data();
    }
    check(awfl_declare_hydrostatic(h_, &f.f, use_gcm_data ? &g : nullptr), "awfl_declare_hydrostatic");
    check(awfl_export_tables(h_, dm.template get<real, 2>("variable_gravity").data(), nullptr, nullptr), "awfl_export_tables");
  }

  // Dycore::compute_mass (Dycore.h:37-60): returns (num_tracers+2) x nens means of q*dz, host memory, row-major
  template <class Coupler>
  std::vector<real> compute_mass(Coupler &coupler) {
    Fields<Coupler> f(coupler, nt_);
    std::vector<real> out((size_t)(nt_ + 2) * coupler.get_nens());
    check(awfl_compute_mass(h_, &f.f, out.data()), "awfl_compute_mass");
    return out;
  }

  // ---- the reference's methods on its HALOED dynamics arrays: real5d (5 | num_tracers, nz+2hs, ny+2hs, nx+2hs, nens) ----
  // Any array type with .data() works: yakl::Array inside a PAM build, the look-alike's views elsewhere.

  // Dycore::compute_mass (Dycore.h:37-60): (num_tracers+2, nens) device array of per-member means of q*dz
  template <class Coupler, class A5>
  typename awfl_b200_detail::MassArray<A5>::type compute_mass(Coupler const &coupler, A5 const &state, A5 const &tracers) const {
    typedef awfl_b200_detail::MassArray<A5> M;
    typename M::type mass = M::make(nt_ + 2, coupler.get_nens());
    check(awfl_compute_mass_dynamics(h_, state.data(), tracers.data(), mass.data()), "awfl_compute_mass_dynamics");
    return mass;
  }

  // Dycore::compute_tendencies (Dycore.h:262-586): tracers_tend carries the FCT start mass in and the tendencies out;
  // state / tracers come back divided by density and multiplied back, ghosts filled (what the SSPRK3 update consumes)
  template <class Coupler, class A5>
  void compute_tendencies(Coupler &coupler, A5 const &state, A5 const &state_tend, A5 const &tracers, A5 const &tracers_tend, real dt) const {
    auto &dm = coupler.get_data_manager_device_readwrite();
    check(awfl_set_variable_gravity(h_, dm.template get<real const, 2>("variable_gravity").data()), "awfl_set_variable_gravity");   // Dycore.h:277
    check(awfl_compute_tendencies_haloed(h_, state.data(), state_tend.data(), tracers.data(), tracers_tend.data(), dt),
          "awfl_compute_tendencies_haloed");
  }

  // Dycore::reconstruct (Dycore.h:591-604) + weno::compute_weno_coefs (WenoLimiter.h:98-181) with caller-supplied matrices:
  // value of the WENO-limited polynomial of the centre cell at its left (ind = 0) or right (ind = 1) edge. Host- and
  // device-callable like the reference's; the time step itself uses the fused row form of csrc/awfl_weno.cuh instead.
  template <class S1, class C2G, class S2C, class WRL, class IDL>
  AWFL_B200_INLINE static real reconstruct(S1 const &stencil, C2G const &coefs_to_gll, S2C const &sten_to_coefs, WRL const &weno_recon_lower,
                                           IDL const &idl, real sigma, int ind) {
    return awfl_b200_detail::reconstruct_generic<ord, real>(stencil, coefs_to_gll, sten_to_coefs, weno_recon_lower, idl, sigma, ind);
  }

  // Dycore::halo_exchange (Dycore.h:608-711): periodic x / y ghosts, wall ghosts in z, on the divided arrays
  template <class Coupler, class A5, class A4>
  void halo_exchange(Coupler const & /*coupler*/, A5 const &state, A5 const &tracers, A4 const &pressure) const {
    check(awfl_halo_exchange(h_, state.data(), tracers.data(), pressure.data()), "awfl_halo_exchange");
  }

  // Dycore::init_supercell (Dycore.h:1096-1276): interior of state / tracers
  template <class Coupler, class A5>
  void init_supercell(Coupler &coupler, A5 &state, A5 &tracers) {
    auto &dm = coupler.get_data_manager_device_readwrite();
    check(awfl_init_supercell(h_, state.data(), tracers.data(), dm.template get<real const, 2>("vertical_interface_height").data()),
          "awfl_init_supercell");
  }

  // Dycore::convert_dynamics_to_coupler (Dycore.h:1281-1331)
  template <class Coupler, class A5>
  void convert_dynamics_to_coupler(Coupler &coupler, A5 const &state, A5 const &tracers) const {
    Fields<Coupler> f(coupler, nt_);
    check(awfl_convert_dynamics_to_coupler(h_, state.data(), tracers.data(), &f.f), "awfl_convert_dynamics_to_coupler");
  }

  // Dycore::convert_coupler_to_dynamics (Dycore.h:1336-1388)
  template <class Coupler, class A5>
  void convert_coupler_to_dynamics(Coupler &coupler, A5 &state, A5 &tracers) const {
    Fields<Coupler> f(coupler, nt_);
    check(awfl_convert_coupler_to_dynamics(h_, &f.f, state.data(), tracers.data()), "awfl_convert_coupler_to_dynamics");
  }

  // Dycore::finalize (Dycore.h:1548)
  template <class Coupler>
  void finalize(Coupler & /*coupler*/) { release(); }

  awfl_handle *handle() { return h_; }

 private:
  awfl_handle *h_;
  int nt_;

  // the coupler's prognostic fields as raw device pointers (pam_coupler.h:259-263 + tracers)
  template <class Coupler>
  struct Fields {
    awfl_fields f;
    std::vector<double *> tr;
    Fields(Coupler &c, int nt) : tr(nt > 0 ? nt : 1) {
      auto &dm = c.get_data_manager_device_readwrite();
      f.density_dry = dm.template get<real, 4>("density_dry").data();
      f.uvel = dm.template get<real, 4>("uvel").data();
      f.vvel = dm.template get<real, 4>("vvel").data();
      f.wvel = dm.template get<real, 4>("wvel").data();
      f.temp = dm.template get<real, 4>("temp").data();
      const std::vector<std::string> names = c.get_tracer_names();
      if ((int)names.size() != nt) awfl_b200_detail::endrun("ERROR: tracers were added after Dycore::init");
      for (int t = 0; t < nt; t++) tr[t] = dm.template get<real, 4>(names[t]).data();
      f.tracers = tr.data();
    }
  };

  // value of a top-level "key : value" line of a YAML file ("" if absent); comments and quotes stripped
  static std::string yaml_scalar(const std::string &path, const std::string &key) {
    std::FILE *f = std::fopen(path.c_str(), "r");
    if (!f) awfl_b200_detail::endrun("ERROR: could not open " + path);
    char line[1024];
    std::string out;
    while (std::fgets(line, sizeof(line), f)) {
      std::string l(line);
      const size_t hash = l.find('#');
      if (hash != std::string::npos) l.erase(hash);
      const size_t colon = l.find(':');
      if (colon == std::string::npos) continue;
      auto trim = [](std::string x) {
        const char *ws = " \t\r\n\"'";
        const size_t b = x.find_first_not_of(ws), e = x.find_last_not_of(ws);
        return b == std::string::npos ? std::string() : x.substr(b, e - b + 1);
      };
      if (trim(l.substr(0, colon)) == key) { out = trim(l.substr(colon + 1)); break; }
    }
    std::fclose(f);
    return out;
  }

  template <class T, class Coupler>
  static void set_default(Coupler &c, const char *key, T v) {
    if (!c.option_exists(key)) c.template set_option<T>(key, v);
  }
  void check(int rc, const char *what) const {
    if (rc != 0) awfl_b200_detail::endrun(std::string(what) + ": " + awfl_last_error(h_));
  }
  void release() {
    if (h_) { awfl_destroy(h_); h_ = nullptr; }
  }
  static int current_device() { return awfl_current_device(); }
};
