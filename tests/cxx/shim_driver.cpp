// Test driver for the C++ host shim (pam_b200/cxx/Dycore.h + pam_coupler_b200.h): the same call sequence as the
// reference's standalone driver (standalone/mmf_simplified/driver.cpp:120-285) -- set options, allocate_coupler_state,
// set_grid, add tracers, dycore.init, then coupler.run_module("dycore", [&](auto &c){ dycore.timeStep(c); }).
//   shim_driver selftest            no GPU needed: names, DataManager/Options errors, init must fail loudly without a GPU
//   shim_driver run IN OUT NSTEPS   reads a case written by tests/test_cxx_shim.py, steps it, writes the fields back
//   shim_driver dynamics IN OUT     the reference's other public methods on caller-owned HALOED arrays (Dycore.h:37,262,608,
//                                   1096,1281,1336): convert in, compute_mass, compute_tendencies, halo_exchange, convert out,
//                                   init_supercell -- every array is dumped for the Python test to compare with the oracle
//   shim_driver reconstruct IN OUT  no GPU needed: the static Dycore::reconstruct (Dycore.h:591-604) on stencils from IN
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pam_coupler_b200.h"
#include "modules_b200.h"

using pam_b200::PamCoupler;

static int selftest() {
  Dycore dycore;
  if (std::string(dycore.dycore_name()) != "SSPRK3+WENO+FV A-grid") { std::printf("FAIL name\n"); return 1; }
  pam_b200::Options o;
  o.set_option<real>("crm_dt", 2.0);
  if (o.get_option<real>("crm_dt") != 2.0 || !o.option_exists("crm_dt")) { std::printf("FAIL options\n"); return 1; }
  int caught = 0;
  try { o.get_option<real>("absent"); } catch (std::string &s) { caught++; }
  try { o.get_option<int>("crm_dt"); } catch (std::string &s) { caught++; }
  if (caught != 2) { std::printf("FAIL option errors\n"); return 1; }
  // Without a device every allocation fails: the coupler must report it (endrun), never fall back to host memory.
  PamCoupler c;
  bool device_ok = true;
  try { c.allocate_coupler_state(8, 1, 8, 2); } catch (std::string &s) { device_ok = false; }
  if (!device_ok) { std::printf("selftest ok (no device: allocation refused loudly)\n"); return 0; }
  // With a device: DataManager error behaviour and a full init
  caught = 0;
  try { c.get_data_manager_device_readwrite().get<real, 4>("nope"); } catch (std::string &s) { caught++; }
  try { c.get_data_manager_device_readwrite().get<real, 3>("uvel"); } catch (std::string &s) { caught++; }
  try { c.get_data_manager_device_readwrite().get<int, 4>("uvel"); } catch (std::string &s) { caught++; }
  if (caught != 3) { std::printf("FAIL DataManager errors\n"); return 1; }
  try { Dycore d; d.init(c); std::printf("FAIL init accepted a coupler without water_vapor\n"); return 1; } catch (std::string &s) {}
  std::printf("selftest ok (device present)\n");
  return 0;
}

static void rd(FILE *f, void *p, size_t n) { if (std::fread(p, 1, n, f) != n) { std::printf("short read\n"); std::exit(2); } }

static int run(const char *in, const char *out, int nsteps, bool mmf) {
  FILE *f = std::fopen(in, "rb");
  if (!f) { std::printf("cannot open %s\n", in); return 2; }
  int hdr[6];   // nz ny nx nens nt use_default_consts
  rd(f, hdr, sizeof(hdr));
  const int nz = hdr[0], ny = hdr[1], nx = hdr[2], nens = hdr[3], nt = hdr[4];
  double par[9];   // xlen ylen crm_dt R_d cp_d R_v cp_v p0 grav
  rd(f, par, sizeof(par));
  std::vector<double> zint(nz + 1);
  rd(f, zint.data(), sizeof(double) * (nz + 1));
  std::vector<std::string> tnames(nt);
  std::vector<int> tpos(nt), tadds(nt);
  for (int t = 0; t < nt; t++) {
    char name[32]; int flags[2];
    rd(f, name, 32); rd(f, flags, sizeof(flags));
    tnames[t] = name; tpos[t] = flags[0]; tadds[t] = flags[1];
  }
  PamCoupler coupler;
  coupler.set_option<real>("crm_dt", par[2]);
  if (!hdr[5]) {
    coupler.set_option<real>("R_d", par[3]); coupler.set_option<real>("cp_d", par[4]); coupler.set_option<real>("R_v", par[5]);
    coupler.set_option<real>("cp_v", par[6]); coupler.set_option<real>("p0", par[7]); coupler.set_option<real>("grav", par[8]);
  }
  coupler.allocate_coupler_state(nz, ny, nx, nens);
  coupler.set_grid(par[0], par[1], zint);
  for (int t = 0; t < nt; t++) coupler.add_tracer(tnames[t], tnames[t], tpos[t] != 0, tadds[t] != 0);   // micro/sgs register first (driver.cpp:189-191)
  Dycore dycore;
  dycore.init(coupler);
  std::vector<std::string> names = {"density_dry", "uvel", "vvel", "wvel", "temp"};
  for (auto &n : tnames) names.push_back(n);
  const size_t ncell = (size_t)nz * ny * nx * nens;
  std::vector<double> buf(ncell);
  auto &dm = coupler.get_data_manager_device_readwrite();
  for (auto &n : names) { rd(f, buf.data(), sizeof(double) * ncell); dm.write<real>(n, buf.data()); }
  std::fclose(f);
  dycore.declare_current_profile_as_hydrostatic(coupler);
  const double dt_cfl = dycore.compute_time_step(coupler);
  if (mmf) {
    // MMF-style loop (driver.cpp:239-250) through modules_b200.h. GCM columns by a rule the Python test repeats:
    // 0.9 x the value of cell (k, j=0, i=0) of the matching CRM field.
    coupler.set_option<real>("gcm_physics_dt", 20.0);
    const char *gn[10] = {"gcm_density_dry", "gcm_uvel", "gcm_vvel", "gcm_temp", "gcm_water_vapor", "gcm_cloud_water", "gcm_cloud_ice",
                          "gcm_num_liq", "gcm_num_ice", "gcm_num_rain"};
    const char *cn[10] = {"density_dry", "uvel", "vvel", "temp", "water_vapor", "cloud_water", "ice", "cloud_water_num", "ice_num", "rain_num"};
    std::vector<double> col((size_t)nz * nens);
    for (int v = 0; v < 10; v++) {
      dm.read<real>(cn[v], buf.data());
      for (int k = 0; k < nz; k++) for (int e = 0; e < nens; e++) col[(size_t)k * nens + e] = 0.9 * buf[(size_t)k * ny * nx * nens + e];
      dm.write<real>(gn[v], col.data());
    }
    modules::compute_gcm_forcing_tendencies(coupler);
    for (int s = 0; s < nsteps; s++) {
      coupler.run_module("apply_gcm_forcing_tendencies", modules::apply_gcm_forcing_tendencies<PamCoupler>);
      coupler.run_module("dycore", [&](PamCoupler &c) { dycore.timeStep(c); });
      coupler.run_module("sponge_layer", modules::sponge_layer<PamCoupler>);
    }
  } else {
    for (int s = 0; s < nsteps; s++) coupler.run_module("dycore", [&](PamCoupler &c) { dycore.timeStep(c); });   // driver.cpp:248
  }
  std::vector<real> mass = dycore.compute_mass(coupler);
  FILE *g = std::fopen(out, "wb");
  std::fwrite(&dt_cfl, sizeof(double), 1, g);
  std::vector<double> gv((size_t)nz * nens);
  dm.read<real>("variable_gravity", gv.data());
  std::fwrite(gv.data(), sizeof(double), gv.size(), g);
  std::fwrite(mass.data(), sizeof(double), mass.size(), g);
  for (auto &n : names) { dm.read<real>(n, buf.data()); std::fwrite(buf.data(), sizeof(double), ncell, g); }
  std::fclose(g);
  dycore.finalize(coupler);
  coupler.finalize();
  std::printf("shim_driver: %d steps of %s done\n", nsteps, dycore.dycore_name());
  return 0;
}

// small fixed-size matrices with the SArray call syntax, filled from the uniform-grid tables of the order compiled in
template <int D0, int D1 = 1, int D2 = 1> struct Mat {
  double v[D0 * D1 * D2];
  double operator()(int i) const { return v[i]; }
  double operator()(int i, int j) const { return v[i * D1 + j]; }
  double operator()(int i, int j, int k) const { return v[(i * D1 + j) * D2 + k]; }
};
#define CAT3_(a, b, c) a##b##c
#define CAT3(a, b, c) CAT3_(a, b, c)
#ifndef MW_ORD
#define MW_ORD_ 5
#else
#define MW_ORD_ MW_ORD
#endif

static int reconstruct_mode(const char *in, const char *out) {
  constexpr int ord = Dycore::ord, hs = Dycore::hs;
  Mat<ord, ord> s2c; Mat<ord, 2> c2g; Mat<hs, hs, hs> wrl; Mat<hs + 1> idl;
  std::memcpy(s2c.v, CAT3(AWFL_S2C_, MW_ORD_, ), sizeof(s2c.v));
  std::memcpy(c2g.v, CAT3(AWFL_C2G_, MW_ORD_, ), sizeof(c2g.v));
  std::memcpy(wrl.v, CAT3(AWFL_WRL_, MW_ORD_, ), sizeof(wrl.v));
  double sum = 0;                                                              // wenoSetIdealSigma, WenoLimiter.h:32-95
  for (int i = 0; i < hs + 1; i++) sum += CAT3(AWFL_IDL_RAW_, MW_ORD_, )[i];
  for (int i = 0; i < hs + 1; i++) idl.v[i] = CAT3(AWFL_IDL_RAW_, MW_ORD_, )[i] / (sum + 1.0e-20);
  const double sigma = CAT3(AWFL_SIGMA_, MW_ORD_, );
  FILE *f = std::fopen(in, "rb");
  if (!f) return 2;
  int n = 0;
  rd(f, &n, sizeof(int));
  std::vector<double> st((size_t)n * ord), res((size_t)2 * n);
  rd(f, st.data(), sizeof(double) * st.size());
  std::fclose(f);
  for (int q = 0; q < n; q++) {
    Mat<ord> u;
    for (int s = 0; s < ord; s++) u.v[s] = st[(size_t)q * ord + s];
    for (int ind = 0; ind < 2; ind++) res[(size_t)ind * n + q] = Dycore::reconstruct(u, c2g, s2c, wrl, idl, sigma, ind);
  }
  FILE *g = std::fopen(out, "wb");
  std::fwrite(res.data(), sizeof(double), res.size(), g);
  std::fclose(g);
  return 0;
}

static int dynamics_mode(const char *in, const char *out) {
  using pam_b200::DeviceArray;
  FILE *f = std::fopen(in, "rb");
  if (!f) { std::printf("cannot open %s\n", in); return 2; }
  int hdr[6];
  rd(f, hdr, sizeof(hdr));
  const int nz = hdr[0], ny = hdr[1], nx = hdr[2], nens = hdr[3], nt = hdr[4];
  double par[9];
  rd(f, par, sizeof(par));
  std::vector<double> zint(nz + 1);
  rd(f, zint.data(), sizeof(double) * (nz + 1));
  std::vector<std::string> tnames(nt);
  std::vector<int> tpos(nt), tadds(nt);
  for (int t = 0; t < nt; t++) {
    char name[32]; int flags[2];
    rd(f, name, 32); rd(f, flags, sizeof(flags));
    tnames[t] = name; tpos[t] = flags[0]; tadds[t] = flags[1];
  }
  PamCoupler coupler;
  coupler.set_option<real>("crm_dt", par[2]);
  coupler.set_option<real>("R_d", par[3]); coupler.set_option<real>("cp_d", par[4]); coupler.set_option<real>("R_v", par[5]);
  coupler.set_option<real>("cp_v", par[6]); coupler.set_option<real>("p0", par[7]); coupler.set_option<real>("grav", par[8]);
  coupler.allocate_coupler_state(nz, ny, nx, nens);
  coupler.set_grid(par[0], par[1], zint);
  for (int t = 0; t < nt; t++) coupler.add_tracer(tnames[t], tnames[t], tpos[t] != 0, tadds[t] != 0);
  Dycore dycore;
  dycore.init(coupler);
  std::vector<std::string> names = {"density_dry", "uvel", "vvel", "wvel", "temp"};
  for (auto &n : tnames) names.push_back(n);
  const size_t ncell = (size_t)nz * ny * nx * nens;
  std::vector<double> buf(ncell), fct((size_t)nt * ncell);
  auto &dm = coupler.get_data_manager_device_readwrite();
  for (size_t v = 0; v < names.size(); v++) {
    rd(f, buf.data(), sizeof(double) * ncell);
    dm.write<real>(names[v], buf.data());
    if (v >= 5) for (size_t q = 0; q < ncell; q++) fct[(v - 5) * ncell + q] = buf[q] > 0 ? buf[q] : 0.;
  }
  std::fclose(f);
  dycore.declare_current_profile_as_hydrostatic(coupler);
  const int hs = Dycore::hs;
  // what Dycore::timeStep declares (Dycore.h:122-126,149-152), zero-filled
  DeviceArray<real, 5> state({Dycore::num_state, nz + 2 * hs, ny + 2 * hs, nx + 2 * hs, nens}), tracers({nt, nz + 2 * hs, ny + 2 * hs, nx + 2 * hs, nens});
  DeviceArray<real, 5> state_tend({Dycore::num_state, nz, ny, nx, nens}), tracers_tend({nt, nz, ny, nx, nens});
  DeviceArray<real, 4> pressure({nz + 2 * hs, ny + 2 * hs, nx + 2 * hs, nens});
  FILE *g = std::fopen(out, "wb");
  auto dump = [&](auto const &a) {
    std::vector<double> h((size_t)a.totElems());
    a.read(h.data());
    std::fwrite(h.data(), sizeof(double), h.size(), g);
  };
  dycore.convert_coupler_to_dynamics(coupler, state, tracers);
  dump(state); dump(tracers);
  auto mass = dycore.compute_mass(coupler, state, tracers);
  dump(mass);
  tracers_tend.write(fct.data());
  dycore.compute_tendencies(coupler, state, state_tend, tracers, tracers_tend, 0.7);
  dump(state); dump(tracers); dump(state_tend); dump(tracers_tend);
  dycore.convert_dynamics_to_coupler(coupler, state, tracers);
  for (auto &n : names) { dm.read<real>(n, buf.data()); std::fwrite(buf.data(), sizeof(double), ncell, g); }
  // halo_exchange on its own: divide the interior first, as compute_tendencies does (the test repeats this in numpy)
  {
    std::vector<double> hst((size_t)state.totElems()), htr((size_t)tracers.totElems());
    state.read(hst.data()); tracers.read(htr.data());
    const size_t nh = hst.size() / 5;
    for (size_t q = 0; q < nh; q++) {
      const double rho = hst[q];
      if (rho == 0.) continue;
      for (int l = 1; l < 5; l++) hst[l * nh + q] /= rho;
      for (int t = 0; t < nt; t++) htr[t * nh + q] /= rho;
    }
    state.write(hst.data()); tracers.write(htr.data());
  }
  dycore.halo_exchange(coupler, state, tracers, pressure);
  dump(state); dump(tracers); dump(pressure);
  dycore.init_supercell(coupler, state, tracers);
  dycore.convert_dynamics_to_coupler(coupler, state, tracers);
  for (auto &n : names) { dm.read<real>(n, buf.data()); std::fwrite(buf.data(), sizeof(double), ncell, g); }
  std::fclose(g);
  dycore.finalize(coupler);
  coupler.finalize();
  std::printf("shim_driver: dynamics-array methods done\n");
  return 0;
}

int main(int argc, char **argv) {
  try {
    if (argc >= 2 && !std::strcmp(argv[1], "selftest")) return selftest();
    if (argc >= 4 && !std::strcmp(argv[1], "reconstruct")) return reconstruct_mode(argv[2], argv[3]);
    if (argc >= 4 && !std::strcmp(argv[1], "dynamics")) return dynamics_mode(argv[2], argv[3]);
    if (argc >= 5 && !std::strcmp(argv[1], "run")) return run(argv[2], argv[3], std::atoi(argv[4]), argc >= 6 && !std::strcmp(argv[5], "mmf"));
  } catch (std::string &s) { std::printf("endrun: %s\n", s.c_str()); return 3; }
  std::printf("usage: shim_driver selftest | run IN OUT NSTEPS [mmf] | dynamics IN OUT | reconstruct IN OUT\n");
  return 64;
}
