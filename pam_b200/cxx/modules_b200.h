// modules_b200.h -- drop-in replacements for the coupler modules a PAM driver runs either side of the dycore
// (pam_core/modules/sponge_layer.h, gcm_forcing.h, perturb_temperature.h, broadcast_initial_gcm_column.h): the same
// functions in the same namespace with the same signatures, re-homed on the C ABI of include/pam_modules_b200.h
// (sm_100a kernels inside libawfl_b200.so; no YAKL parallel_for, no CPU fallback).
//
//     coupler.run_module( "apply_gcm_forcing_tendencies" , modules::apply_gcm_forcing_tendencies<pam::PamCoupler> );
//     coupler.run_module( "sponge_layer"                 , modules::sponge_layer<pam::PamCoupler>                 );
//
// Templated on the coupler type like pam_b200/cxx/Dycore.h (works with pam::PamCoupler and with pam_coupler_b200.h).
// The scratch workspace of a coupler shape is created on first use and kept for the life of the process.
#pragma once
#include <cstdio>
#include <map>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/awfl_b200.h"
#include "../../include/pam_modules_b200.h"

#ifndef PAM_B200_REAL_DEFINED
#define PAM_B200_REAL_DEFINED
typedef double real;   // pam_core/pam_const.h:22
#endif

namespace modules {
namespace b200_detail {

[[noreturn]] inline void endrun(const std::string &msg) {          // pam_core/pam_const.h:249-252
  std::fprintf(stderr, "%s\n", msg.c_str());
  throw std::string(msg);
}

template <class Coupler>
inline pamm_workspace *workspace(Coupler &c) {
  static std::map<std::tuple<int, int, int, int, int>, pamm_workspace *> cache;
  const auto key = std::make_tuple(c.get_nz(), c.get_ny(), c.get_nx(), c.get_nens(), awfl_current_device());
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  pamm_workspace *w = nullptr;
  if (pamm_create(c.get_nz(), c.get_ny(), c.get_nx(), c.get_nens(), awfl_current_device(), nullptr, &w) != 0)
    endrun(std::string("pamm_create: ") + pamm_last_error(nullptr));
  cache[key] = w;
  return w;
}

inline void check(pamm_workspace *w, int rc, const char *what) {
  if (rc != 0) endrun(std::string(what) + ": " + pamm_last_error(w));
}

template <class DM> inline pamm_crm_fields crm_fields(DM &dm) {
  pamm_crm_fields c;
  c.density_dry = dm.template get<real, 4>("density_dry").data();
  c.uvel = dm.template get<real, 4>("uvel").data();
  c.vvel = dm.template get<real, 4>("vvel").data();
  c.temp = dm.template get<real, 4>("temp").data();
  c.water_vapor = dm.template get<real, 4>("water_vapor").data();
  c.cloud_water = dm.template get<real, 4>("cloud_water").data();
  c.ice = dm.template get<real, 4>("ice").data();
  c.cloud_water_num = dm.template get<real, 4>("cloud_water_num").data();
  c.ice_num = dm.template get<real, 4>("ice_num").data();
  c.rain_num = dm.template get<real, 4>("rain_num").data();
  return c;
}
template <class DM> inline pamm_gcm_columns gcm_columns(DM &dm) {
  pamm_gcm_columns g;
  g.gcm_density_dry = dm.template get<real, 2>("gcm_density_dry").data();
  g.gcm_uvel = dm.template get<real, 2>("gcm_uvel").data();
  g.gcm_vvel = dm.template get<real, 2>("gcm_vvel").data();
  g.gcm_temp = dm.template get<real, 2>("gcm_temp").data();
  g.gcm_water_vapor = dm.template get<real, 2>("gcm_water_vapor").data();
  g.gcm_cloud_water = dm.template get<real, 2>("gcm_cloud_water").data();
  g.gcm_cloud_ice = dm.template get<real, 2>("gcm_cloud_ice").data();
  g.gcm_num_liq = dm.template get<real, 2>("gcm_num_liq").data();
  g.gcm_num_ice = dm.template get<real, 2>("gcm_num_ice").data();
  g.gcm_num_rain = dm.template get<real, 2>("gcm_num_rain").data();
  return g;
}
template <class DM> inline pamm_forcing_tend forcing_tend(DM &dm) {
  pamm_forcing_tend t;
  t.rho_d = dm.template get<real, 2>("gcm_forcing_tend_rho_d").data();
  t.uvel = dm.template get<real, 2>("gcm_forcing_tend_uvel").data();
  t.vvel = dm.template get<real, 2>("gcm_forcing_tend_vvel").data();
  t.temp = dm.template get<real, 2>("gcm_forcing_tend_temp").data();
  t.qtot = dm.template get<real, 2>("gcm_forcing_tend_qtot").data();
  t.qv = dm.template get<real, 2>("gcm_forcing_tend_qv").data();
  t.ql = dm.template get<real, 2>("gcm_forcing_tend_ql").data();
  t.qi = dm.template get<real, 2>("gcm_forcing_tend_qi").data();
  t.rho_v = dm.template get<real, 2>("gcm_forcing_tend_rho_v").data();
  t.rho_l = dm.template get<real, 2>("gcm_forcing_tend_rho_l").data();
  t.rho_i = dm.template get<real, 2>("gcm_forcing_tend_rho_i").data();
  t.nc = dm.template get<real, 2>("gcm_forcing_tend_nc").data();
  t.ni = dm.template get<real, 2>("gcm_forcing_tend_ni").data();
  t.nr = dm.template get<real, 2>("gcm_forcing_tend_nr").data();
  return t;
}

}  // namespace b200_detail

// modules::sponge_layer (pam_core/modules/sponge_layer.h:9-96)
template <class Coupler>
inline void sponge_layer(Coupler &coupler) {
  using namespace b200_detail;
  const int num_layers = coupler.option_exists("sponge_num_layers") ? coupler.template get_option<int>("sponge_num_layers") : 5;
  const real time_scale = coupler.option_exists("sponge_time_scale") ? coupler.template get_option<real>("sponge_time_scale") : 60;
  auto &dm = coupler.get_data_manager_device_readwrite();
  std::vector<double *> f;
  for (const char *n : {"density_dry", "uvel", "vvel", "wvel", "temp"}) f.push_back(dm.template get<real, 4>(n).data());
  for (const std::string &n : coupler.get_tracer_names()) f.push_back(dm.template get<real, 4>(n).data());
  pamm_workspace *w = workspace(coupler);
  check(w, pamm_sponge_layer(w, f.data(), (int)f.size(), dm.template get<real, 2>("vertical_interface_height").data(),
                             dm.template get<real, 2>("vertical_midpoint_height").data(), coupler.template get_option<real>("crm_dt"),
                             time_scale, num_layers), "pamm_sponge_layer");
}

// modules::compute_gcm_forcing_tendencies (pam_core/modules/gcm_forcing.h:18-205)
template <class Coupler>
inline void compute_gcm_forcing_tendencies(Coupler &coupler) {
  using namespace b200_detail;
  auto &dm = coupler.get_data_manager_device_readwrite();
  const int nz = coupler.get_nz(), nens = coupler.get_nens();
  if (!dm.entry_exists("gcm_forcing_tend_uvel")) {                                        // gcm_forcing.h:134-149
    for (const char *n : {"rho_d", "uvel", "vvel", "temp", "qtot", "qv", "ql", "qi", "rho_v", "rho_l", "rho_i", "nc", "ni", "nr"})
      dm.template register_and_allocate<real>(std::string("gcm_forcing_tend_") + n, "GCM forcing", {nz, nens});
  }
  const pamm_crm_fields c = crm_fields(dm);
  const pamm_gcm_columns g = gcm_columns(dm);
  const pamm_forcing_tend t = forcing_tend(dm);
  pamm_workspace *w = workspace(coupler);
  check(w, pamm_compute_gcm_forcing_tendencies(w, &c, &g, &t, coupler.template get_option<real>("gcm_physics_dt")),
        "pamm_compute_gcm_forcing_tendencies");
}

// modules::apply_gcm_forcing_tendencies + fill_holes (pam_core/modules/gcm_forcing.h:208-436)
template <class Coupler>
inline void apply_gcm_forcing_tendencies(Coupler &coupler) {
  using namespace b200_detail;
  auto &dm = coupler.get_data_manager_device_readwrite();
  const pamm_crm_fields c = crm_fields(dm);
  const pamm_gcm_columns g = gcm_columns(dm);
  const pamm_forcing_tend t = forcing_tend(dm);
  pamm_workspace *w = workspace(coupler);
  check(w, pamm_apply_gcm_forcing_tendencies(w, &c, &g, &t, dm.template get<real, 2>("vertical_cell_dz").data(),
                                             coupler.template get_option<real>("crm_dt"), coupler.template get_option<real>("gcm_physics_dt")),
        "pamm_apply_gcm_forcing_tendencies");
}

// modules::broadcast_initial_gcm_column (pam_core/modules/broadcast_initial_gcm_column.h:7-41)
template <class Coupler>
inline void broadcast_initial_gcm_column(Coupler &coupler) {
  using namespace b200_detail;
  auto &dm = coupler.get_data_manager_device_readwrite();
  const double *cols[6];
  double *flds[6];
  const char *cn[6] = {"gcm_density_dry", "gcm_uvel", "gcm_vvel", "gcm_wvel", "gcm_temp", "gcm_water_vapor"};
  const char *fn[6] = {"density_dry", "uvel", "vvel", "wvel", "temp", "water_vapor"};
  for (int i = 0; i < 6; i++) { cols[i] = dm.template get<real, 2>(cn[i]).data(); flds[i] = dm.template get<real, 4>(fn[i]).data(); }
  pamm_workspace *w = workspace(coupler);
  check(w, pamm_broadcast_initial_gcm_column(w, cols, flds), "pamm_broadcast_initial_gcm_column");
}

// modules::perturb_temperature (pam_core/modules/perturb_temperature.h:10-63). id: DEVICE array of nens ints
// (the reference takes an intConst1d; pass its .data()).
template <class Coupler>
inline void perturb_temperature(Coupler &coupler, const int *id_device, int id_size, real magnitude = 0.1) {
  using namespace b200_detail;
  if (id_size != coupler.get_nens()) endrun("ERROR: size of id array must be the same as nens");     // perturb_temperature.h:20
  auto &dm = coupler.get_data_manager_device_readwrite();
  pamm_workspace *w = workspace(coupler);
  check(w, pamm_perturb_temperature(w, dm.template get<real, 4>("temp").data(), id_device, magnitude), "pamm_perturb_temperature");
}

}  // namespace modules
