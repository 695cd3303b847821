// pam_coupler_b200.h -- header-only look-alike of the part of pam::PamCoupler / DataManager / Options that the AWFL
// dycore touches (pam_core/pam_coupler.h:59-356, pam_core/DataManager.h:90-312, pam_core/Options.h:63-131), for C++
// callers and tests that do not build against PAM + YAKL. Same method names, template signatures and error behaviour
// (endrun prints and throws std::string, pam_core/pam_const.h:249-252). Device memory comes from the C ABI's runtime
// helpers, so this header needs no CUDA include. Inside a real PAM build use PAM's own pam_coupler.h: Dycore.h accepts
// either.
#pragma once
#include <cstring>
#include <initializer_list>
#include <map>
#include <string>
#include <type_traits>
#include <typeinfo>
#include <vector>

#include "Dycore.h"

namespace pam_b200 {

using awfl_b200_detail::endrun;

// Non-owning view returned by DataManager::get (the reference returns a non-owning yakl::Array, DataManager.h:292,310)
template <class T, int N>
struct ArrayView {
  T *ptr;
  int dims[N];
  T *data() const { return ptr; }
  int extent(int i) const { return dims[i]; }
  long totElems() const { long n = 1; for (int i = 0; i < N; i++) n *= dims[i]; return n; }
};

// Owning device array created by a caller (the look-alike of `real5d state("state",5,nz+2*hs,...)`): zero-filled
// device memory from the C ABI's runtime helpers, freed with the last copy's owner. Move-only.
template <class T, int N>
class DeviceArray {
 public:
  DeviceArray() : ptr_(nullptr) { for (int i = 0; i < N; i++) dims_[i] = 0; }
  DeviceArray(std::initializer_list<int> dims) : ptr_(nullptr) {
    int i = 0;
    for (int d : dims) if (i < N) dims_[i++] = d;
    void *p = nullptr;
    if (awfl_device_alloc(&p, (unsigned long)totElems() * sizeof(T)) != 0) endrun("ERROR: device allocation failed");
    ptr_ = (T *)p;
  }
  DeviceArray(DeviceArray &&o) noexcept : ptr_(o.ptr_) { for (int i = 0; i < N; i++) dims_[i] = o.dims_[i]; o.ptr_ = nullptr; }
  DeviceArray &operator=(DeviceArray &&o) noexcept {
    if (this != &o) { awfl_device_free(ptr_); ptr_ = o.ptr_; for (int i = 0; i < N; i++) dims_[i] = o.dims_[i]; o.ptr_ = nullptr; }
    return *this;
  }
  DeviceArray(const DeviceArray &) = delete;
  DeviceArray &operator=(const DeviceArray &) = delete;
  ~DeviceArray() { awfl_device_free(ptr_); }
  T *data() const { return ptr_; }
  int extent(int i) const { return dims_[i]; }
  long totElems() const { long n = 1; for (int i = 0; i < N; i++) n *= dims_[i]; return n; }
  void read(T *host) const { if (awfl_copy_to_host(host, ptr_, (unsigned long)totElems() * sizeof(T))) endrun("ERROR: copy to host failed"); }
  void write(const T *host) { if (awfl_copy_to_device(ptr_, host, (unsigned long)totElems() * sizeof(T))) endrun("ERROR: copy to device failed"); }

 private:
  T *ptr_;
  int dims_[N];
};

class DataManager {
 public:
  DataManager() {}
  ~DataManager() { finalize(); }
  DataManager(const DataManager &) = delete;

  // DataManager.h:90-153
  template <class T>
  void register_and_allocate(std::string name, std::string desc, std::initializer_list<int> dims,
                             std::initializer_list<std::string> /*dim_names*/ = {}) {
    if (entries_.count(name)) endrun("ERROR: Duplicate entry name: " + name);        // DataManager.h:104-108
    Entry e;
    e.desc = desc; e.dims.assign(dims.begin(), dims.end()); e.type = typeid(T).hash_code(); e.dirty = false;
    long n = 1;
    for (int d : e.dims) n *= d;
    if (n <= 0) endrun("ERROR: non-positive dimension for " + name);
    e.bytes = (unsigned long)n * sizeof(T);
    if (awfl_device_alloc(&e.ptr, e.bytes) != 0) endrun("ERROR: device allocation failed for " + name);
    entries_[name] = e;
  }
  void unregister_and_deallocate(std::string name) {
    auto it = entries_.find(name);
    if (it == entries_.end()) return;
    awfl_device_free(it->second.ptr);
    entries_.erase(it);
  }
  bool entry_exists(std::string name) const { return entries_.count(name) > 0; }

  // DataManager.h:285-312: a non-const T marks the entry dirty
  template <class T, int N>
  ArrayView<T, N> get(std::string name) {
    Entry &e = find(name);
    typedef typename std::remove_const<T>::type TNC;
    if (e.type != typeid(TNC).hash_code()) endrun("ERROR: Requested type is different from entry type for " + name);   // DataManager.h:576-583
    if ((int)e.dims.size() != N) endrun("ERROR: Requested dimensions is different from entry dimensions for " + name); // DataManager.h:569-575
    if (!std::is_const<T>::value) e.dirty = true;                                                                     // DataManager.h:306
    ArrayView<T, N> v;
    v.ptr = (T *)e.ptr;
    for (int i = 0; i < N; i++) v.dims[i] = e.dims[i];
    return v;
  }
  template <class T, int N>
  ArrayView<T, N> get(std::string name) const { return const_cast<DataManager *>(this)->get<T, N>(name); }
  template <class T>
  ArrayView<T, 1> get_collapsed(std::string name) {
    Entry &e = find(name);
    if (!std::is_const<T>::value) e.dirty = true;
    ArrayView<T, 1> v;
    v.ptr = (T *)e.ptr;
    long n = 1;
    for (int d : e.dims) n *= d;
    v.dims[0] = (int)n;
    return v;
  }
  bool entry_is_dirty(std::string name) { return find(name).dirty; }
  void clean_all_entries() { for (auto &kv : entries_) kv.second.dirty = false; }
  std::vector<std::string> get_dirty_entries() const {
    std::vector<std::string> r;
    for (auto &kv : entries_) if (kv.second.dirty) r.push_back(kv.first);
    return r;
  }
  // host <-> device transfer of a whole entry (the reference goes through yakl::Array::deep_copy_to)
  template <class T> void write(std::string name, const T *host) {
    Entry &e = find(name);
    if (awfl_copy_to_device(e.ptr, host, e.bytes)) endrun("ERROR: copy to device failed for " + name);
  }
  template <class T> void read(std::string name, T *host) {
    Entry &e = find(name);
    if (awfl_copy_to_host(host, e.ptr, e.bytes)) endrun("ERROR: copy to host failed for " + name);
  }
  void finalize() {                                                                  // DataManager.h:596-602
    for (auto &kv : entries_) awfl_device_free(kv.second.ptr);
    entries_.clear();
  }

 private:
  struct Entry { std::string desc; std::vector<int> dims; size_t type; void *ptr; unsigned long bytes; bool dirty; };
  std::map<std::string, Entry> entries_;
  Entry &find(const std::string &name) {
    auto it = entries_.find(name);
    if (it == entries_.end()) endrun("ERROR: Could not find entry: " + name);        // DataManager.h:553-560
    return it->second;
  }
};

// Options.h:63-131 (typed key/value store)
class Options {
 public:
  template <class T> void set_option(std::string key, T v) {
    Val x; x.type = typeid(T).hash_code(); std::memset(x.raw, 0, sizeof(x.raw));
    if constexpr (std::is_same<T, std::string>::value) { x.str = v; }                // Options.h stores std::string too
    else {
      static_assert(sizeof(T) <= 16, "option type too large");
      std::memcpy(x.raw, &v, sizeof(T));
    }
    vals_[key] = x;
  }
  template <class T> T get_option(std::string key) const {
    auto it = vals_.find(key);
    if (it == vals_.end()) endrun("ERROR: option key does not exist: " + key);       // Options.h:84-87
    if (it->second.type != typeid(T).hash_code()) endrun("ERROR: Requesting option with the wrong type: " + key);
    if constexpr (std::is_same<T, std::string>::value) { return it->second.str; }
    else { T v; std::memcpy(&v, it->second.raw, sizeof(T)); return v; }
  }
  bool option_exists(std::string key) const { return vals_.count(key) > 0; }
  void delete_option(std::string key) { vals_.erase(key); }

 private:
  struct Val { size_t type; unsigned char raw[16]; std::string str; };
  std::map<std::string, Val> vals_;
};

class PamCoupler {
 public:
  PamCoupler() : xlen_(0), ylen_(0), nz_(0), ny_(0), nx_(0), nens_(0) {}
  real get_xlen() const { return xlen_; }
  real get_ylen() const { return ylen_; }
  real get_dx() const { return xlen_ / nx_; }
  real get_dy() const { return ylen_ / ny_; }
  int get_nx() const { return nx_; }
  int get_ny() const { return ny_; }
  int get_nz() const { return nz_; }
  int get_nens() const { return nens_; }
  DataManager const &get_data_manager_device_readonly() const { return dm_; }
  DataManager &get_data_manager_device_readwrite() { return dm_; }
  template <class T> void set_option(std::string key, T v) { opt_.set_option<T>(key, v); }
  template <class T> void add_option(std::string key, T v) { opt_.set_option<T>(key, v); }
  template <class T> T get_option(std::string key) const { return opt_.get_option<T>(key); }
  bool option_exists(std::string key) const { return opt_.option_exists(key); }
  void delete_option(std::string key) { opt_.delete_option(key); }

  // pam_coupler.h:139-160
  template <class F> void run_module(std::string /*name*/, F const &f) { f(*this); }

  // pam_coupler.h:255-356 (the entries the dycore path touches)
  void allocate_coupler_state(int nz, int ny, int nx, int nens) {
    nz_ = nz; ny_ = ny; nx_ = nx; nens_ = nens;
    for (const char *n : {"density_dry", "uvel", "vvel", "wvel", "temp"}) dm_.register_and_allocate<real>(n, n, {nz, ny, nx, nens});
    dm_.register_and_allocate<real>("vertical_interface_height", "", {nz + 1, nens});
    dm_.register_and_allocate<real>("vertical_cell_dz", "", {nz, nens});
    dm_.register_and_allocate<real>("vertical_midpoint_height", "", {nz, nens});
    for (const char *n : {"gcm_density_dry", "gcm_uvel", "gcm_vvel", "gcm_wvel", "gcm_temp", "gcm_water_vapor", "gcm_cloud_water", "gcm_cloud_ice",
                          "gcm_num_liq", "gcm_num_ice", "gcm_num_rain"})   // pam_coupler.h:268-281
      dm_.register_and_allocate<real>(n, "", {nz, nens});
  }
  // pam_coupler.h:184-202: zint is a HOST array of nz+1 interface heights, the same for every ensemble member
  void set_grid(real xlen, real ylen, const std::vector<real> &zint) {
    if ((int)zint.size() != nz_ + 1) endrun("ERROR: zint must hold nz+1 interface heights");
    xlen_ = xlen; ylen_ = ylen;
    std::vector<real> zi((size_t)(nz_ + 1) * nens_), dz((size_t)nz_ * nens_), zm((size_t)nz_ * nens_);
    for (int k = 0; k <= nz_; k++) for (int e = 0; e < nens_; e++) zi[(size_t)k * nens_ + e] = zint[k];
    for (int k = 0; k < nz_; k++) for (int e = 0; e < nens_; e++) {
      dz[(size_t)k * nens_ + e] = zint[k + 1] - zint[k];
      zm[(size_t)k * nens_ + e] = 0.5 * (zint[k] + zint[k + 1]);
    }
    dm_.write<real>("vertical_interface_height", zi.data());
    dm_.write<real>("vertical_cell_dz", dz.data());
    dm_.write<real>("vertical_midpoint_height", zm.data());
  }
  // pam_coupler.h:206-251
  void add_tracer(std::string name, std::string desc, bool positive, bool adds_mass) {
    dm_.register_and_allocate<real>(name, desc, {nz_, ny_, nx_, nens_});
    tracers_.push_back({name, desc, positive, adds_mass});
  }
  int get_num_tracers() const { return (int)tracers_.size(); }
  std::vector<std::string> get_tracer_names() const {
    std::vector<std::string> r;
    for (auto &t : tracers_) r.push_back(t.name);
    return r;
  }
  void get_tracer_info(std::string name, std::string &desc, bool &found, bool &positive, bool &adds_mass) const {
    found = false;
    for (auto &t : tracers_) if (t.name == name) { desc = t.desc; found = true; positive = t.positive; adds_mass = t.adds_mass; }
  }
  bool tracer_exists(std::string name) const {
    for (auto &t : tracers_) if (t.name == name) return true;
    return false;
  }
  void finalize() { dm_.finalize(); }

 private:
  struct Tracer { std::string name, desc; bool positive, adds_mass; };
  real xlen_, ylen_;
  int nz_, ny_, nx_, nens_;
  DataManager dm_;
  Options opt_;
  std::vector<Tracer> tracers_;
};

}  // namespace pam_b200

// Dycore::compute_mass(coupler, state, tracers) returns an owning (num_tracers+2, nens) device array of this header's kind
namespace awfl_b200_detail {
template <class T> struct MassArray<pam_b200::DeviceArray<T, 5>> {
  typedef pam_b200::DeviceArray<double, 2> type;
  static type make(int nv, int nens) { return type({nv, nens}); }
};
template <class T> struct MassArray<pam_b200::ArrayView<T, 5>> : MassArray<pam_b200::DeviceArray<double, 5>> {};
}  // namespace awfl_b200_detail
