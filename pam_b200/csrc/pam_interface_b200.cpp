// The GCM-facing host registry of PAM, pam_core/pam_interface (pam_interface.h + pam_interface_extern_c.cpp:5-274): the
// extern "C" symbols that pam_fortran_interface.F90 binds, so that an E3SM-MMF build can link libawfl_b200.so in place of
// the reference's pam_interface objects. Host-side only, like the reference ("the pam_interface
// routines only deal with host-side data", pam_interface.h:9-12): named arrays (owned or mirrored from the GCM) and typed
// options, one registry per host thread (pam_interface.h:16-32). Two additions carry a registered host array to / from
// device memory through the library's own copy entry points (awfl_copy_to_device / awfl_copy_to_host).
//
// Error behaviour as pam_core/pam_const.h:249-252: print the message to stderr, throw std::string.
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <typeinfo>
#include <vector>

#include "../../include/awfl_b200.h"
#include "../../include/pam_interface_b200.h"

namespace {

[[noreturn]] void endrun(const std::string &msg) {
  std::fprintf(stderr, "%s\n", msg.c_str());
  throw std::string(msg);
}

struct Entry {
  std::string desc;
  std::vector<int> dims;
  size_t type;          // typeid hash of the element type
  size_t elem;          // element size in bytes
  void *ptr;
  bool owned, readonly;
  size_t count() const { size_t n = 1; for (int d : dims) n *= (size_t)d; return n; }
};
struct Option { size_t type; unsigned char raw[16]; std::string str; };
struct Registry {
  std::map<std::string, Entry> arrays;
  std::map<std::string, Option> options;
  std::map<std::string, int> dimensions;
  ~Registry() { for (auto &kv : arrays) if (kv.second.owned) std::free(kv.second.ptr); }
};

std::mutex g_mutex;
std::map<std::thread::id, Registry> g_registries;    // pam_interface.h:16 (one coupler per host thread)

Registry &reg() {
  std::lock_guard<std::mutex> lock(g_mutex);
  return g_registries[std::this_thread::get_id()];
}

template <class T>
void add_array(const char *name, const char *desc, const int *dims, int ndims, T *ptr, bool readonly) {
  if (ndims < 1 || ndims > 7) endrun("ERROR: arrays must have between 1 and 7 dimensions: " + std::string(name));
  Registry &r = reg();
  if (r.arrays.count(name)) endrun("ERROR: Duplicate entry name: " + std::string(name));                 // DataManager.h:104-108
  Entry e;
  e.desc = desc ? desc : ""; e.dims.assign(dims, dims + ndims); e.type = typeid(T).hash_code(); e.elem = sizeof(T);
  e.readonly = readonly; e.owned = ptr == nullptr;
  for (int d : e.dims) if (d < 1) endrun("ERROR: non-positive dimension for " + std::string(name));
  e.ptr = ptr ? (void *)ptr : std::calloc(e.count(), sizeof(T));
  if (!e.ptr) endrun("ERROR: allocation failed for " + std::string(name));
  r.arrays[name] = e;
}

template <class T>
void get_array(const char *name, T *&ptr, int *dims, int ndims) {
  Registry &r = reg();
  auto it = r.arrays.find(name);
  if (it == r.arrays.end()) endrun("ERROR: Could not find entry: " + std::string(name));                  // DataManager.h:553-560
  const Entry &e = it->second;
  if (e.type != typeid(T).hash_code()) endrun("ERROR: Requested type is different from entry type for " + std::string(name));
  if ((int)e.dims.size() != ndims) endrun("ERROR: Requested dimensions is different from entry dimensions for " + std::string(name));
  for (int i = 0; i < ndims; i++) dims[i] = e.dims[i];
  ptr = (T *)e.ptr;
}

template <class T>
void set_option(const char *name, T v) {
  Option o; o.type = typeid(T).hash_code(); std::memset(o.raw, 0, sizeof(o.raw));
  static_assert(sizeof(T) <= 16, "option type too large");
  std::memcpy(o.raw, &v, sizeof(T));
  reg().options[name] = o;
}
template <class T>
T get_option(const char *name) {
  Registry &r = reg();
  auto it = r.options.find(name);
  if (it == r.options.end()) endrun("ERROR: option key does not exist: " + std::string(name));            // Options.h:84-87
  if (it->second.type != typeid(T).hash_code()) endrun("ERROR: Requesting option with the wrong type: " + std::string(name));
  T v; std::memcpy(&v, it->second.raw, sizeof(T));
  return v;
}
const std::string &get_option_string(const char *name) {
  Registry &r = reg();
  auto it = r.options.find(name);
  if (it == r.options.end()) endrun("ERROR: option key does not exist: " + std::string(name));
  if (it->second.type != typeid(std::string).hash_code()) endrun("ERROR: Requesting option with the wrong type: " + std::string(name));
  return it->second.str;
}

}  // namespace

#define PAMI_TYPED(SUFFIX, T)                                                                                                      \
  extern "C" void pam_interface_mirror_array_readonly_##SUFFIX(char const *name, char const *desc, int *dims, int ndims, T *ptr) {  \
    if (!ptr) endrun("ERROR: null pointer mirrored as " + std::string(name));                                                       \
    add_array<T>(name, desc, dims, ndims, ptr, true);                                                                               \
  }                                                                                                                                 \
  extern "C" void pam_interface_mirror_array_readwrite_##SUFFIX(char const *name, char const *desc, int *dims, int ndims, T *ptr) { \
    if (!ptr) endrun("ERROR: null pointer mirrored as " + std::string(name));                                                       \
    add_array<T>(name, desc, dims, ndims, ptr, false);                                                                              \
  }                                                                                                                                 \
  extern "C" void pam_interface_create_array_##SUFFIX(char const *name, char const *desc, int *dims, int ndims) {                   \
    add_array<T>(name, desc, dims, ndims, (T *)nullptr, false);                                                                     \
  }                                                                                                                                 \
  extern "C" void pam_interface_get_array_##SUFFIX(char const *name, T *&ptr, int *dims, int ndims) { get_array<T>(name, ptr, dims, ndims); } \
  extern "C" void pam_interface_set_option_##SUFFIX(char const *name, T value) { set_option<T>(name, value); }                      \
  extern "C" void pam_interface_get_option_##SUFFIX(char const *name, T &value) { value = get_option<T>(name); }

PAMI_TYPED(bool, bool)
PAMI_TYPED(int, int)
PAMI_TYPED(float, float)
PAMI_TYPED(double, double)

extern "C" void pam_interface_finalize() {                               // pam_interface.h:21 couplers.clear()
  std::lock_guard<std::mutex> lock(g_mutex);
  g_registries.clear();
}
extern "C" void pam_interface_set_option_string(char const *name, char const *value) {
  Option o; o.type = typeid(std::string).hash_code(); std::memset(o.raw, 0, sizeof(o.raw)); o.str = value;
  reg().options[name] = o;
}
extern "C" void pam_interface_get_option_stringlen(char const *name, int &len) { len = (int)get_option_string(name).size(); }
extern "C" void pam_interface_get_option_string(char const *name, char *value) {
  const std::string &s = get_option_string(name);
  for (size_t i = 0; i < s.size(); i++) value[i] = s[i];                 // no terminator: the Fortran side knows the length
}
extern "C" void pam_interface_option_exists(char const *name, bool &exists) { exists = reg().options.count(name) > 0; }
extern "C" void pam_interface_remove_option(char const *name) { reg().options.erase(name); }
extern "C" void pam_interface_make_readonly(char const *name) {
  Registry &r = reg();
  auto it = r.arrays.find(name);
  if (it == r.arrays.end()) endrun("ERROR: Could not find entry: " + std::string(name));
  it->second.readonly = true;
}
extern "C" void pam_interface_destroy_array(char const *name) {
  Registry &r = reg();
  auto it = r.arrays.find(name);
  if (it == r.arrays.end()) return;                                      // DataManager::unregister_and_deallocate ignores unknown names
  if (it->second.owned) std::free(it->second.ptr);
  r.arrays.erase(it);
}
extern "C" void pam_interface_array_exists(char const *name, bool &exists) { exists = reg().arrays.count(name) > 0; }
extern "C" void pam_interface_register_dimension(char const *name, int len) {
  Registry &r = reg();
  auto it = r.dimensions.find(name);
  if (it != r.dimensions.end() && it->second != len) endrun("ERROR: dimension already exists with another length: " + std::string(name));
  r.dimensions[name] = len;
}

// ---- host mirror <-> device (B200 additions) ----
extern "C" int pam_interface_b200_array_bytes(char const *name, unsigned long *bytes) {
  Registry &r = reg();
  auto it = r.arrays.find(name);
  if (it == r.arrays.end()) return 1;
  *bytes = (unsigned long)(it->second.count() * it->second.elem);
  return 0;
}
extern "C" int pam_interface_b200_copy_to_device(char const *name, void *device_ptr) {
  Registry &r = reg();
  auto it = r.arrays.find(name);
  if (it == r.arrays.end() || !device_ptr) return 1;
  return awfl_copy_to_device(device_ptr, it->second.ptr, (unsigned long)(it->second.count() * it->second.elem));
}
extern "C" int pam_interface_b200_copy_from_device(char const *name, const void *device_ptr) {
  Registry &r = reg();
  auto it = r.arrays.find(name);
  if (it == r.arrays.end() || !device_ptr) return 1;
  if (it->second.readonly) endrun("ERROR: Requesting a read-only entry for writing: " + std::string(name));
  return awfl_copy_to_host(it->second.ptr, device_ptr, (unsigned long)(it->second.count() * it->second.elem));
}
