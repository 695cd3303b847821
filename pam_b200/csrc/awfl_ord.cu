// One translation unit per reconstruction order (-DAWFL_ORD_INST=3|5|7|9): instantiates the order-templated kernels of
// awfl_sweep.cuh and defines the host launchers declared in awfl_types.h.
#include "awfl_sweep.cuh"
#include "awfl_march.cuh"

#ifndef AWFL_ORD_INST
#error "compile with -DAWFL_ORD_INST=3|5|7|9"
#endif

namespace awfl {

// A marching kernel and its dynamic shared-memory ring. The opt-in to > 48 KB of dynamic shared memory and the request
// for the largest shared-memory carve-out are per device and per kernel: OrdLaunch<ORD>::configure() sets them for every
// instantiation when a handle is created on a device (no lazily filled process-wide set: handles on several devices or
// host threads share nothing).
template <class... KArgs>
static void configure_march(void (*kernel)(KArgs...), size_t bytes) {
  cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}
template <class... KArgs, class... Args>
static void launch_march(void (*kernel)(KArgs...), dim3 blocks, size_t bytes, cudaStream_t s, const Args &...args) {
  kernel<<<blocks, MT, bytes, s>>>(args...);
}
#define AWFL_LAUNCH_MARCH(KERNEL, BYTES, ...) launch_march(KERNEL, blocks, BYTES, s, __VA_ARGS__)

// lines_per_warp > 1 (tracer march): adjacent levels for the horizontal marches (first line index, inside blockIdx.z),
// adjacent columns for the vertical march (second line index = blockIdx.y)
static dim3 march_grid(const Geom &g, int dir, int sbits, int labits, int ntr, int lines_per_warp) {
  const unsigned netile = (unsigned)(g.ne + 31) / 32;   // member tiles of the slab
  unsigned lb = dir == 0 ? g.ny : g.nx;
  if (dir == 2) lb = (lb + lines_per_warp - 1) / lines_per_warp;
  return dim3(netile, lb, (unsigned)ntr << (labits + sbits));
}

template <int ORD>
void OrdLaunch<ORD>::march_state(cudaStream_t s, const Geom &g, const Phys &ph, const SweepArgs &a, int dir, int mode, int vt) {
  const dim3 blocks = march_grid(g, dir, a.sbits, a.labits, 1, 1);
  constexpr size_t B = MarchCfg<ORD>::SMEM_STATE;
  // x and y sweeps only ever accumulate (MODE 0); the z sweep is always the last of a stage (MODE 1 or 2). A 2-D run
  // (ny == 1, no y sweep) takes the kernels that leave the reconstruction of v out (RingF in awfl_march.cuh) unless the
  // handle asks for the general ones
  const bool s2d = g.sim2d && !a.general_2d, trq = s2d && a.fused_tracer;
#define AWFL_STATE_CASE(D, M, V)                                                              \
  do {                                                                                        \
    if (trq) AWFL_LAUNCH_MARCH((k_march_state<ORD, D, M, V, true, true>), B, g, ph, a);       \
    else if (s2d) AWFL_LAUNCH_MARCH((k_march_state<ORD, D, M, V, true, false>), B, g, ph, a); \
    else AWFL_LAUNCH_MARCH((k_march_state<ORD, D, M, V, false, false>), B, g, ph, a);         \
  } while (0)
  if (dir == 0) AWFL_STATE_CASE(0, 0, 0);
  else if (dir == 1) AWFL_LAUNCH_MARCH((k_march_state<ORD, 1, 0, 0, false, false>), B, g, ph, a);
  else if (mode == 1) {
    if (vt == 1) AWFL_STATE_CASE(2, 1, 1);
    else if (vt == 3) AWFL_STATE_CASE(2, 1, 3);
    else AWFL_STATE_CASE(2, 1, 2);
  } else {
    if (vt == 1) AWFL_STATE_CASE(2, 2, 1);
    else AWFL_STATE_CASE(2, 2, 2);
  }
#undef AWFL_STATE_CASE
}

template <int ORD>
void OrdLaunch<ORD>::march_tracer(cudaStream_t s, const Geom &g, const TracerSweepArgs &a, int dir, int vt, bool pair) {
  // pair: a.ntr (even) tracers, two per warp on one line; else two adjacent lines of one tracer per warp
  const dim3 blocks = pair ? march_grid(g, dir, a.sbits, a.labits, a.ntr / 2, 1) : march_grid(g, dir, a.sbits, a.labits, a.ntr, 2);
  constexpr size_t B = MarchCfg<ORD>::SMEM_TRACER;
#define AWFL_TRACER_CASE(D, V)                                                       \
  do {                                                                               \
    if (pair) AWFL_LAUNCH_MARCH((k_march_tracer<ORD, D, V, true>), B, g, a);         \
    else AWFL_LAUNCH_MARCH((k_march_tracer<ORD, D, V, false>), B, g, a);             \
  } while (0)
  if (dir == 0) AWFL_TRACER_CASE(0, 0);
  else if (dir == 1) AWFL_TRACER_CASE(1, 0);
  else if (vt == 1) AWFL_TRACER_CASE(2, 1);
  else if (vt == 3) AWFL_TRACER_CASE(2, 3);
  else AWFL_TRACER_CASE(2, 2);
#undef AWFL_TRACER_CASE
}

template <int ORD>
void OrdLaunch<ORD>::configure() {
  constexpr size_t BS = MarchCfg<ORD>::SMEM_STATE, BT = MarchCfg<ORD>::SMEM_TRACER;
#define AWFL_CONFIGURE_2D(D, M, V)                                   \
  configure_march(k_march_state<ORD, D, M, V, false, false>, BS);    \
  configure_march(k_march_state<ORD, D, M, V, true, false>, BS);     \
  configure_march(k_march_state<ORD, D, M, V, true, true>, BS)
  AWFL_CONFIGURE_2D(0, 0, 0);
  configure_march(k_march_state<ORD, 1, 0, 0, false, false>, BS);
  AWFL_CONFIGURE_2D(2, 1, 1);
  AWFL_CONFIGURE_2D(2, 1, 2);
  AWFL_CONFIGURE_2D(2, 1, 3);
  AWFL_CONFIGURE_2D(2, 2, 1);
  AWFL_CONFIGURE_2D(2, 2, 2);
#undef AWFL_CONFIGURE_2D
#define AWFL_CONFIGURE_TR(D, V)                                  \
  configure_march(k_march_tracer<ORD, D, V, false>, BT);         \
  configure_march(k_march_tracer<ORD, D, V, true>, BT)
  AWFL_CONFIGURE_TR(0, 0);
  AWFL_CONFIGURE_TR(1, 0);
  AWFL_CONFIGURE_TR(2, 1);
  AWFL_CONFIGURE_TR(2, 2);
  AWFL_CONFIGURE_TR(2, 3);
#undef AWFL_CONFIGURE_TR
}

template <int ORD>
int OrdLaunch<ORD>::upload_cvtab(const double *host, size_t ndoubles, const int *level_offsets, size_t nlevels) {
  if (ndoubles > (size_t)AWFL_CVTAB_DOUBLES || nlevels > (size_t)AWFL_CVLVL_MAX) return (int)cudaErrorInvalidValue;
  cudaError_t e = cudaMemcpyToSymbol(c_vtab, host, ndoubles * sizeof(double), 0, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_vlvl, level_offsets, nlevels * sizeof(int), 0, cudaMemcpyHostToDevice);
  return (int)e;
}

template <int ORD>
int OrdLaunch<ORD>::upload_cvpage(cudaStream_t s, const double *host_pinned, size_t ndoubles, const int *level_offsets_pinned, size_t nlevels) {
  if (ndoubles > (size_t)AWFL_CVTAB_DOUBLES || nlevels > (size_t)AWFL_CVLVL_MAX) return (int)cudaErrorInvalidValue;
  cudaError_t e = cudaMemcpyToSymbolAsync(c_vtab, host_pinned, ndoubles * sizeof(double), 0, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaMemcpyToSymbolAsync(c_vlvl, level_offsets_pinned, nlevels * sizeof(int), 0, cudaMemcpyHostToDevice, s);
  return (int)e;
}

template <int ORD>
size_t OrdLaunch<ORD>::cvtab_capacity() { return (size_t)AWFL_CVTAB_DOUBLES; }
template <int ORD>
size_t OrdLaunch<ORD>::cvtab_max_levels() { return (size_t)AWFL_CVLVL_MAX; }

template <int ORD>
void OrdLaunch<ORD>::hydro_pint(cudaStream_t s, const Geom &g, const Phys &ph, const HydroArgs &a, bool shared) {
  const long ncol = (long)g.ny * g.nx * g.nens;
  const unsigned blocks = (unsigned)((ncol + 127) / 128);
  if (shared) k_hydro_pint<ORD, true><<<blocks, 128, 0, s>>>(g, ph, a);
  else k_hydro_pint<ORD, false><<<blocks, 128, 0, s>>>(g, ph, a);
}

template <int ORD>
void OrdLaunch<ORD>::reconstruct(cudaStream_t s, const double *st, int n, int ind, int level, int iens, int nens, const double *vtab,
                                 bool shared, double *out) {
  const unsigned blocks = (unsigned)((n + 127) / 128);
  if (shared) k_reconstruct<ORD, true><<<blocks, 128, 0, s>>>(st, n, ind, level, iens, nens, vtab, out);
  else k_reconstruct<ORD, false><<<blocks, 128, 0, s>>>(st, n, ind, level, iens, nens, vtab, out);
}

template struct OrdLaunch<AWFL_ORD_INST>;

}  // namespace awfl
