// Plain structs shared by the per-order kernel translation units and the host side.
#pragma once
#include <cuda_runtime.h>

namespace awfl {

enum { idR = 0, idU = 1, idV = 2, idW = 3, idT = 4, NS = 5 };
static constexpr int MT = 32;  // threads per block of the marching kernels: ONE warp (work item) per block; lane = ensemble member

struct Geom {
  int nz, ny, nx, nens, nt;
  int sim2d;
  long ncell;     // nz*ny*nx*nens
  double dx, dy;
  // Member slab a launch works on: members [e0, e0+ne) of the nens stored ones (e0 a multiple of 32). The whole ensemble
  // is (0, nens); awfl_time_step_host runs the step slab by slab so that host<->device copies of one slab overlap the
  // kernels of another (CRMs are independent: no kernel reads across members).
  int e0, ne;
};
struct Phys {
  double R_d, R_v, C0, gamma, grav;
};

// State march arguments (k_march_state)
struct SweepArgs {
  const double *qin;      // stage input, N fields
  const double *qbase;    // step-start state (MODE 1, a != 0)
  double *qout;           // stage output (MODE 1) or total tendency (MODE 2), 5 fields
  double *tend;           // partial tendencies, 5 fields
  double *mflux;          // interface mass flux of this direction
  const double *dz;       // (nz,nens)
  const double *grav_var; // (nz,nens)
  const double *vtab;     // vertical matrices (global-memory copy)
  const double *rdz;      // 1/dz (nz,nens)
  double *pcache;         // pressure of the stage input, one field: written by the x march, read by the y and z marches
  long vst;               // element stride of vtab: nens (member-dependent grid) or 1 (shared grid)
  double a, b, c;         // SSPRK3 stage coefficients (c = b*dt)
  int first;              // 1: this is the first sweep of the stage (tend is written, not accumulated)
  int sbits, labits;      // grid encoding (decode_item): 2^sbits segments per line, labits bits for the first line index
  int seg_only;           // >= 0: only this segment of every line works, the other blocks leave at once (paged vertical tables)
  double *tflux;          // fused_tracer: interface fluxes of the first tracer in this direction
  int fused_tracer;       // 2-D: the state march carries the first tracer (its flux goes to tflux; the tracer march skips it)
  int general_2d;         // 2-D run on the general kernels, which reconstruct v too (AWFL_GENERAL_2D=1 at create time: A/B and agreement tests)
};

// Tracer march arguments (k_march_tracer)
struct TracerSweepArgs {
  const double *qin;    // stage input, N fields (rho = field 0, tracers from field 5)
  const double *mflux;  // interface mass flux of this direction
  double *tflux;        // nt fields of interface fluxes for this direction
  const double *vtab;
  long vst;             // element stride of vtab: nens or 1
  int sbits, labits;    // grid encoding (decode_item)
  int seg_only;         // see SweepArgs
  int tr0, ntr;         // tracers [tr0, tr0 + ntr) of this launch (gridDim.z = ntr << (labits + sbits) must stay <= 65535)
};

// declare_current_profile_as_hydrostatic arguments
struct HydroArgs {
  const double *q;        // dycore state (rho, ..., rho*theta) or NULL when gcm columns are used
  const double *g_rd, *g_t, *g_rv, *g_rc, *g_ri;
  const double *dz, *vtab;
  double *pint;           // (nz+1,ny,nx,nens)
  double *grav_var;
};

// Per-order launchers, defined in awfl_ord.cu (one translation unit per reconstruction order so they build in parallel).
template <int ORD> struct OrdLaunch {
  // cell-centred marching kernels (awfl_march.cuh); vt: 0 constant horizontal, 1 constant-bank vertical, 2 global vertical
  static void march_state(cudaStream_t s, const Geom &g, const Phys &ph, const SweepArgs &a, int dir, int mode, int vt);
  static void march_tracer(cudaStream_t s, const Geom &g, const TracerSweepArgs &a, int dir, int vt, bool pair);
  static void configure();   // per-device kernel attributes (shared-memory rings) of every marching kernel of this order
  static int upload_cvtab(const double *host, size_t ndoubles, const int *level_offsets, size_t nlevels);   // synchronous; returns a cudaError_t
  // one PAGE of the table (the levels a segment of the vertical marches needs), asynchronously on the launching stream
  static int upload_cvpage(cudaStream_t s, const double *host_pinned, size_t ndoubles, const int *level_offsets_pinned, size_t nlevels);
  static size_t cvtab_capacity();
  static size_t cvtab_max_levels();
  static void hydro_pint(cudaStream_t s, const Geom &g, const Phys &ph, const HydroArgs &a, bool shared);
  static void reconstruct(cudaStream_t s, const double *st, int n, int ind, int level, int iens, int nens, const double *vtab, bool shared, double *out);
};

}  // namespace awfl
