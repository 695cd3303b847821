/* awfl_b200.h -- C ABI of the B200-native AWFL dycore time step (libawfl_b200.so).
 *
 * Drop-in boundary for ONE hot path of E3SM-Project/PAM: the AWFL dynamical-core time step
 * (reference: dynamics/awfl/Dycore.h, class Dycore). Every entry point below replaces one public method
 * of the reference's `class Dycore` (header-only C++ over YAKL parallel_for); the C++ shim in
 * pam_b200/cxx/Dycore.h keeps the reference's method names and pulls raw device pointers out of the
 * coupler's DataManager to call these. No torch / YAKL types cross this boundary: plain pointers and sizes.
 *
 * Conventions
 *   - all field pointers are DEVICE pointers to FP64 arrays of shape (nz,ny,nx,nens), C order, nens fastest
 *     (pam_core/pam_const.h:30-37, pam_core/pam_coupler.h:255-263) unless a parameter says "host";
 *   - every function returns 0 on success, non-zero on error; awfl_last_error() gives the message
 *     (the reference throws std::string through endrun(), pam_core/pam_const.h:249-252; the shim converts);
 *   - a handle is bound to one device and one stream; ONE handle must not be used from two host threads at once,
 *     but different handles are independent: any number of them may live in one process, on the same or on different
 *     devices, and be driven from different host threads concurrently (one coupler per host thread in the reference,
 *     pam_core/pam_interface/pam_interface.h:16-32). Every entry point makes the handle's device current for the
 *     duration of the call and restores the caller's;
 *   - there is no CPU fallback: every entry point launches sm_100a kernels or fails.
 */
#ifndef AWFL_B200_H
#define AWFL_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct awfl_handle awfl_handle;

/* Options resolved once by Dycore::init (dynamics/awfl/Dycore.h:866-981) + grid (pam_coupler.h:184-202). */
typedef struct {
  int nz, ny, nx, nens;          /* coupler.get_nz/ny/nx/nens                                        */
  int num_tracers;               /* coupler.get_num_tracers()                                        */
  int ord;                       /* reconstruction order: 3,5,7,9 (the reference's compile-time MW_ORD, Dycore.h:17-21) */
  double xlen, ylen;             /* domain lengths [m]; dx = xlen/nx, dy = ylen/ny                   */
  double R_d, cp_d, R_v, cp_v, p0, grav; /* Dycore.h:871-876 (defaults 287,1003,461,1859,1e5,9.81)   */
  double gamma_d, kappa_d, C0;   /* Dycore.h:885-891; pass 0 to have them derived from the above     */
  int idWV;                      /* index of "water_vapor" among the tracers (Dycore.h:974)          */
  const unsigned char *tracer_positive;  /* host, num_tracers flags (Dycore.h:968)                   */
  const unsigned char *tracer_adds_mass; /* host, num_tracers flags (Dycore.h:969)                   */
  const double *dz_host;         /* host copy of "vertical_cell_dz" (nz,nens)                        */
  int device;                    /* CUDA device ordinal                                              */
  void *stream;                  /* cudaStream_t to enqueue on (NULL = the legacy default stream)       */
} awfl_config;

/* The coupler's prognostic fields (pam_coupler.h:259-263 + every registered tracer). */
typedef struct {
  double *density_dry, *uvel, *vvel, *wvel, *temp;
  double *const *tracers;        /* host array of num_tracers device pointers                        */
} awfl_fields;

/* Optional large-scale columns for declare_hydrostatic(use_gcm_data=true) (Dycore.h:1416-1420), each (nz,nens) */
typedef struct {
  const double *gcm_density_dry, *gcm_temp, *gcm_water_vapor, *gcm_cloud_water, *gcm_cloud_ice;
} awfl_gcm_columns;

/* Dycore::init (Dycore.h:835-1091): resolves constants, builds the per-level vertical WENO matrices on the host
 * by the reference's algorithm (Dycore.h:897-939, TransformMatrices_variable.h:10-69), uploads them, allocates
 * the persistent workspace (no allocation happens inside a time step). */
int awfl_create(const awfl_config *cfg, awfl_handle **out);

/* Dycore::finalize (Dycore.h:1548) + destructor. Never frees caller-owned fields. */
void awfl_destroy(awfl_handle *h);

/* Message of the last failing call on this handle (or of awfl_create when h == NULL). */
const char *awfl_last_error(const awfl_handle *h);

/* Dycore::dycore_name (Dycore.h:1544). */
const char *awfl_dycore_name(void);

/* Dycore::timeStep (Dycore.h:107-255): coupler -> dycore state, clip tracers, CFL dt, ncycles SSPRK3 sub-cycles,
 * dycore state -> coupler. Fields are updated in place. ncycles_override > 0 fixes the sub-cycle count and
 * removes the only host sync (the 8-byte dt read-back); ncycles_out (may be NULL) receives the count used. */
int awfl_time_step(awfl_handle *h, const awfl_fields *f, double crm_dt, int ncycles_override, int *ncycles_out);

/* Same step through HOST buffers (pageable or pinned): copies the 5+num_tracers fields to the device, steps,
 * copies them back -- the end-to-end path of a caller whose coupler lives in host memory. host_fields[0..4] =
 * density_dry,uvel,vvel,wvel,temp, then the tracers. */
int awfl_time_step_host(awfl_handle *h, double *const *host_fields, double crm_dt, int ncycles_override, int *ncycles_out);

/* Number of member slabs awfl_time_step_host pipelines (copy of one slab overlaps the kernels of another; CRMs are
 * independent, so results do not depend on it). 0 (default, or the environment variable AWFL_HOST_SLABS) = chosen from
 * the grid so that the slabs in flight fill the GPU once; 1 = copy everything, step, copy everything back.
 * awfl_host_slabs returns the count the next call will use. */
int awfl_set_host_slabs(awfl_handle *h, int nslabs);
int awfl_host_slabs(const awfl_handle *h);

/* awfl_time_step_host starts its slabs with a predicted sub-cycle count (the previous call's) before the CFL dt of the
 * whole ensemble is known and redoes the step, un-overlapped, when the prediction was wrong: how often that happened. */
long awfl_host_mispredictions(const awfl_handle *h);

/* Dycore::compute_time_step (Dycore.h:65-102): min over cells of cfl*d/(|vel|+cs); result written to *dt_host. */
int awfl_compute_time_step(awfl_handle *h, const awfl_fields *f, double cfl, double *dt_host);

/* Dycore::compute_tendencies (Dycore.h:262-586) on UN-haloed conserved arrays, for tendency-level parity tests.
 * state (5,nz,ny,nx,nens) = rho, rho*u, rho*v, rho*w, rho*theta; tracers / fct_start (num_tracers,...);
 * outputs state_tend (5,...) and tracers_tend (num_tracers,...). All device pointers. Inputs are not modified. */
int awfl_compute_tendencies(awfl_handle *h, const double *state, const double *tracers, const double *fct_start,
                            double dt, double *state_tend, double *tracers_tend);

/* ---- The reference's HALOED dynamics arrays (the arguments of the other public Dycore methods) ----
 * state   (5, nz+2hs, ny+2hs, nx+2hs, nens) = rho, rho*u, rho*v, rho*w, rho*theta     hs = (ord+1)/2
 * tracers (num_tracers, nz+2hs, ny+2hs, nx+2hs, nens)
 * pressure (nz+2hs, ny+2hs, nx+2hs, nens);  state_tend (5,nz,ny,nx,nens), tracers_tend (num_tracers,nz,ny,nx,nens)
 * All DEVICE pointers, C order, nens fastest -- the layout of the reference's real5d / real4d arrays. The time step
 * itself never builds these arrays; the entry points below give a caller who holds them the reference's methods with
 * the reference's effects on that memory. Work is enqueued on the handle's stream (no host synchronisation). */

/* Dycore::convert_coupler_to_dynamics (Dycore.h:1336-1388): coupler fields -> interior of state / tracers. */
int awfl_convert_coupler_to_dynamics(awfl_handle *h, const awfl_fields *f, double *state, double *tracers);

/* Dycore::convert_dynamics_to_coupler (Dycore.h:1281-1331): interior of state / tracers -> coupler fields. */
int awfl_convert_dynamics_to_coupler(awfl_handle *h, const double *state, const double *tracers, const awfl_fields *f);

/* Dycore::halo_exchange (Dycore.h:608-711) on arrays whose velocities, theta and tracers are already divided by density:
 * periodic x (and, in 3-D, y) ghosts, wall ghosts in z (w = 0, the others zero-gradient, density by hydrostatic
 * extrapolation at constant theta, pressure from it). Ghost theta is the boundary cell's (the intended value where the
 * reference has a read/write race). Corner ghosts are not touched, as in the reference. */
int awfl_halo_exchange(awfl_handle *h, double *state, double *tracers, double *pressure);

/* Dycore::compute_tendencies (Dycore.h:262-586) with the reference's own signature: tracers_tend holds the FCT start
 * mass on entry (Dycore.h:158,173,197) and the tracer tendencies on return; state / tracers come back as the reference
 * leaves them -- interior divided by density and multiplied back ((x/rho)*rho, Dycore.h:316-320,527-532: the SSPRK3
 * update consumes exactly that), ghosts filled as by halo_exchange. Uses the handle's variable_gravity. */
int awfl_compute_tendencies_haloed(awfl_handle *h, double *state, double *state_tend, double *tracers, double *tracers_tend, double dt);

/* Dycore::compute_mass (Dycore.h:37-60) with the reference's arguments: out_device (num_tracers+2, nens) DEVICE array
 * (the reference returns a device real2d). */
int awfl_compute_mass_dynamics(awfl_handle *h, const double *state, const double *tracers, double *out_device);

/* awfl_compute_mass_dynamics without a handle and WITHOUT ANY ALLOCATION. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it). Calls no cudaMalloc / cudaFree / allocating library routine and does not synchronise:
 * two kernels go to `stream` (a cudaStream_t, NULL = the default stream).
 *   state (5,nz+2hs,ny+2hs,nx+2hs,nens), tracers (num_tracers,...): DEVICE haloed arrays as above, hs = (ord+1)/2;
 *   dz (nz,nens): DEVICE cell heights; out_device (num_tracers+2,nens): DEVICE result, same values as
 *   awfl_compute_mass_dynamics (same kernels, same summation order).
 * The de-haloed copy of the fields goes to a static device array of AWFL_NOALLOC_MAX_DOUBLES doubles (16 MiB) that every
 * GPU which loads the library holds until the process exits: (5 + num_tracers) * nz * ny * nx * nens beyond that is
 * rejected (return 1, text in awfl_last_error(NULL)). The arrays must live on the CURRENT device. All calls on a device
 * share the static array: calls that overlap on the GPU (different streams without a dependency) race; calls on one
 * stream, or otherwise ordered, do not. Nothing in it has to start at zero.
 * awfl_noalloc_setup() resolves the array's device address for the current device (done by the first call otherwise):
 * call it once per device before a graph capture begins. */
#define AWFL_NOALLOC_MAX_DOUBLES (1ul << 21)
int awfl_noalloc_setup(void);
unsigned long awfl_noalloc_max_doubles(void);
int awfl_compute_mass_dynamics_noalloc(int nz, int ny, int nx, int nens, int num_tracers, int ord, const double *state,
                                       const double *tracers, const double *dz, double *out_device, void *stream);

/* Dycore::init_supercell (Dycore.h:1096-1276): the supercell sounding as interior values of state / tracers.
 * zint: DEVICE "vertical_interface_height" (nz+1,nens). Init-time call: allocates scratch and synchronises. */
int awfl_init_supercell(awfl_handle *h, double *state, double *tracers, const double *zint);

/* Dycore::declare_current_profile_as_hydrostatic (Dycore.h:1392-1504), grav_balance branch: fills the handle's
 * "variable_gravity" (nz,nens). gcm == NULL <=> use_gcm_data == false. Deterministic (no atomics). */
int awfl_declare_hydrostatic(awfl_handle *h, const awfl_fields *f, const awfl_gcm_columns *gcm);

/* The idealized initial states of Dycore::init (Dycore.h:985-1091, chosen there by "initData" of the standalone
 * YAML file): data_spec 0 = rising thermal (:1017-1086), 1 = supercell sounding (Dycore::init_supercell,
 * :1096-1276), each followed by convert_dynamics_to_coupler (:1281-1331): the coupler fields are overwritten.
 * zint: DEVICE "vertical_interface_height" (nz+1,nens). Init-time call: allocates scratch and synchronises. */
int awfl_init_idealized(awfl_handle *h, const awfl_fields *f, const double *zint, int data_spec);

/* Dycore::compute_mass (Dycore.h:37-60): out_host[(num_tracers+2)*nens] = mean over cells of q*dz for each tracer,
 * rho and rho*theta. */
int awfl_compute_mass(awfl_handle *h, const awfl_fields *f, double *out_host);

/* Dycore-owned DataManager entries other modules may read (Dycore.h:868,897-898): device pointers owned by the
 * handle. variable_gravity (nz,nens); vert_sten_to_coefs (nz+2,ord,ord,nens); vert_weno_recon_lower
 * (nz+2,hs,hs,hs,nens) with hs=(ord+1)/2. */
double *awfl_variable_gravity(awfl_handle *h);
const double *awfl_vert_sten_to_coefs(awfl_handle *h);
const double *awfl_vert_weno_recon_lower(awfl_handle *h);

/* The shim keeps the DataManager entries as the source of truth, like the reference (which fetches "variable_gravity"
 * from the DataManager in every compute_tendencies call, Dycore.h:277): awfl_set_variable_gravity copies a caller-owned
 * device array (nz,nens) into the handle before a step; awfl_export_tables copies the handle's variable_gravity,
 * vert_sten_to_coefs and vert_weno_recon_lower into caller-owned device arrays (any pointer may be NULL). Both are
 * device-to-device copies on the handle's stream. */
int awfl_set_variable_gravity(awfl_handle *h, const double *variable_gravity);
int awfl_export_tables(awfl_handle *h, double *variable_gravity, double *vert_sten_to_coefs, double *vert_weno_recon_lower);

/* Small runtime helpers so that a host shim needs no CUDA header of its own: synchronous device-to-host /
 * host-to-device copies (used at init for the (nz,nens) vertical grid and by tests) and the current device ordinal. */
int awfl_copy_to_host(void *dst_host, const void *src_device, unsigned long bytes);
int awfl_copy_to_device(void *dst_device, const void *src_host, unsigned long bytes);
int awfl_current_device(void);
int awfl_device_alloc(void **ptr, unsigned long bytes);   /* zero-filled device memory */
void awfl_device_free(void *ptr);

/* Resolved option value by name ("R_d","cp_d","R_v","cp_v","p0","grav","cv_d","gamma_d","kappa_d","C0"); NaN if unknown. */
double awfl_get_option(const awfl_handle *h, const char *name);

/* Table-level test hook for Dycore::reconstruct (Dycore.h:591-604) on the device: n stencils of `ord` values
 * (device), ind = 0 left edge / 1 right edge. level < 0: constant horizontal matrices; else the vertical matrices
 * of (level, iens). Static: needs no handle state besides the tables. */
int awfl_reconstruct(awfl_handle *h, const double *stencils, int n, int ind, int level, int iens, double *out);

/* Where the z marches of this handle read their per-level vertical matrices: 1 = the constant bank of (device, ord),
 * claimed at awfl_create and shared by handles whose tables are bit-identical; 3 = a global-memory table with unit
 * stride (shared vertical grid, but the constant bank is held by a handle with another grid, or the table does not fit);
 * 2 = a member-strided global table (member-dependent grid); 4 = the constant bank used page by page (a shared grid whose
 * table does not fit as a whole, e.g. ord 9 at 58 levels: the z marches run segment by segment behind the upload of the
 * levels each segment reads; needs the bank for itself). Results do not depend on it (same arithmetic). */
int awfl_table_home(const awfl_handle *h);

/* Diagnostic: which state marches of this handle carry the first tracer (2-D runs: its flux, the tracer part of
 * Dycore.h:381-385 / :513-517, then comes out of the state march and the tracer march runs for the other tracers only).
 * Bit 0 = x march, bit 1 = z march; 0 for 3-D handles and handles without tracers. Environment AWFL_FUSED_TRACER at create
 * time overrides the per-order default. */
int awfl_fused_tracer(const awfl_handle *h);

/* Debug aid: fills every scratch buffer of the handle with NaN bit patterns (on the handle's stream). A later call that
 * read scratch memory before writing it returns NaN -- the check compute-sanitizer's initcheck would make. */
int awfl_poison_workspace(awfl_handle *h);

/* Number of kernel launches issued by this handle since creation (bench.py reports the per-step count). */
long awfl_launch_count(const awfl_handle *h);

/* Per-kernel-class device timing with CUDA events on the handle's stream. Between begin and end every kernel launch is
 * bracketed by an event pair; end synchronises and returns, per class, the summed milliseconds and the launch count.
 * Classes (AWFL_KC_COUNT entries): */
enum { AWFL_KC_IN = 0, AWFL_KC_OUT, AWFL_KC_STATE_X, AWFL_KC_STATE_Y, AWFL_KC_STATE_Z, AWFL_KC_TRACER_X, AWFL_KC_TRACER_Y,
       AWFL_KC_TRACER_Z, AWFL_KC_FCT, AWFL_KC_TRACER_UPD, AWFL_KC_OTHER, AWFL_KC_COUNT };
int awfl_profile_begin(awfl_handle *h);
int awfl_profile_end(awfl_handle *h, double *ms_by_class, long *launches_by_class);

/* FP64 FMA-pipe microbenchmark: runs `iters` dependent-chain DFMA rounds on every SM and returns the achieved
 * TFLOP/s (2 flops per FMA) in *tflops -- the measured denominator of the FP64 roofline. */
int awfl_measure_fp64_peak(awfl_handle *h, int iters, double *tflops);

#ifdef __cplusplus
}
#endif
#endif
