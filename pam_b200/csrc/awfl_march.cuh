// Cell-centred marching sweep kernels of the AWFL time step (sm_100a). Included only by awfl_ord.cu.
//
// One warp = 32 consecutive ensemble members of one line segment (every global access is a coalesced 256-byte line);
// the warp marches cell by cell along the sweep direction. The stencil window of each thread lives in a shared-memory
// ring (one private column per thread: no bank conflicts, no barriers), which keeps the register footprint small enough
// for 16-24 resident warps per SM -- the FP64 pipe, not latency, is then the limit (profiles/r01a: the register-window
// kernels ran 8 warps/SM at 51 % / 33 % FP64-pipe utilisation, stalled on fixed-latency dependencies).
//
// Cell-centred formulation of Dycore.h:334-519: per marched cell c ONE set of WENO polynomials is built from the window
// c-hs+1..c+hs-1 -- the acoustic pair (rho*u_n, p) and the four advected quantities (u, v, w, theta). Interface c takes
// its left states from the saved right-edge values of cell c-1 and its right states from the left-edge values of cell
// c; the upwind advective value is a select between the two (Dycore.h:368). All six reconstructions of a step use the
// matrices of the same level, so the vertical per-level matrices are warp-uniform constant-bank loads.
#pragma once
#include <math.h>
#include "awfl_types.h"
#include "awfl_weno.cuh"

namespace awfl {

// Ring fields in marching order: the acoustic pair first, then the advected quantities, then density. The RECONSTRUCTED
// fields are contiguous from 0. In a 2-D run (S2D) v is not one of them: the reference zeroes the v tendency there
// (Dycore.h:565), so the flux of v -- one of the six reconstructions per cell and direction -- never reaches a result.
// v then sits behind the reconstructed fields, only for the stage combination of its round-tripped value. The slot v
// leaves free in the pairs / triples of reconstructions can carry the mixing ratio of the FIRST TRACER (TRQ): its
// interface flux (the tracer part of Dycore.h:381-385 / :513-517) then comes out of the state march and the tracer march
// runs for the other tracers only.
enum { RF_COUNT = 8 };
template <bool S2D, bool TRQ> struct RingF {
  static_assert(S2D || !TRQ, "the state march carries a tracer only in its 2-D form");
  static constexpr int M = 0, P = 1, U = 2, W = S2D ? 3 : 4, T = S2D ? 4 : 5, Q = 5, V = S2D ? (TRQ ? 6 : 5) : 3, R = 7;
  static constexpr int NVS = S2D ? 5 : 6;                    // reconstructed state quantities: M, P, U, [V,] W, T
  static constexpr int NVAR = NVS + (TRQ ? 1 : 0);           // ... + the first tracer's mixing ratio
};

#ifndef AWFL_MINB5
#define AWFL_MINB5 12
#endif
#ifndef AWFL_MINB5_Z
#define AWFL_MINB5_Z 16
#endif
#ifndef AWFL_MINB5_TR
#define AWFL_MINB5_TR 16
#endif
// Loop over the variable pairs of a cell: unrolled in the horizontal marches (three call sites of weno_ring; the variable
// index becomes a compile-time constant, so the per-variable branches, the field offsets and the argument set-up fold:
// x march 6.78 -> 6.34 ms), rolled in the vertical march (unrolled with the inlined reconstruction the three copies share
// their table loads and spill 0.9 KB; unrolled around a call 8.88 vs 7.33 ms)
#ifndef AWFL_V0_UNROLL_H
#define AWFL_V0_UNROLL_H 3
#endif
#ifndef AWFL_V0_UNROLL_Z
#define AWFL_V0_UNROLL_Z 1
#endif
#ifndef AWFL_NV
#define AWFL_NV 2      // reconstructions evaluated together in the vertical state march (and everywhere above ord 5)
#endif
#ifndef AWFL_NV_H
#define AWFL_NV_H 3    // ... in the horizontal state marches up to ord 5 (two calls per cell instead of three: x 6.33 -> 6.20 ms;
                       // three at a time spill in the vertical march: 9.55 vs 7.33 ms)
#endif

template <int ORD> struct MarchCfg {
  // resident one-warp blocks per SM
  static constexpr int MINB = ORD == 3 ? 16 : (ORD == 5 ? AWFL_MINB5 : 8);      // horizontal marches
  static constexpr int MINB_Z = ORD == 3 ? 16 : (ORD == 5 ? AWFL_MINB5_Z : 8);  // vertical march (measured: ord 5 prefers 16 x 128 registers)
  static constexpr int MINB_TR = ORD == 3 ? 20 : (ORD == 5 ? AWFL_MINB5_TR : 8);
  // per thread: the window ring (RF_COUNT fields x ORD cells) + the previous interface's 5 fluxes + 4 right-edge values
  // + a staging area for the read-modify-write operands (partial tendency, base state) of the group being reconstructed
  static constexpr int COLS_STATE = RF_COUNT * ORD + NS + 4 + 2 * NS;   // staging: (tend, base) of the 5 state variables
  static constexpr size_t SMEM_STATE = (size_t)COLS_STATE * MT * sizeof(double);
  static constexpr size_t SMEM_TRACER = (size_t)(2 * (ORD + 3)) * MT * sizeof(double);   // two lines per warp
};

// x^g for x > 0 (normal range): exp(g*log(x)). Relative error <~ 4e-15 for the arguments met here (|g log x| < 40),
// against 94 FP64-pipe instructions for the correctly rounded pow(): 49. Stated in DESIGN.md (tolerance budget).
// (CUDA's exp and log carry their coefficients as immediates -- two UMOV each, a fifth of the x march's non-FP64
// instructions. A hand-written pair with constant-bank coefficients removed those UMOVs and was SLOWER, inlined (spills) or
// called, plain or Estrin-form polynomial: x march 5.96 -> 6.33 ... 7.68 ms, profiles/r02i_pow_variants.txt,
// r02j_pow_variants.txt. The library pair's dependency chains are shorter; the UMOVs ride the uniform datapath.)
__device__ __forceinline__ double pow_pos(double x, double g) { return exp(g * log(x)); }

struct Raw { double r, ru, rv, rw, rt, p, tq; };

// 8-byte asynchronous global -> shared copy (LDGSTS). Unlike a load into a register it is not tracked by the register
// scoreboard, so it stays in flight across the call to weno_ring (a CALL waits for every pending register load:
// profiles/r01c measured 28 % of the z march's stall samples there).
__device__ __forceinline__ void cp_async8(double *smem_dst, const double *gsrc) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// wait until at most `left` of the most recently committed groups are still in flight (left <= 2)
__device__ __forceinline__ void cp_async_wait_left(int left) {
  if (left >= 2) asm volatile("cp.async.wait_group 2;" ::: "memory");
  else if (left == 1) asm volatile("cp.async.wait_group 1;" ::: "memory");
  else asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// NV reconstructions on consecutive ring columns. A real function call on purpose: inlined into the march loops, ptxas
// hoists the ~80 loop-invariant constant-bank loads of the WENO tables out of the loop, runs out of uniform registers and
// spills them (measured: 500 B of spills in the tracer march). Behind a call boundary the tables are read where they
// are used.
template <int NV> struct EOn { double E[NV], O[NV]; };
template <int ORD> __device__ __forceinline__ int ring_slot(int head, int s) {
  const int t = head + s;
  return (t >= ORD ? t - ORD : t) * MT;
}
template <int ORD, int NV, class M>
__device__ __noinline__ EOn<NV> weno_ring(const double *fb, int head, M m) {
  double u[NV][ORD];
#pragma unroll
  for (int v = 0; v < NV; v++) {
#pragma unroll
    for (int s = 0; s < ORD; s++) u[v][s] = fb[(v * ORD) * MT + ring_slot<ORD>(head, s)];
  }
  EOn<NV> r;
  weno_EO_n<ORD, NV>(u, m, r.E, r.O);
  return r;
}

// Inlined variant for the tracer march with per-level matrices: their addresses change every step, so nothing can be
// hoisted, and inlined the level index stays in a uniform register (behind the call it arrives in a vector register and
// every matrix element becomes a vector LDC: profiles/r01c, 13 % of the z tracer march's stall samples).
template <int ORD, int NV, class M>
__device__ __forceinline__ EOn<NV> weno_ring_inl(const double *fb, int head, M m) {
  double u[NV][ORD];
#pragma unroll
  for (int v = 0; v < NV; v++) {
#pragma unroll
    for (int s = 0; s < ORD; s++) u[v][s] = fb[(v * ORD) * MT + ring_slot<ORD>(head, s)];
  }
  EOn<NV> r;
  weno_EO_n<ORD, NV>(u, m, r.E, r.O);
  return r;
}

// Slow path of the tracer march: one reconstruction with the overflow-proof weights (weno_EO_n<.., SAFE = true>), taken
// when the fast form returned a non-finite value. Never inlined: it must not cost the fast path a register.
template <int ORD, class M>
__device__ __noinline__ EOn<1> weno_ring_safe(const double *fb, int head, M m) {
  double u[1][ORD];
#pragma unroll
  for (int s = 0; s < ORD; s++) u[0][s] = fb[ring_slot<ORD>(head, s)];
  EOn<1> r;
  weno_EO_n<ORD, 1, M, true>(u, m, r.E, r.O);
  return r;
}

// Matrix provider of a step: VT 0 = constant horizontal tables, 1 = vertical tables in the constant bank (all members
// share one vertical grid), 2 = vertical tables in global memory with a run-time element stride (member-dependent grid),
// 3 = in global memory with unit stride (shared grid whose table is too large for the constant bank).
template <int ORD, int VT> struct MatSel;
template <int ORD> struct MatSel<ORD, 0> {
  using type = HMat<ORD>;
  static __device__ __forceinline__ type get(const double *, long, long, int) { return type(); }
};
template <int ORD> struct MatSel<ORD, 1> {
  using type = CMat<ORD>;
  static __device__ __forceinline__ type get(const double *, long, long, int level) { type m; m.base = c_vlvl[level]; return m; }
};
template <int ORD> struct MatSel<ORD, 2> {
  using type = VMat<ORD>;
  static __device__ __forceinline__ type get(const double *vtab, long vst, long voff, int level) {
    type m; m.p = vtab + (long)level * VMat<ORD>::NEL * vst + voff; m.st = vst; return m;
  }
};

template <int ORD> struct MatSel<ORD, 3> {
  using type = UMat<ORD>;
  static __device__ __forceinline__ type get(const double *vtab, long, long, int level) {
    type m; m.p = vtab + (long)level * UMat<ORD>::NEL; return m;
  }
};

// ---------------------------------------------------------------------------------------------------------------
// State march: Dycore.h:310-321 (primitives, pressure), :629-710 (periodic wrap / wall ghost cells), :334-519 (one
// direction of the flux kernel), :553-566 (flux divergence, gravity) and, on the last sweep of a stage, the SSPRK3
// stage combination :162-165 / :182-187 / :206-211.
//   MODE 0: write (first sweep) or accumulate this direction's flux divergence into tend[5]
//   MODE 1: last sweep: finish the tendency and write qout = a*qbase + b*roundtrip(qin) + c*tend
//   MODE 2: last sweep: finish the tendency and write it to qout (tendency-level parity entry point)
// ---------------------------------------------------------------------------------------------------------------
// Register budget of the state march (one-warp blocks: 65536 / (32 * regs) resident warps per SM). Measured on B200
// (profiles/r01g): ord 5 x/y march 168 (12 warps), z march 144 (14 warps; at 128 ptxas spills the loop control, and
// the reloads stall every R2UR of the uniform march position on the load scoreboard), ord 3 128, ord 7/9 255.
#ifndef AWFL_MAXNREG_Z
#define AWFL_MAXNREG_Z 144
#endif
#ifndef AWFL_MAXNREG_H
#define AWFL_MAXNREG_H 168
#endif
// The 2-D vertical march that carries the first tracer (TRQ) keeps the overflow-proof reconstruction behind a call inside
// its rolled loop: at 144 registers it spills 184 B and runs 9.2 ms on C2, at 168 none and 6.97 ms (r02G).
#ifndef AWFL_MAXNREG_ZQ
#define AWFL_MAXNREG_ZQ 168
#endif
#ifndef AWFL_MAXNREG_ZQ3
#define AWFL_MAXNREG_ZQ3 128
#endif
#define AWFL_STATE_BOUNDS __launch_bounds__(MT) __maxnreg__(ORD == 3 ? (DIR == 2 && TRQ ? AWFL_MAXNREG_ZQ3 : 128) : (ORD == 5 ? (DIR == 2 ? (TRQ ? AWFL_MAXNREG_ZQ : AWFL_MAXNREG_Z) : AWFL_MAXNREG_H) : 255))
template <int ORD, int DIR, int MODE, int VT, bool S2D, bool TRQ>
__global__ void AWFL_STATE_BOUNDS
k_march_state(const Geom g, const Phys ph, const SweepArgs a) {
  using F = RingF<S2D, TRQ>;
  constexpr int NVAR = F::NVAR, NVS = F::NVS;
  // 2-D: v is only needed where the stage is combined (its round-tripped value), not by a sweep that accumulates tendencies
  constexpr bool NEED_V = !(S2D && MODE == 0);
  constexpr int HS = (ORD + 1) / 2;
  constexpr double cs = 350.0;             // Dycore.h:335
  constexpr double rcs = 1.0 / 350.0;
  constexpr int NORMAL = DIR == 0 ? idU : (DIR == 1 ? idV : idW);
  constexpr int NV = (DIR != 2 && ORD <= 5) ? AWFL_NV_H : AWFL_NV;             // reconstructions per group
  // unrolling of the loop over variable groups. The vertical march stays rolled at ord >= 5 (unrolled, the copies share their
  // per-level table loads and spill 0.8 / 1.6 KB at ord 5 / 7); ord 3 has 17 table values per level and unrolls without
  // spills: z march 4.64 -> 4.14 ms on C4 (profiles/r02q_ord3_variants.txt)
  constexpr int V0_UNROLL = DIR == 2 ? (ORD == 3 ? 3 : AWFL_V0_UNROLL_Z) : AWFL_V0_UNROLL_H;
  extern __shared__ double ring_smem[];
  Item it;
  // one warp per block: the march position depends on blockIdx only, so the compiler keeps the loop control and the
  // per-level matrix addresses in uniform registers
  if (!decode_item<DIR>(g, a.sbits, a.labits, it, DIR == 2 ? a.seg_only : -1)) return;
  const int n = DIR == 0 ? g.nx : (DIR == 1 ? g.ny : g.nz);
  const long sd = stride_of(DIR, g);
  // cells of one field: fits 32 bits (checked by awfl_create), which lets a field offset k * nc be ONE 32 x 32 -> 64 bit
  // multiply-add instead of a 64-bit product
  const size_t nc = (unsigned)g.ncell;
  const double *__restrict__ q = a.qin;
  double *const rb = ring_smem + threadIdx.x;            // this thread's column: entry e at rb[e*MT]
  double *const fpv = rb + (RF_COUNT * ORD) * MT;        // previous interface: fluxes [0..4], right-edge values [5..8]
  double *const stg = fpv + (NS + 4) * MT;               // staged read-modify-write operands of the current group
  const long voff = a.vst > 1 ? it.e : 0;

  // Cells are fetched in marching order through a cursor that moves one cell per fetch -- across the periodic seam in the
  // horizontal marches, clamped at the walls in the vertical one (wall ghosts derive from the boundary cell) -- instead of
  // a 64-bit index product per cell and field.
  auto wrapped = [&](int c) -> int {
    if (DIR == 2) return c < 0 ? 0 : (c >= n ? n - 1 : c);
    return c < 0 ? c + n : (c >= n ? c - n : c);                      // periodic, Dycore.h:629-657
  };
  int fc = it.c0 - 1 - (HS - 1);                                      // next cell to fetch (unwrapped)
  const long seam_back = (long)(n - 1) * sd;
  const double *qp = q + it.base + (long)wrapped(fc) * sd;
  const long pdelta = a.pcache - q;                                   // the pressure of a cell sits at the same offset in its own field
  auto fetch_next = [&]() -> Raw {
    Raw x;
    x.r = __ldg(qp); x.ru = __ldg(qp + nc); x.rw = __ldg(qp + 3 * nc);
    x.rv = NEED_V ? __ldg(qp + 2 * nc) : 0.0;
    x.rt = __ldg(qp + 4 * nc);
    x.tq = TRQ ? __ldg(qp + 5 * nc) : 0.0;                              // the first tracer: field NS of the stage input
    // the pressure C0 (rho theta)^gamma of a cell is the same in every direction of a stage: the x march (always the
    // first sweep) computes it, the other marches read it back -- one x^gamma (~160 instructions) per cell and stage
    // instead of one per direction
    x.p = DIR == 0 ? 0.0 : __ldg(qp + pdelta);
    long step;
    if (DIR == 2) step = (fc >= 0 && fc < n - 1) ? sd : 0;            // clamped at the walls
    else step = wrapped(fc) == n - 1 ? -seam_back : sd;               // back across the seam
    qp += step;
    fc++;
    return x;
  };
  // conserved -> primitives + pressure (Dycore.h:310-321); wall ghosts per Dycore.h:662-710 with the intended
  // semantics of SURVEY.md F2 (ghost theta = theta of the boundary cell)
  auto put = [&](const Raw &x, int c, int so) {
    const double rinv = rcp_pos(x.r);
    double r = x.r, w = x.rw * rinv;
    double m = DIR == 0 ? x.ru : (DIR == 1 ? x.rv : x.rw);
    double p;
    if (DIR == 0) {
      p = ph.C0 * pow_pos(x.rt, ph.gamma);
      if (c >= 0 && c < n) a.pcache[it.base + (long)c * sd] = p;
    } else {
      p = x.p;
    }
    if (DIR == 2 && (c < 0 || c >= n)) {
      // hydrostatic constant-theta extrapolation: rho_g^(gamma-1) = rho0^(gamma-1) +- grav (gamma-1) dz0 dist / (gamma C0 theta0^gamma)
      // (Dycore.h:683-690, :697-704), with C0 theta0^gamma = p0 / rho0^gamma
      const int dist = c < 0 ? -c : c - n + 1;
      const double sgn = c < 0 ? 1.0 : -1.0;
      const double dz0 = __ldg(a.dz + (long)(c < 0 ? 0 : n - 1) * g.nens + it.e);
      const double A = pow_pos(x.r, ph.gamma);
      const double coef = ph.grav * (ph.gamma - 1.0) * dz0 * A / (ph.gamma * p);
      const double X = fma(sgn * coef, (double)dist, A * rinv);
      const double rg = pow_pos(X, 1.0 / (ph.gamma - 1.0));
      p = (p / A) * rg * X;            // C0 (rho_g theta0)^gamma, Dycore.h:691-694
      r = rg; w = 0.0; m = 0.0;
    }
    rb[(F::M * ORD) * MT + so] = m;
    rb[(F::P * ORD) * MT + so] = p;
    rb[(F::U * ORD) * MT + so] = x.ru * rinv;
    if (NEED_V) rb[(F::V * ORD) * MT + so] = x.rv * rinv;
    rb[(F::W * ORD) * MT + so] = w;
    rb[(F::T * ORD) * MT + so] = x.rt * rinv;
    // mixing ratio, Dycore.h:320; a wall ghost repeats the boundary cell's (zero gradient, Dycore.h:674-677: rinv is still that cell's)
    if (TRQ) rb[(F::Q * ORD) * MT + so] = x.tq * rinv;
    rb[(F::R * ORD) * MT + so] = r;
  };

  int head = 0;                       // ring slot of the oldest of the ORD most recent cells
  const double rdd_h = DIR == 0 ? 1.0 / g.dx : 1.0 / g.dy;
  double mL = 0.0, pL = 0.0, mRs = 0.0, mLns = 0.0;
  const int jfirst = it.c0 - 1 - (HS - 1), jlast = it.c1 + HS - 1;
  Raw cur = fetch_next();

#pragma unroll 1
  for (int j = jfirst; j <= jlast; j++) {
    Raw nxt = cur;
    if (j < jlast) nxt = fetch_next();
    const int c = j - (HS - 1);        // cell whose window c-HS+1..c+HS-1 will be complete after this put
    const bool upd_on = c > it.c0;             // cell c-1 has both interfaces
    const int cc = c - 1;
    const long idx = it.base + (long)cc * sd;
    const int k = DIR == 2 ? cc : it.la;
    double rdd = rdd_h, gv = 0.0;
    if (upd_on) {                      // per-level scalars of the update: issued before the conversion below hides them
      if (DIR == 2) rdd = __ldg(a.rdz + (long)k * g.nens + it.e);
      if (MODE != 0) gv = __ldg(a.grav_var + (long)k * g.nens + it.e);
    }
    put(cur, j, head * MT);            // newest cell replaces the oldest
    head = head + 1 == ORD ? 0 : head + 1;
    cur = nxt;
    if (c < it.c0 - 1) continue;
    const typename MatSel<ORD, VT>::type mat = MatSel<ORD, VT>::get(a.vtab, a.vst, voff, c + 1);
    const bool flux_on = c >= it.c0;           // interface c belongs to this segment
    const bool wall = (DIR == 2) && (c == 0 || c == n);   // Dycore.h:477,482,496
    const int so = ring_slot<ORD>(head, HS - 2);   // ring slot of cell c-1
    const double rho = (upd_on && MODE != 0) ? rb[(F::R * ORD) * MT + so] : 0.0;
    double rn = 0.0, pp = 0.0;
    bool up = false;
    // Read-modify-write operands of cell c-1 (partial tendency, step-start state) arrive by asynchronous copies. Where the
    // loop over the variable groups is unrolled (horizontal marches) ALL of them are started here, one commit group per
    // variable group, and land while the groups are reconstructed; the rolled vertical march keeps one group ahead (all
    // started here it spills 0.8 KB; started right before the group's own reconstruction the copies were late: 4 % of the
    // z march's stall samples, 7.31 -> 7.12 ms)
    constexpr bool PRE_ALL = V0_UNROLL > 1;
    const bool ld_t = upd_on && !a.first, ld_q = upd_on && MODE == 1 && a.a != 0.0;
    // Cursors over the five state variables l = 0..4 of cell c-1, advanced once per variable. (Addressed as base + l * stride
    // with the run-time l of the rolled vertical march, every access cost 8-9 integer instructions -- the 64-bit l * ncell
    // and the shared-memory slots: a quarter of the z march's non-FP64 instructions, profiles/r02d_state_z_line_mix.txt.)
    const double *pf_t = a.tend + idx, *pf_q = a.qbase + idx;   // operands still to be staged: partial tendency, step-start state
    double *pf_s = stg;                                        // ... and where they are staged: (tend, base) pairs
    // 2-D: the step-start v of cell c-1 for its stage combination (its field is skipped by the cursors); loaded here, used last
    double qb_v = 0.0;
    if (S2D && ld_q) qb_v = a.qbase[idV * nc + idx];           // plain load: qout may alias qbase, read before the write below
    // the reconstructed variables go NV at a time; where NV does not divide their number (2-D) a shorter last group follows
    constexpr int VFULL = (NVAR / NV) * NV, NTAIL = NVAR - VFULL;
    // state field behind reconstructed variable v >= 1: rho, u, [v,] w, theta -- in 2-D the field cursors step over v
    auto start_copies = [&](int g0) {
#pragma unroll
      for (int i = 0; i < NV; i++) {
        const int v = g0 + i;
        if (v > 0 && v < NVS) {
          if (ld_t) cp_async8(pf_s, pf_t);
          if (ld_q) cp_async8(pf_s + MT, pf_q);   // qout may alias qbase: read before the write below
          pf_s += 2 * MT;
          if (S2D && v == 2) { pf_t += 2 * nc; pf_q += 2 * nc; } else { pf_t += nc; pf_q += nc; }
        }
      }
    };
    if (PRE_ALL && (ld_t || ld_q)) {
#pragma unroll
      for (int g0 = 0; g0 < NVAR; g0 += NV) { start_copies(g0); cp_async_commit(); }
    }
    if (!PRE_ALL && (ld_t || ld_q)) { start_copies(0); cp_async_commit(); }
    double *po = (MODE == 0 ? a.tend : a.qout) + idx;           // where the state field of the current variable goes (cell c-1)
    const double *sg = stg;                                     // its staged operands
    double *fl = fpv;                                           // its flux through the previous interface
    double *fe = fpv + (NS - 1) * MT;                           // its right-edge value in the previous cell (advected variables)
    const double *rl = rb + ((F::U - 1) * ORD) * MT + so;       // its primitive value of cell c-1 in the ring (advected variables)
    // one reconstructed variable: v = 0 rho*u_n, 1 p (the acoustic pair closes interface c), then the advected u, [v,] w, theta
    auto handle = [&](int v, double E, double O) {
      if (v == 0) { mRs = E - O; mLns = E + O; return; }
      if (TRQ && v == NVS) {
        // the first tracer: flux = mass flux x upwind edge value of the mixing ratio, zero through the walls
        // (Dycore.h:381-385 / :513-517, :477, :496); no operands, no update here -- the FCT kernels finish the tracers
        double *const feq = fpv + (NS + 3) * MT;               // the right-edge slot v leaves free
        double Fq = rn * (up ? feq[0] : E - O);
        feq[0] = E + O;
        if (wall) Fq = 0.0;
        if (flux_on && (c < it.c1 || (DIR == 2 && c == n))) a.tflux[it.base + (long)c * sd] = Fq;
        return;
      }
      // this variable's cursors; the shared ones move on to the next variable
      double *const po_l = po, *const fl_l = fl, *const fe_l = fe;
      const double *const sg_l = sg, *const rl_l = rl;
      po += (S2D && v == 2) ? 2 * nc : nc;
      sg += 2 * MT; fl += MT; fe += MT; rl += ORD * MT;
      int l;
      double Fx;
      if (v == 1) {
        // interface c: characteristic upwinding, Dycore.h:361-365
        const double pR = E - O;
        double mR = mRs;
        if (wall) { mL = 0.0; mR = 0.0; }
        const double w1 = 0.5 * (pR - cs * mR);
        const double w2 = 0.5 * (pL + cs * mL);
        pp = w1 + w2;
        rn = wall ? 0.0 : (w2 - w1) * rcs;
        up = rn > 0.0;               // Dycore.h:368: the upwind cell is c-1 when the mass flux is positive
        mL = mLns; pL = E + O;
        // interface mass flux for the tracer march (x,y: n entries, seam stored at 0; z: n+1 entries)
        if (flux_on && (c < it.c1 || (DIR == 2 && c == n))) a.mflux[it.base + (long)c * sd] = rn;
        l = idR; Fx = rn;
      } else {
        l = (S2D && v >= 3) ? v : v - 1;                        // idU..idT
        Fx = rn * (up ? fe_l[0] : E - O);
        fe_l[0] = E + O;
        if (l == NORMAL) Fx += pp;
      }
      if (!flux_on) return;
      if (upd_on) {
        const double div = (Fx - fl_l[0]) * rdd;
        double T = (ld_t ? sg_l[0] : 0.0) - div;
        if (MODE == 0) po_l[0] = T;
        else {
          if (l == idW) T = fma(-gv, rho, T);                 // Dycore.h:560
          if (!S2D && l == idV && g.sim2d) T = 0.0;           // Dycore.h:565 (a 2-D run on the general kernel)
          if (MODE == 2) po_l[0] = T;
          else {
            // round-tripped stage input (x/rho)*rho, Dycore.h:316-319 + :527-530
            const double rt = l == idR ? rho : rl_l[0] * rho;
            const double o = fma(a.a, ld_q ? sg_l[MT] : 0.0, a.b * rt);
            po_l[0] = fma(a.c, T, o);
          }
        }
      }
      fl_l[0] = Fx;
    };
    // reconstructed quantities in ring order, NV at a time: their reconstructions together (shared table loads), then each
    // variable's fluxes / update
#pragma unroll(V0_UNROLL)
    for (int v0 = 0; v0 < VFULL; v0 += NV) {
      if (!PRE_ALL && (ld_t || ld_q)) { if (v0 + NV < NVAR) start_copies(v0 + NV); cp_async_commit(); }   // one group ahead
      // vertical marches on a shared grid: inlined (uniform per-level table loads are not hoisted out of the cell loop, and
      // without the call the ~100 argument / return moves per cell go away: z march 7.80 -> 7.33 ms). Horizontal marches
      // and member-dependent grids keep the real call (see weno_ring; inlined they spill: measured 7.11 vs 6.78 ms in x,
      // 0.5-2 KB of spills with the strided tables)
      EOn<NV> eo;
      if constexpr (VT == 1 || VT == 3) eo = weno_ring_inl<ORD, NV>(rb + (v0 * ORD) * MT, head, mat);
      else eo = weno_ring<ORD, NV>(rb + (v0 * ORD) * MT, head, mat);
      if (ld_t || ld_q) {
        if (PRE_ALL) cp_async_wait_left((NVAR - 1 - v0) / NV);   // the later groups may still be in flight
        else cp_async_wait_left(1);                              // all but the group just started
      }
      if (TRQ && v0 + NV == NVAR) {
        // tracers carry arbitrary units (number concentrations ...): a magnitude that overflows the fast weights gives NaN
        // or inf and goes through the overflow-proof ones, as in the tracer march
        if (!(fabs(eo.E[NV - 1]) <= 1.7976931348623157e308)) {
          const EOn<1> sf = weno_ring_safe<ORD>(rb + (F::Q * ORD) * MT, head, mat);
          eo.E[NV - 1] = sf.E[0]; eo.O[NV - 1] = sf.O[0];
        }
      }
#pragma unroll
      for (int i = 0; i < NV; i++) handle(v0 + i, eo.E[i], eo.O[i]);
    }
    if constexpr (NTAIL > 0) {
      // 2-D: the shorter last group, always through the real call (a second inlined reconstruction next to the loop's spills
      // 0.8-4.7 KB in the vertical marches)
      if (!PRE_ALL && (ld_t || ld_q)) cp_async_commit();
      const EOn<NTAIL> eo = weno_ring<ORD, NTAIL>(rb + (VFULL * ORD) * MT, head, mat);
      if (ld_t || ld_q) cp_async_wait_left(0);
#pragma unroll
      for (int i = 0; i < NTAIL; i++) handle(VFULL + i, eo.E[i], eo.O[i]);
    }
    if (S2D && upd_on && MODE != 0) {
      // v of cell c-1 in a 2-D run: tendency 0 (Dycore.h:565), i.e. the stage combination of its round-tripped value
      double *const pv = a.qout + idV * nc + idx;
      if (MODE == 2) pv[0] = 0.0;
      else pv[0] = fma(a.a, qb_v, a.b * (rb[(F::V * ORD) * MT + so] * rho));
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Tracer march: the tracer part of Dycore.h:381-385 / :434-438 / :513-517. flux = massflux * upwind WENO edge value of
// (tracer/rho), cell-centred like the state march. Work item = (tracer, line, segment, ensemble tile).
// ---------------------------------------------------------------------------------------------------------------
// PAIR = false: a warp marches two adjacent lines of ONE tracer; PAIR = true: one line of TWO tracers (tr, tr+1), which
// share the line's density, its reciprocal and the mass flux -- loaded and computed once per cell instead of per tracer.
template <int ORD, int DIR, int VT, bool PAIR>
__global__ void __launch_bounds__(MT, MarchCfg<ORD>::MINB_TR)
k_march_tracer(const Geom g, const TracerSweepArgs a) {
  constexpr int HS = (ORD + 1) / 2;
  constexpr int NL = 2;                                       // reconstructions marched together by one warp
  extern __shared__ double ring_smem[];
  Item it;                                                    // one warp per block (see k_march_state)
  // the two lines of a warp are adjacent levels k, k+1 for the horizontal marches and adjacent columns i, i+1 for the
  // vertical march (an odd count repeats the last line: the same values are written twice)
  if (!decode_item<DIR, (PAIR || DIR == 2) ? 1 : NL, (!PAIR && DIR == 2) ? NL : 1>(g, a.sbits, a.labits, it, DIR == 2 ? a.seg_only : -1)) return;
  const int tr = a.tr0 + (PAIR ? 2 * it.tr : it.tr);
  if ((PAIR ? 2 * it.tr + 1 : it.tr) >= a.ntr || tr + (PAIR ? 1 : 0) >= g.nt) return;   // pair launches carry an even count
  const int n = DIR == 0 ? g.nx : (DIR == 1 ? g.ny : g.nz);
  const long sd = stride_of(DIR, g);
  const size_t nc = (unsigned)g.ncell;                        // < 2^32, see k_march_state
  const double *__restrict__ rho = a.qin;
  const size_t fstride = DIR == 2 ? nc + nc / g.nz : nc;      // one tracer's interface fluxes of this direction
  const double *__restrict__ tq = a.qin + (size_t)(NS + tr) * nc;   // first (or only) tracer of the warp
  double *__restrict__ fl = a.tflux + (size_t)tr * fstride;
  // offsets of reconstruction v: the line (base) and, in a pair, the tracer (tb into the tracer fields, fb into the fluxes)
  long base[NL], tb[NL], fb[NL];
  base[0] = it.base;
  if (PAIR) base[1] = it.base;
  else if (DIR == 2) base[1] = it.lb + 1 < g.nx ? it.base + g.nens : it.base;
  else base[1] = it.la + 1 < g.nz ? it.base + (long)g.ny * g.nx * g.nens : it.base;
#pragma unroll
  for (int v = 0; v < NL; v++) {
    tb[v] = PAIR ? base[v] + (long)(v * nc) : base[v];
    fb[v] = PAIR ? base[v] + (long)(v * fstride) : base[v];
  }
  double *const rb = ring_smem + threadIdx.x;                 // ring of line v at rb[(v*ORD + slot)*MT]
  double *const stg = rb + (NL * ORD) * MT;                   // staged next cell (tracer, rho) and next mass flux, per line
  const long voff = a.vst > 1 ? it.e : 0;
  auto cell_off = [&](int c) -> long {
    int cw;
    if (DIR == 2) cw = c < 0 ? 0 : (c >= n ? n - 1 : c);       // wall: zero-gradient mixing ratio, Dycore.h:674-677
    else cw = c < 0 ? c + n : (c >= n ? c - n : c);            // periodic, Dycore.h:629-657
    return (long)cw * sd;
  };
  int head = 0;
  {
    const int cs0 = it.c0 - 1 - (HS - 1);
#pragma unroll
    for (int s = 0; s < ORD; s++) {
      const long o = cell_off(cs0 + s);
      if (PAIR) {
        const double rinv = rcp_pos(__ldg(rho + base[0] + o));   // Dycore.h:320
#pragma unroll
        for (int v = 0; v < NL; v++) rb[(v * ORD + s) * MT] = __ldg(tq + tb[v] + o) * rinv;
      } else {
#pragma unroll
        for (int v = 0; v < NL; v++) rb[(v * ORD + s) * MT] = __ldg(tq + tb[v] + o) * rcp_pos(__ldg(rho + base[v] + o));   // Dycore.h:320
      }
    }
  }
  // staging of the next cell: PAIR: [tracer 0, tracer 1, rho, mass flux]; else per line [tracer, rho, mass flux]
  double Rp[NL] = {0.0, 0.0}, rn[NL] = {0.0, 0.0};
#pragma unroll 1
  for (int c = it.c0 - 1; c < it.c1; c++) {
    const bool more = c + 1 < it.c1;
    if (more) {                          // asynchronous: these stay in flight across the call (see cp_async8)
      const long o = cell_off(c + HS), om = (long)(c + 1) * sd;
      if (PAIR) {
        cp_async8(stg, tq + tb[0] + o);
        cp_async8(stg + MT, tq + tb[1] + o);
        cp_async8(stg + 2 * MT, rho + base[0] + o);
        cp_async8(stg + 3 * MT, a.mflux + base[0] + om);
      } else {
#pragma unroll
        for (int v = 0; v < NL; v++) {
          cp_async8(stg + (3 * v) * MT, tq + tb[v] + o);
          cp_async8(stg + (3 * v + 1) * MT, rho + base[v] + o);
          cp_async8(stg + (3 * v + 2) * MT, a.mflux + base[v] + om);
        }
      }
    }
    const typename MatSel<ORD, VT>::type mat = MatSel<ORD, VT>::get(a.vtab, a.vst, voff, c + 1);
    EOn<NL> eo;
    if constexpr (VT == 0) eo = weno_ring<ORD, NL>(rb, head, mat); else eo = weno_ring_inl<ORD, NL>(rb, head, mat);
    // tracers carry arbitrary units (number concentrations ...): a magnitude that overflows the fast weights gives NaN or
    // inf -- one test for both reconstructions (a sum that overflows by itself only sends finite values through the slow
    // path too)
    if (!(fabs(eo.E[0]) + fabs(eo.E[1]) <= 1.7976931348623157e308)) {
#pragma unroll
      for (int v = 0; v < NL; v++) {
        const EOn<1> sf = weno_ring_safe<ORD>(rb + (v * ORD) * MT, head, mat);
        eo.E[v] = sf.E[0]; eo.O[v] = sf.O[0];
      }
    }
#pragma unroll
    for (int v = 0; v < NL; v++) {
      const double E = eo.E[v], O = eo.O[v];
      if (c >= it.c0) {
        double flux = rn[v] * (rn[v] > 0.0 ? Rp[v] : E - O);
        if (DIR == 2 && c == 0) flux = 0.0;                     // bottom wall: rw = 0 (Dycore.h:477,482)
        fl[fb[v] + (long)c * sd] = flux;
      }
      Rp[v] = E + O;
    }
    if (more) {
      cp_async_wait();
      if (PAIR) {
        const double rinv = rcp_pos(stg[2 * MT]);                 // Dycore.h:320
        rb[head * MT] = stg[0] * rinv;
        rb[(ORD + head) * MT] = stg[MT] * rinv;
        rn[0] = rn[1] = stg[3 * MT];                              // mass flux of interface c+1
      } else {
#pragma unroll
        for (int v = 0; v < NL; v++) {
          rb[(v * ORD + head) * MT] = stg[(3 * v) * MT] * rcp_pos(stg[(3 * v + 1) * MT]);   // Dycore.h:320
          rn[v] = stg[(3 * v + 2) * MT];                            // mass flux of interface c+1
        }
      }
      head = head + 1 == ORD ? 0 : head + 1;
    }
  }
  if (DIR == 2 && it.c1 == n) {                                 // top wall: rw = 0 (Dycore.h:496)
#pragma unroll
    for (int v = 0; v < NL; v++) fl[fb[v] + (long)n * sd] = 0.0;
  }
}

}  // namespace awfl
